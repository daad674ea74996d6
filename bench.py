#!/usr/bin/env python
"""bench.py — Offline utterances/s of the int8 RNN-T hot path on N B200s (+ encoder LSTM-step us).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--scaling weak|strong]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json config 3): one STEP = one Offline query of synthetic utterances (dev-clean-like length
distribution, 240-dim stacked filterbank features, random-init weights with synthetic calibration ranges), sorted by
length descending like PytorchSUT.issue_queries; the sorted utterances are dealt round-robin to the ranks and every rank
cuts its share into ragged batches of <= --batch (2048 on B200; the reference's CPU run uses 256).  Weak scaling
(default): 2513 utterances per GPU per step.  Strong scaling (--scaling strong): ONE query of 24576 samples
(configs/mlperf.conf:63) for the whole job.  The only collective is the single all_gather of (id, length, tokens) that
ends a query.

  value  utterances/s with the query's features already in HBM (device-timed with CUDA events, max over ranks); no
         per-kernel instrumentation runs inside this timed region
  e2e    the same metric through PytorchSUT + the LoadGen shim with HOST (pinned) features: host->device copies and
         the device->host token read are inside the timed region
  roofline      the persistent recurrent LSTM kernel (the dominant kernel), bracketed by CUDA events in a separate,
                untimed pass over the same step; bytes = TRUE (frame, utterance) pairs x the per-unit figure
  lstm_step_us_c2 / encoder_microbench   BASELINE config 2: encoder only, 256 x 500 frames
  parity        transcripts of the cpu_baseline's utterances decoded by the GPU path in the same batches, compared with
                the CPU oracle's (identical utterances, token agreement, WER, top-2 margin attribution)
  cpu_baseline  the CPU oracle (a restatement of the reference's AMX path) on the host cores, bounded sample
`--impl reference` prints the reference arm: the same CPU path timed on the host cores on a bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np
import torch

UTTS_PER_GPU = 2513       # configs/mlperf.conf:13 performance_sample_count of the reference
BATCH = 2048              # SUT batch size on B200 (the reference CPU run uses 256, run.sh:71): more 128-row tiles per CTA in the recurrent kernel; <= 2240 decode lanes per launch, multiple of 128 for ragged tile skipping
QUERY_TOKEN_CAP = 256     # token columns of the transcript gather: the longest utterance has 250 stacked frames; a longer row is clipped in the gather (never observed: ~0.27 tokens per stacked frame)
DECODE_MARGIN_TOL = 0.3   # one bf16 ulp of the prediction state moves a logit of the synthetic joint by up to 8 * 8 * 2^-8 = 0.25 (tests/test_gpu_pipeline.py)
METRIC = "offline_utterances_per_second"
UNIT = "utterances/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--utts", type=int, default=UTTS_PER_GPU, help="utterances per GPU and step")
    ap.add_argument("--batch", type=int, default=BATCH)
    ap.add_argument("--streams", type=int, default=2, help="compute streams the batches of a step alternate between (PytorchSUT: RNNT_SUT_COMPUTE_STREAMS)")
    ap.add_argument("--cpu-seconds", type=float, default=20.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-microbench", action="store_true", help="skip config 2 (encoder only, 256 x 500 frames)")
    ap.add_argument("--no-parity", action="store_true", help="skip the post-run transcript check against the CPU oracle")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--query", type=int, default=24576, help="samples of the one query of --scaling strong (configs/mlperf.conf:63)")
    return ap.parse_args()


def dist_env():
    return int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))


class ClockSampler:
    """SM clock and throttle reasons DURING the timed region (B200_PROFILING.md recipe).  NVML in a thread of this
    process, one sample every 10 ms (a `nvidia-smi -lms` child takes longer to start on an 8-GPU box than five steps take);
    the nvidia-smi child is the fallback when the NVML binding is missing."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")
    NVML_REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"),
                    (0x4, "sw_power_cap"))

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.proc = None
        self.path = None
        self.thread = None
        self.stop_flag = threading.Event()
        self.sm, self.reason_bits, self.smax = [], 0, None
        self.nvml = self.handle = None
        try:                                    # handle looked up before the timed region starts
            import pynvml
            pynvml.nvmlInit()
            try:
                uuid = str(torch.cuda.get_device_properties(gpu_index).uuid)
                uuid = uuid if uuid.startswith("GPU-") else "GPU-" + uuid
                try:
                    self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid)
                except TypeError:
                    self.handle = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode())
            except Exception:
                self.handle = pynvml.nvmlDeviceGetHandleByIndex(gpu_index)
            self.smax = float(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.nvml = pynvml
        except Exception:
            self.nvml = self.handle = None

    def _sample_nvml(self):
        n = self.nvml
        reasons = getattr(n, "nvmlDeviceGetCurrentClocksEventReasons", None) or n.nvmlDeviceGetCurrentClocksThrottleReasons
        while True:
            try:
                self.sm.append(float(n.nvmlDeviceGetClockInfo(self.handle, n.NVML_CLOCK_SM)))
                self.reason_bits |= int(reasons(self.handle))
            except Exception:
                pass
            if self.stop_flag.wait(0.010):
                return

    def start(self):
        if self.nvml is not None:
            self.thread = threading.Thread(target=self._sample_nvml, daemon=True)
            self.thread.start()
            return
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu_index)], stdout=open(self.path, "w"),
                                         stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.thread is not None:
            self.stop_flag.set()
            self.thread.join(timeout=5)
            if self.sm:
                out.update(sm_mhz=float(np.median(self.sm)), sm_max_mhz=self.smax, samples=len(self.sm), source="nvml",
                           reasons=sorted(name for bit, name in self.NVML_REASONS if self.reason_bits & bit))
            return out
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, reasons, smax = [], set(), None
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 9:
                    continue
                try:
                    sm.append(float(parts[1]))
                    smax = float(parts[2])
                except ValueError:
                    continue
                for name, val in zip(names, parts[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=smax, reasons=sorted(reasons), samples=len(sm),
                       source="nvidia-smi")
        return out


# ------------------------------------------------------------------------------------------------ workload
def build_query(world, rank, utts_per_gpu, batch, total=None):
    """Global query = `total` utterances (default world * utts_per_gpu; lengths: fixed-seed gamma fit of dev-clean, the
    2513 distinct utterances replicated beyond like MLPerf does), sorted by length descending
    (PytorchSUT.issue_queries) and dealt round-robin to the ranks; each rank cuts its share into batches of `batch`
    (SURVEY.md 8(e)).  Returns this rank's list of batches, each a list of (utt_id, length), and the query's frames."""
    from rnnt_inference_b200 import sharding, synthetic as S
    total = world * utts_per_gpu if total is None else total
    lens = S.utterance_lengths(UTTS_PER_GPU)
    ulen = lens[np.arange(total) % UTTS_PER_GPU]
    return sharding.shard_batches(ulen, batch, world, rank), int(ulen.sum())


def make_model(device, batch=BATCH):
    from rnnt_inference_b200 import synthetic as S
    from rnnt_inference_b200.decoder import GreedyDecoder
    from rnnt_inference_b200.modeling_rnnt import RNNT
    sd = S.make_checkpoint()
    rnnt = RNNT(sd, "quant", True, device=device).eval()
    return sd, GreedyDecoder(rnnt, "quant", True, -1, batch)


def features_for(batches):
    """Host features of every distinct utterance this rank touches (replicas share one tensor)."""
    from rnnt_inference_b200 import synthetic as S
    feats = {}
    for b in batches:
        for u, n in b:
            key = u % UTTS_PER_GPU
            if key not in feats:
                feats[key] = S.make_features(n, key)
    return feats


def stage_on_device(batches, feats, device):
    """Device-resident form of each batch: packed ragged fp32 [sum T_i, 240], offsets, lens (N padded to 32)."""
    staged = []
    for b in batches:
        n_pad = (len(b) + 31) // 32 * 32
        lens = [n for _, n in b]
        packed = torch.cat([feats[u % UTTS_PER_GPU] for u, _ in b], 0).to(device)
        offsets = torch.zeros(n_pad, dtype=torch.int64)
        offsets[:len(b)] = torch.tensor(np.concatenate([[0], np.cumsum(lens[:-1])]))
        lens_t = torch.zeros(n_pad, dtype=torch.int32)
        lens_t[:len(b)] = torch.tensor(lens, dtype=torch.int32)
        ids = torch.tensor([u for u, _ in b], dtype=torch.int64, device=device)
        staged.append((packed, offsets.to(device), lens_t.to(device), max(lens), len(b), lens, ids))
    return staged


_batches_run = [0]


def run_step_device(decoder, staged, world=1, n_max=None, streams=None):
    """One step = one pass over this rank's query, inputs already in HBM; with N > 1 the step ends with the ONE
    all_gather of (sample id, length, tokens) from all ranks (the path's only collective; n_max = rows per rank and the
    token width are known from the query, so no shape exchange).  With `streams`, consecutive batches rotate over them like
    the chunks of PytorchSUT.issue_queries (the caller forks / joins them around the timed region; with N > 1 they are
    joined before the gather).  Returns the emitted-token count as a list of device scalars, one per batch."""
    from rnnt_inference_b200 import sharding
    totals, outs = [], []
    cap = max((m + 1) // 2 for _p, _o, _l, m, _n, _h, _i in staged) if staged else 1
    main = torch.cuda.current_stream()
    for packed, offsets, lens, max_len, n_actual, lens_host, ids in staged:
        _batches_run[0] += 1
        with torch.cuda.stream(streams[_batches_run[0] % len(streams)] if streams else main):
            res, res_len = decoder.forward_ragged(packed, offsets, lens, max_len, lens_host)
            totals.append(res_len[:n_actual].sum())
            if world > 1:
                tok = torch.full((n_actual, cap), -1, dtype=torch.int32, device=res.device)
                w = min(cap, res.shape[1])
                tok[:, :w] = res[:n_actual, :w]
                outs.append((ids, res_len[:n_actual].to(torch.int32), tok))
    if world > 1 and streams:
        for st in streams:
            main.wait_stream(st)
    if world > 1:
        dev = staged[0][0].device if staged else torch.device("cuda", torch.cuda.current_device())
        cat = lambda k, dt, shape: torch.cat([o[k] for o in outs], 0) if outs else torch.empty(shape, dtype=dt, device=dev)
        sharding.gather_transcripts(cat(0, torch.int64, (0,)), cat(1, torch.int32, (0,)), cat(2, torch.int32, (0, cap)), world,
                                    n_max=n_max, l_max=QUERY_TOKEN_CAP)
        if streams:                      # the next step's batches reuse the gathered buffers' streams
            for st in streams:
                st.wait_stream(main)
    return totals


# ------------------------------------------------------------------------------------------------ roofline
NCU_TRAFFIC_RATIO = 9.626 / 12.362


def recurrent_roofline(events, peaks):
    """Algorithmic HBM bytes of the recurrent kernel per launch (DESIGN.md / SURVEY.md 8(d)): per (timestep, utterance)
    G int32 16384 + h_in 1024 + c read+write 4096 + h_out 1024 + y 1024 (int8) or 2048 (bf16) = 23552 / 24576 bytes."""
    recs = events.get("lstm_recurrent", [])
    if not recs:
        return None, None
    tot_ms, tot_bytes, tot_ops, items = 0.0, 0.0, 0.0, 0
    for start, end, (rows, tile_items, last) in recs:   # rows = TRUE (frame, utterance) pairs of the launch (ops.lstm_layer_i8)
        tot_ms += start.elapsed_time(end)
        tot_bytes += float(rows) * (24576 if last else 23552)
        tot_ops += float(rows) * 2 * 1024 * 4096
        items += tile_items
    launches = len(recs)
    achieved = tot_bytes / (tot_ms * 1e-3) / 1e9
    peak = peaks.get("hbm_gbs", 6650.0)
    # traffic: dram__bytes_read.sum + dram__bytes_write.sum of ONE `ncu --set full` launch (first pre-RNN layer of the bench's
    # 2048-utterance batch, profiles/r02/ncu_raw_summary.md): 9.031 + 0.596 GB against 12.362 GB algorithmic (524 866 true
    # frames x 23 552 B; part of G is still in L2 from the GEMM that wrote it).  The measured ratio is applied to this
    # run's mean launch so the figure is per launch like `achieved`.
    traffic = round(NCU_TRAFFIC_RATIO * tot_bytes / launches)
    roof = {"kernel": "lstm_recurrent_kernel", "bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
            "frac": round(achieved / peak, 4), "traffic": traffic,
            "bytes_basis": "true (frame, utterance) pairs x 23552 B (24576 B for the bf16 last layer); tile padding and rows past an utterance's end are not counted",
            "traffic_source": "ncu --set full (profiles/r02/ncu_raw_summary.md), pre-RNN layer 0 of the 2048-utterance batch: 9.626 GB dram r+w / 12.362 GB algorithmic = 0.779, scaled to this run's mean launch",
            "peak_source": peaks.get("_source", "fallback"),
            "launches": launches, "avg_launch_ms": round(tot_ms / launches, 4),
            "algorithmic_bytes_per_launch": round(tot_bytes / launches), "tensor_top_s": round(tot_ops / (tot_ms * 1e-3) / 1e12, 1)}
    return roof, tot_ms * 1e3 / items     # second value: mean us per (layer, timestep, 128-row tile) item


def kernel_shares(events):
    out = {}
    for name, recs in events.items():
        out[name] = round(sum(s.elapsed_time(e) for s, e, _ in recs), 3)
    if "greedy_decode" in events:       # one launch per batch: (lanes, stacked frames of the longest, ms)
        out["greedy_decode_per_batch"] = [(u[1], u[0], round(s.elapsed_time(e), 3)) for s, e, u in events["greedy_decode"]]
    return out


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            p = json.load(open(path))
            p["_source"] = "MEASURED_PEAKS.json (burst copy bandwidth)"
            return p
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "_source": "fallback (B200_PROFILING.md)"}


# ------------------------------------------------------------------------------------------------ CPU baseline
def host_threads():
    """Threads the CPU arm may use: every core this process is allowed on (torch.distributed.run exports
    OMP_NUM_THREADS=1, which must not turn the baseline into a single-thread run)."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_baseline(sd, world_batches, mean_frames, budget_s, keep_chunks=False):
    """The CPU oracle (restatement of the reference's int8/bf16 path, OpenMP over the batch) on the host cores, on
    contiguous chunks of the length-sorted query taken at spread quantiles.  Throughput is frames-weighted:
    (frames processed / seconds) / mean frames per utterance of the query, so a short sample is not biased by the
    lengths it happened to contain.  keep_chunks: also return [(utterances, tokens, decision trace)] per chunk for the
    transcript check of the GPU path."""
    from oracle import oracle as O
    from oracle import pipeline as OP
    from rnnt_inference_b200 import synthetic as S
    cores = O.set_num_threads(host_threads())
    chunk = int(min(256, max(32, cores)))
    flat = [u for b in world_batches for u in b]          # (utt_id, length), sorted by length descending
    params = OP.prepare_params(sd)
    frames, utts, elapsed, used, kept = 0, 0, 0.0, [], []
    for q in (0.5, 0.25, 0.75, 0.125, 0.625, 0.375, 0.875):
        start = int(q * max(0, len(flat) - chunk))
        part = flat[start:start + chunk]
        feats = [S.make_features(n, u % UTTS_PER_GPU) for u, n in part]
        x, xl = OP.pad_batch(feats, pad_to=32)
        trace = O.new_decision_trace(x.shape[1], int(xl.max()) + 96) if keep_chunks else None
        t0 = time.perf_counter()
        res, res_len, _ = OP.greedy_decode_batch(params, x, xl, trace=trace)
        elapsed += time.perf_counter() - t0
        frames += int(sum(n for _, n in part))
        utts += len(part)
        used.append(q)
        if keep_chunks:
            kept.append((part, [list(res[i, :res_len[i]]) for i in range(len(part))], trace))
        if elapsed >= budget_s:
            break
    value = (frames / elapsed) / mean_frames
    out = {"value": round(value, 3), "unit": UNIT, "cores": cores, "kind": "port",
           "sample": f"{utts} utterances ({frames} frames) in {len(used)} length-sorted chunks of {chunk} at quantiles {used}, "
                     f"{elapsed:.1f} s; frames-weighted to the query's mean of {mean_frames:.1f} frames/utterance"}
    return (out, kept) if keep_chunks else out


def parity_check(decoder, chunks, device):
    """Decode the cpu_baseline's chunks on the GPU in the same batches (same utterances, same order, N padded to 32) and
    compare the transcripts with the oracle's: identical utterances, token agreement, WER against the oracle, and the
    top-2 logit margin at the decision every difference starts at (the hoisted f.W_trans runs on the tensor cores, so a
    near-tie may flip — SURVEY 7.2)."""
    from rnnt_inference_b200 import eval_accuracy as E
    from rnnt_inference_b200 import synthetic as S
    test, truth, traces = [], [], []
    for part, toks, trace in chunks:
        feats = {u % UTTS_PER_GPU: S.make_features(n, u % UTTS_PER_GPU) for u, n in part}
        (packed, offsets, lens, max_len, n_actual, lens_host, _ids), = stage_on_device([part], feats, device)
        with torch.no_grad():
            res, res_len = decoder.forward_ragged(packed, offsets, lens, max_len, lens_host)
        res, res_len = res.cpu().numpy(), res_len.cpu().numpy()
        test += [list(res[i, :res_len[i]]) for i in range(n_actual)]
        truth += toks
        traces.append({k: v[:n_actual] for k, v in trace.items()})
    m = max(t["sym"].shape[1] for t in traces)
    pad = lambda a, fill: np.pad(a, ((0, 0), (0, m - a.shape[1])), constant_values=fill)
    trace = {"sym": np.concatenate([pad(t["sym"], -1) for t in traces]), "margin": np.concatenate([pad(t["margin"], 0) for t in traces]),
             "count": np.concatenate([t["count"] for t in traces])}
    rep = E.compare_transcripts(test, truth, trace, margin_tol=DECODE_MARGIN_TOL)
    return {"utterances": rep["utterances"], "identical_utterances": rep["identical_utterances"],
            "tokens": rep["tokens_truth"], "token_agreement": round(rep["token_agreement"], 6),
            "wer_vs_oracle": round(rep["wer_vs_truth"], 6), "differing": [(d["utterance"], d["first_differing_token"], round(d.get("margin", -1), 5)) for d in rep["differences"]][:16],
            "differences_not_at_a_near_tie": len(rep["unexplained"]), "margin_tol": DECODE_MARGIN_TOL,
            "margin_quantiles": {str(k): round(v, 5) for k, v in rep.get("margin_quantiles", {}).items()}}


def reference_arm(args):
    """`--impl reference`: the reference's CPU path (restated: its AMX/oneDNN plugin cannot be built offline, see
    DESIGN.md) timed on the host cores with all threads; every step is a bounded sample of the same workload."""
    rank, _local, world = dist_env()
    if rank != 0:
        return
    from rnnt_inference_b200 import synthetic as S
    batches, total_frames = build_query(1, 0, args.utts, args.batch)
    mean_frames = total_frames / args.utts
    sd = S.make_checkpoint()
    per_step = max(3.0, min(args.cpu_seconds, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(min(args.warmup, 1)):
        cpu_baseline(sd, batches, mean_frames, 0.0)
    print(f"[bench] reference arm: {host_threads()} host threads (OMP_NUM_THREADS in the environment was "
          f"{os.environ.get('OMP_NUM_THREADS', 'unset')}; the oracle's team size is set explicitly)", file=sys.stderr)
    vals, last = [], None
    t0 = time.perf_counter()
    for _ in range(args.steps):
        last = cpu_baseline(sd, batches, mean_frames, per_step)
        vals.append(last["value"])
    ms = (time.perf_counter() - t0) * 1e3 / max(1, args.steps)
    value = float(np.mean(vals))
    last["value"] = round(value, 3)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms, 1), "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
        "dtype": "int8", "data": "synthetic",
        "config": {"workload": f"C3 Offline: {args.utts} synthetic dev-clean-length utterances per GPU, sorted, batch {args.batch}, full int8 RNN-T greedy decode",
                   "note": "CPU restatement of the reference path (oracle/), bounded sample per step"},
        "cpu_baseline": last,
        "e2e": {"value": round(value, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------ e2e through the SUT
def make_sut(sd, batches, feats, device, batch):
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    xs, xl = [], []
    for b in batches:
        for u, n in b:
            xs.append(feats[u % UTTS_PER_GPU])
            xl.append(torch.tensor(n, dtype=torch.int64))
    sut = PytorchSUT(sd, {"x": xs, "x_lens": xl}, batch, "quant", None, enable_bf16=True, scenario="Offline", device=device)
    return sut, len(xs)


def e2e_strong(sd, args, device, world, rank, barrier, reduce_scalar):
    """End-to-end leg of --scaling strong with N > 1: ONE LoadGen-facing SUT on rank 0 (sharding.DistributedOfflineSUT).
    Rank 0 receives the Offline query of args.query samples from the LoadGen shim, broadcasts it, every rank decodes its
    share of the length-sorted order from its own pinned features through its PytorchSUT pipeline, one all_gather
    brings the transcripts back and rank 0 calls QuerySamplesComplete.  Timed on rank 0 around StartTest."""
    from rnnt_inference_b200 import loadgen_shim as lg
    from rnnt_inference_b200 import sharding, synthetic as S
    from rnnt_inference_b200.pytorch_sut import PytorchSUT
    total = args.query
    lens = S.utterance_lengths(UTTS_PER_GPU)
    ulen = [int(v) for v in lens[np.arange(total) % UTTS_PER_GPU]]
    order = np.argsort(-np.asarray(ulen), kind="stable")[rank::world]
    mine = sorted(set(int(i) for i in order))
    feats = {}
    for i in mine:
        key = i % UTTS_PER_GPU
        if key not in feats:
            feats[key] = S.make_features(ulen[i], key)
    placeholder = torch.zeros((1, 240))
    xs = [feats.get(i % UTTS_PER_GPU, placeholder) if (i % UTTS_PER_GPU) in feats else placeholder for i in range(total)]
    data = {"x": xs, "x_lens": [torch.tensor(n if (i % UTTS_PER_GPU) in feats else 1, dtype=torch.int64) for i, n in enumerate(ulen)]}
    sut = PytorchSUT(sd, data, args.batch, "quant", None, enable_bf16=True, scenario="Offline", device=device, preload=False)
    sut.qsl.load_query_samples(mine)
    collected = []
    sut.response_hook = lambda samples, tokens, counts: collected.append(([s.id for s in samples], np.array(tokens[:len(samples)], copy=True), list(counts)))

    def run_shard(samples):
        collected.clear()
        sut.issue_queries(list(samples))
        n = sum(len(c[0]) for c in collected)
        ids = torch.from_numpy(np.concatenate([np.asarray(c[0], dtype=np.int64) for c in collected]))
        cnt = torch.from_numpy(np.concatenate([np.asarray(c[2], dtype=np.int32) for c in collected]))
        tok = np.full((n, QUERY_TOKEN_CAP), -1, dtype=np.int32)
        r = 0
        for _ids, t, _c in collected:
            w = min(t.shape[1], QUERY_TOKEN_CAP)
            tok[r:r + t.shape[0], :w] = t[:, :w]
            r += t.shape[0]
        return ids.to(device), cnt.to(device), torch.from_numpy(tok).to(device)

    front = sharding.DistributedOfflineSUT(run_shard, ulen, device, l_max=QUERY_TOKEN_CAP, complete=lg.QuerySamplesComplete,
                                           make_sample=lg.QuerySample, make_response=lg.QuerySampleResponse)
    k_e2e = max(1, min(args.steps, 3))
    result = None
    if rank == 0:
        handle = lg.ConstructSUT(front.issue_queries, front.flush_queries)
        qsl = lg.ConstructQSL(total, total, lambda ids: None, lambda ids: None)
        settings = lg.TestSettings()
        settings.scenario = lg.TestScenario.Offline
        settings.min_query_count = total
        lg.StartTest(handle, qsl, settings)                 # warm-up query
        h0, d0 = sut.h2d_bytes, sut.d2h_bytes
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        done = 0
        for _ in range(k_e2e):
            done += lg.StartTest(handle, qsl, settings)["completed"]
        sec = time.perf_counter() - t0
        front.shutdown()
        result = {"value": round(done / sec, 1), "unit": UNIT, "steps": k_e2e, "query_samples": total,
                  "h2d_bytes_per_step_rank0": int((sut.h2d_bytes - h0) / k_e2e), "d2h_bytes_per_step_rank0": int((sut.d2h_bytes - d0) / k_e2e),
                  "api": "one LoadGen-facing DistributedOfflineSUT on rank 0 (loadgen_shim.StartTest); ranks shard the sorted query, one all_gather of transcripts"}
    else:
        front.serve()
    barrier()
    h2d = reduce_scalar(float(sut.h2d_bytes), dist_sum())
    d2h = reduce_scalar(float(sut.d2h_bytes), dist_sum())
    if result is not None:
        result["h2d_bytes_per_step"] = int(h2d / (k_e2e + 1))
        result["d2h_bytes_per_step"] = int(d2h / (k_e2e + 1))
    return result


def dist_sum():
    import torch.distributed as dist
    return dist.ReduceOp.SUM


def run_e2e_query(sut, count):
    """One Offline query through the LoadGen shim: host features -> transcripts on the host."""
    from rnnt_inference_b200 import loadgen_shim as lg
    settings = lg.TestSettings()
    settings.scenario = lg.TestScenario.Offline
    settings.min_query_count = count
    return lg.StartTest(sut.sut, sut.qsl.qsl, settings)


def main():
    args = parse_args()
    if args.impl == "reference":
        return reference_arm(args)
    rank, local, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    numa = None
    if world > 1 and os.environ.get("RNNT_NUMA_BIND", "1") != "0":
        from rnnt_inference_b200 import sharding as _sh
        numa = _sh.bind_to_gpu_numa_node(local)       # before any pinned allocation of this rank
    import torch.distributed as dist
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce_scalar(v, op):
        if world == 1:
            return v
        t = torch.tensor([v], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=op)
        return float(t.item())

    from rnnt_inference_b200 import ops
    sd, decoder = make_model(device, args.batch)
    strong = args.scaling == "strong"
    total = args.query if strong else None
    batches, _ = build_query(world, rank, args.utts, args.batch, total)
    n_max = ((total if strong else world * args.utts) + world - 1) // world
    feats = features_for(batches)
    staged = stage_on_device(batches, feats, device)
    my_utts = sum(len(b) for b in batches)
    my_frames = sum(n for b in batches for _, n in b)

    # ---- device-resident leg: W warm-up steps, then exactly K timed steps (no per-kernel instrumentation inside)
    streams = [torch.cuda.Stream(device) for _ in range(args.streams)] if args.streams > 1 else None
    main_stream = torch.cuda.current_stream()

    def fork():
        for st in streams or []:
            st.wait_stream(main_stream)

    def join():
        for st in streams or []:
            main_stream.wait_stream(st)

    with torch.no_grad():
        fork()
        for _ in range(args.warmup):
            run_step_device(decoder, staged, world, n_max, streams)
        join()
        ops.telemetry.events = None
        launches0 = ops.telemetry.launches
        sampler = ClockSampler(local)
        barrier()
        sampler.start()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        fork()
        tokens = None
        for _ in range(args.steps):
            tokens = run_step_device(decoder, staged, world, n_max, streams)
        join()
        ev1.record()
        barrier()
        clocks = sampler.stop()
        launches = ops.telemetry.launches - launches0
        # ---- the same step once more, untimed, with every kernel of the library bracketed by CUDA events
        ops.telemetry.events = {}
        run_step_device(decoder, staged, world, n_max)
        torch.cuda.synchronize()
        events = ops.telemetry.events
        ops.telemetry.events = None
    from rnnt_inference_b200 import _lib
    _lib.check_device_status()
    ms_total = reduce_scalar(ev0.elapsed_time(ev1), dist.ReduceOp.MAX if world > 1 else None)
    utts_total = reduce_scalar(float(my_utts), dist.ReduceOp.SUM if world > 1 else None)
    ms_per_step = ms_total / args.steps
    value = utts_total * args.steps / (ms_total * 1e-3)
    roof, step_us = recurrent_roofline(events, load_peaks())
    shares = kernel_shares(events)
    tokens_per_step = int(sum(int(t.item()) for t in tokens)) if tokens else 0

    # ---- e2e leg: PytorchSUT + LoadGen shim, host features
    e2e = None
    if not args.no_e2e and strong and world > 1:
        e2e = e2e_strong(sd, args, device, world, rank, barrier, reduce_scalar)
    elif not args.no_e2e:
        sut, count = make_sut(sd, batches, feats, device, args.batch)
        for _ in range(2):
            run_e2e_query(sut, count)
        k_e2e = max(1, min(args.steps, 5))
        h0, d0 = sut.h2d_bytes, sut.d2h_bytes
        barrier()
        t0 = time.perf_counter()
        done = 0
        for _ in range(k_e2e):
            r = run_e2e_query(sut, count)
            done += r["completed"]
        torch.cuda.synchronize()
        sec = time.perf_counter() - t0
        sec = reduce_scalar(sec, dist.ReduceOp.MAX if world > 1 else None)
        done_all = reduce_scalar(float(done), dist.ReduceOp.SUM if world > 1 else None)
        h2d = reduce_scalar(float(sut.h2d_bytes - h0), dist.ReduceOp.SUM if world > 1 else None)
        d2h = reduce_scalar(float(sut.d2h_bytes - d0), dist.ReduceOp.SUM if world > 1 else None)
        e2e = {"value": round(done_all / sec, 1), "unit": UNIT, "h2d_bytes_per_step": int(h2d / k_e2e),
               "d2h_bytes_per_step": int(d2h / k_e2e), "steps": k_e2e, "api": "PytorchSUT.issue_queries via loadgen_shim.StartTest",
               "chunks": f"sorted query cut by sharding.plan_chunks(batch {args.batch}, head_fraction {sut.head_fraction}): the first chunk's "
                         "compute covers the second chunk's host->device copy"}

    micro = encoder_microbench(decoder, device) if (rank == 0 and not args.no_microbench) else None
    cpu = parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        full, total_frames = build_query(1, 0, args.utts, args.batch)
        cpu, chunks = cpu_baseline(sd, full, total_frames / args.utts, args.cpu_seconds, keep_chunks=True)
        if not args.no_parity:
            parity = parity_check(decoder, chunks, device)
            _lib.check_device_status()
    if rank == 0:
        workload = (f"C3 Offline, strong scaling: ONE query of {total} synthetic dev-clean-length utterances for the whole job"
                    if strong else f"C3 Offline: {args.utts} synthetic dev-clean-length utterances per GPU")
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round(ms_per_step, 2), "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "int8", "data": "synthetic",
            "config": {"workload": workload + f", sorted, batch {args.batch}, full int8 RNN-T greedy decode",
                       "utterances_per_gpu_per_step": my_utts, "frames_per_gpu_per_step": my_frames, "batch": args.batch,
                       "batches_per_step": len(batches), "compute_streams": args.streams, "tokens_per_step_rank0": tokens_per_step,
                       "tokens_per_stacked_frame": round(tokens_per_step / max(1, sum((n + 1) // 2 for b in batches for _, n in b)), 4),
                       "l2": "inputs and per-layer int32 projections (up to 2 GB) exceed the 126 MB L2; no explicit flush",
                       "parallelism": f"dp{world}: length-sorted utterances dealt round-robin to ranks, transcripts gathered with one NCCL all_gather per query",
                       "numa_binding_rank0": (f"node {numa[0]}, {numa[1]} cores" if numa else None)},
            "lstm_step_us": round(step_us, 3) if step_us else None,
            "lstm_step_us_note": "recurrent kernel, mean per (layer, timestep, 128-row batch tile) item of this workload",
            "lstm_step_us_c2": micro["lstm_step_us"] if micro else None,
            "kernel_ms": shares, "kernel_ms_note": "one untimed, event-bracketed pass over the same step",
            "gpu_launches": launches, "clocks": clocks, "e2e": e2e, "roofline": roof, "cpu_baseline": cpu, "parity": parity}
        if micro:
            line["encoder_microbench"] = micro
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def encoder_microbench(decoder, device, n=256, t=500, reps=3):
    """BASELINE config 2: int8 encoder only, batch 256 x 500 frames; mean us per (layer, timestep) of the recurrent kernel."""
    from rnnt_inference_b200 import ops
    trans = decoder.rnnt.transcription
    x = torch.randn((t, n, 240), device=device).clamp_(-5.8, 5.8)
    lens = torch.full((n,), t, dtype=torch.int32, device=device)

    def run():
        hid = 1024
        z = lambda dt: [torch.zeros((n, hid), dtype=dt, device=device) for _ in range(5)]
        hx, cx = z(torch.int8), z(torch.float16)
        return trans(x, lens, hx[:2], cx[:2], hx[2:], cx[2:])

    with torch.no_grad():
        for _ in range(2):
            run()
        ops.telemetry.events = {}
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        ev0.record()
        for _ in range(reps):
            run()
        ev1.record()
        torch.cuda.synchronize()
        events = ops.telemetry.events
        ops.telemetry.events = None
    roof, item_us = recurrent_roofline(events, load_peaks())
    step_us = item_us * ((n + 127) // 128)          # one (layer, timestep) of the whole 256-utterance batch = 2 tile items
    shares = kernel_shares(events)
    gemm_ms = shares.get("gemm_i8", 0.0) / reps
    gemm_ops = 2.0 * n * (t * (4096 * 256 + 4096 * 1024) + (t // 2) * (4096 * 2048 + 2 * 4096 * 1024))
    return {"workload": f"C2 encoder only, {n} x {t} frames", "encoder_ms": round(ev0.elapsed_time(ev1) / reps, 3),
            "lstm_step_us": round(step_us, 3), "lstm_step_us_note": "recurrent kernel time / (layers x timesteps): one LSTM step of the 256-utterance batch",
            "recurrent_roofline": roof, "gemm_ms": round(gemm_ms, 3),
            "gemm_top_s": round(gemm_ops / (gemm_ms * 1e-3) / 1e12, 1) if gemm_ms else None,
            "kernel_ms": {k: round(v / reps, 3) for k, v in shares.items()}}


if __name__ == "__main__":
    main()
