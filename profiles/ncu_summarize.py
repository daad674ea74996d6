"""Turns the artefacts of profiles/ncu_capture.sh into the tracked summaries.
Usage: python profiles/ncu_summarize.py [gpurun_out] [profiles/r02] [suffix]
  launches.csv            -> <out>/launches<suffix>.csv (copy) + launches<suffix>_summary.md (time and share per kernel)
  prof_<k>.ncu-rep        -> <out>/ncu_raw_summary<suffix>.md (the metrics DESIGN.md cites, read with `ncu -i ... --page raw --csv`)"""
import csv
import io
import os
import shutil
import subprocess
import sys
from collections import OrderedDict

src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
out = sys.argv[2] if len(sys.argv) > 2 else "profiles/r02"
suffix = sys.argv[3] if len(sys.argv) > 3 else "_final"

METRICS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_tag_requests.avg.pct_of_peak_sustained_elapsed", "lts__t_requests_srcunit_tex_op_read.sum",
    "lts__t_requests_srcunit_tex_op_write.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
    "sm__mem_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_tc_wavefronts_mem_shared.sum",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__registers_per_thread", "smsp__inst_executed.sum", "sm__cycles_elapsed.max", "launch__grid_size",
    "launch__block_size", "launch__cluster_size",
]

lc = os.path.join(src, "launches.csv")
if os.path.exists(lc):
    shutil.copy(lc, os.path.join(out, "launches%s.csv" % suffix))
    lines = [l for l in open(lc) if l.startswith('"')]
    rows = list(csv.reader(io.StringIO("".join(lines))))
    hdr, rows = rows[0], rows[1:]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = OrderedDict()
    for r in rows:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
        name = r[ki].split("(")[0][:110]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(out, "launches%s_summary.md" % suffix), "w") as f:
        f.write("# ncu launch list of `bench.py --steps 1 --warmup 1` (three passes over the step: warm-up, timed, instrumented; "
                "cold-cache, serialised: compare shares)\n\nlaunches %d, total %.2f ms\n\n| kernel | launches | ms | share |\n|---|---|---|---|\n"
                % (len(rows), tot))
        for k, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write("| `%s` | %d | %.3f | %.1f %% |\n" % (k, n, ms, 100 * ms / tot))
    print("launch list:", len(rows), "launches", round(tot, 2), "ms")

with open(os.path.join(out, "ncu_raw_summary%s.md" % suffix), "w") as f:
    f.write("# ncu --set full captures, one launch each inside `bench.py --steps 1 --warmup 1` (profiles/ncu_capture.sh)\n\n"
            "`prof_rec`: lstm_recurrent_kernel<1>, first pre-RNN layer of the 2048-utterance batch (T = 500, ragged tiles); `prof_dec`: "
            "greedy_decode_kernel<true>, 2048 lanes; `prof_gemm`: the projection GEMM of the second pre-RNN layer (K = 1024) of that "
            "batch; `prof_gemm_bf16`: its hoisted joint product.  Raw `.ncu-rep` files: gpurun_out/ (scratch).\n")
    for k in ("rec", "dec", "gemm", "gemm_bf16"):
        rep = os.path.join(src, "prof_%s.ncu-rep" % k)
        if not os.path.exists(rep):
            continue
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        if len(rows) < 3:
            continue
        hdr, units, vals = rows[0], rows[1], rows[2]
        name = vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else k
        f.write("\n## prof_%s: `%s`\n\n" % (k, name[:70]))
        for m in METRICS:
            if m in hdr:
                i = hdr.index(m)
                f.write("* `%s` = %s %s\n" % (m, vals[i], units[i]))
print("written", out)
