"""Data-parallel sharding of an Offline query over the GPUs of one box.

Utterances are independent, so the path shards without any collective on the data path: every rank sorts the whole
query by length (descending, stable — `PytorchSUT.issue_queries`, models/pytorch_sut.py:58-65), keeps the samples
`rank, rank + world, ...` of that order — so every rank sees the same length distribution and the same amount of
work, which dealing whole batches of 1024 would not give — and cuts its share into batches (the reference's C++
harness deals sorted work to its instances likewise, csrc/torch_sut.cpp:88-122).  The only exchange is ONE all_gather of (sample id, length, tokens) at the end of a query,
O(100 KB): `gather_transcripts` (NCCL for CUDA tensors, gloo for CPU tensors).
"""
import numpy as np
import torch
import torch.distributed as dist


def bind_to_gpu_numa_node(device_index):
    """One process per GPU: run this process on the CPU cores of the NUMA node its GPU hangs off, BEFORE it allocates its
    page-locked feature arena, so that the arena lands in that node's memory (first touch) and the per-batch host->device
    copies of the 8 ranks do not all cross the socket interconnect.  Returns (node, number of cores) or None when the
    topology cannot be read (then nothing changes)."""
    import os
    try:
        props = torch.cuda.get_device_properties(device_index)
        bdf = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read().strip())
        if node < 0:
            return None
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node, len(cpus)
    except (OSError, ValueError, AttributeError, RuntimeError):
        return None


def plan_chunks(lens_desc, batch_size, head_fraction=0.15, min_frames=65536):
    """Chunks of one pipelined Offline query, as (lo, hi) ranges into the length-sorted (descending) sample list, in the
    order they RUN (shortest first).

    Plain chunking — the reference's `samples[i:i + batch_size]` (models/pytorch_sut.py:62), run back to front — leaves
    one host->device copy that nothing hides: the copy of the first full batch behind the compute of the short
    remainder (2513 utterances at batch 2048: 9 ms of copy behind 4 ms of compute).  Copying a frame costs about a
    quarter of computing it, so the short end of the query — the remainder plus the last full batch — is cut instead
    where the head (run first; its own copy is the exposed one) holds `head_fraction` of that region's frames: its
    compute then covers the copy of the rest of the region, and every later full batch is covered by its predecessor
    (measured on B200, 2513 utterances: 53.1 K utterances/s end to end with plain chunks, 56.0 K for 0.12 - 0.18, 53.5 K
    at 0.25 where the smaller batches cost more device time than the copy they hide).
    Regions with fewer than `min_frames` frames (copy under a millisecond) and head_fraction <= 0 keep the plain chunks.
    Chunks are independent, so the plan changes no transcript except through the reference's batch-scoped quirk (a lane
    capped on its last frame emits until its batch's loop ends), exactly as the choice of batch_size does."""
    n = len(lens_desc)
    if n == 0:
        return []
    plain = [(i, min(i + batch_size, n)) for i in range(0, n, batch_size)][::-1]
    if head_fraction <= 0 or n < 2:
        return plain
    rem = n % batch_size
    region = rem + batch_size if n > batch_size else n          # the shortest `region` samples: [n - region, n)
    if rem == 0 and n >= batch_size:
        region = batch_size
    lo = n - region
    frames = np.asarray(lens_desc[lo:], dtype=np.int64)
    total = int(frames.sum())
    if total < min_frames:
        return plain
    # head = the shortest samples whose frames reach head_fraction of the region; both parts within batch_size
    tail_cum = np.cumsum(frames[::-1])
    k = int(np.searchsorted(tail_cum, head_fraction * total)) + 1
    k = max(k, region - batch_size, 1)
    k = min(k, batch_size, region - 1)
    full = [(i, i + batch_size) for i in range(0, lo, batch_size)][::-1]
    return [(n - k, n), (lo, n - k)] + full


def shard_batches(lengths, batch_size, world=1, rank=0):
    """lengths: per-sample feature lengths of the WHOLE query.  Returns this rank's batches, each a list of
    (sample_index, length), in the order they are to be run (longest first)."""
    lengths = np.asarray(lengths)
    order = np.argsort(-lengths, kind="stable")[rank::world]
    batches = [order[i:i + batch_size] for i in range(0, len(order), batch_size)]
    return [[(int(i), int(lengths[i])) for i in b] for b in batches]


def gather_transcripts(ids, lens, tokens, world=1, group=None, n_max=None, l_max=None):
    """ids int64 {n}, lens int32 {n}, tokens int32 {n, L} of this rank (same device on every rank).
    Returns (ids, lens, tokens) of ALL ranks concatenated in rank order, tokens padded with -1.

    ONE collective: every rank packs its rows into an int32 buffer [n_max, 3 + l_max] = (id low word, id high word,
    length, tokens...) and a single all_gather moves it (one gather per query, O(100 KB), latency-bound).
    n_max / l_max (rows per rank, token columns) are known to every rank from the query itself — the shard sizes and the
    longest utterance — so no shape exchange is needed; when the caller does not pass them they are agreed on with one
    extra small all_gather.  Rows beyond a rank's own count carry id -1 and are dropped."""
    if world == 1:
        return ids, lens, tokens
    dev = ids.device
    if n_max is None or l_max is None:
        shape = torch.tensor([tokens.shape[0], tokens.shape[1]], dtype=torch.int64, device=dev)
        shapes = [torch.empty_like(shape) for _ in range(world)]
        dist.all_gather(shapes, shape, group=group)
        n_max = int(max(s[0].item() for s in shapes))
        l_max = int(max(s[1].item() for s in shapes))
    n, width = tokens.shape[0], min(tokens.shape[1], l_max)
    if n > n_max:
        raise RuntimeError(f"gather_transcripts: {n} rows on this rank but n_max = {n_max}")
    mine = torch.full((n_max, 3 + l_max), -1, dtype=torch.int32, device=dev)
    mine[:n, 0:2] = ids.to(torch.int64).view(torch.int32).view(n, 2)
    mine[:n, 2] = lens.to(torch.int32)
    mine[:n, 3:3 + width] = tokens[:, :width]
    rows = torch.empty((world * n_max, 3 + l_max), dtype=torch.int32, device=dev)     # rank-major concatenation
    dist.all_gather_into_tensor(rows, mine, group=group)
    all_ids = rows[:, 0:2].contiguous().view(torch.int64).view(-1)
    keep = all_ids >= 0
    return all_ids[keep], rows[keep, 2].contiguous(), rows[keep, 3:].contiguous()


class DistributedOfflineSUT:
    """ONE LoadGen-facing SUT for the whole box (the reference's C++ harness owns all its instances behind
    one SUT likewise, csrc/torch_sut.cpp:88-122).  Built on every rank over the same dataset:
      * rank 0 receives the Offline query from LoadGen, broadcasts its (sample id, QSL index) pairs;
      * every rank sorts the query by length exactly like PytorchSUT.issue_queries and decodes the samples
        rank, rank + world, ... of that order on its own GPU (`run_shard`);
      * ONE all_gather brings (sample id, length, tokens) of all ranks together and rank 0 answers LoadGen.
    Ranks other than 0 sit in `serve()` until rank 0 calls `shutdown()`.

    run_shard(samples) -> (ids int64 {n}, lens int32 {n}, tokens int32 {n, L}) on `device`; `lengths[index]` are the
    feature lengths of the QSL (the sort key); `complete(responses)` defaults to lg.QuerySamplesComplete."""

    def __init__(self, run_shard, lengths, device, l_max=256, complete=None, make_sample=None, make_response=None):
        self.run_shard, self.lengths, self.device, self.l_max = run_shard, np.asarray(list(lengths), dtype=np.int64), device, int(l_max)
        self.rank, self.world = dist.get_rank(), dist.get_world_size()
        self.complete, self.make_sample, self.make_response = complete, make_sample, make_response
        self.queries = 0

    def _round(self, pairs):
        """pairs: int64 {n, 2} (id, index) on rank 0, None elsewhere.  Returns the gathered (ids, lens, tokens)."""
        head = torch.tensor([pairs.shape[0] if self.rank == 0 else 0], dtype=torch.int64, device=self.device)
        dist.broadcast(head, 0)
        n = int(head.item())
        if n == 0:
            return None
        if self.rank != 0:
            pairs = torch.empty((n, 2), dtype=torch.int64, device=self.device)
        dist.broadcast(pairs, 0)
        host = pairs.cpu().numpy()
        lens = self.lengths[host[:, 1]]
        mine = np.argsort(-lens, kind="stable")[self.rank::self.world]
        samples = [self.make_sample(i, x) for i, x in zip(host[mine, 0].tolist(), host[mine, 1].tolist())]
        ids, out_lens, tokens = self.run_shard(samples)
        self.queries += 1
        return gather_transcripts(ids, out_lens, tokens, self.world, n_max=(n + self.world - 1) // self.world, l_max=self.l_max)

    def issue_queries(self, samples):
        """LoadGen entry point (rank 0 only)."""
        assert self.rank == 0
        # per-sample Python is the front's cost at 24 576 samples per query: plain comprehensions and array arithmetic only
        pairs = np.empty((len(samples), 2), dtype=np.int64)
        pairs[:, 0] = [s.id for s in samples]
        pairs[:, 1] = [s.index for s in samples]
        ids, lens, tokens = self._round(torch.from_numpy(pairs).to(self.device))
        ids, lens = ids.cpu().numpy(), lens.cpu().numpy()
        tokens = np.ascontiguousarray(tokens.cpu().numpy())
        base, stride = tokens.ctypes.data, tokens.strides[0]
        ptrs = (base + stride * np.arange(len(ids), dtype=np.int64)).tolist()
        sizes = (np.minimum(lens, tokens.shape[1]).astype(np.int64) * 4).tolist()
        make = self.make_response
        self.complete([make(i, p, z) for i, p, z in zip(ids.tolist(), ptrs, sizes)])
        self._keep = tokens          # the responses point into it until LoadGen has copied them

    def flush_queries(self):
        pass

    def serve(self):
        """Ranks != 0: take part in every query until shutdown."""
        assert self.rank != 0
        while self._round(None) is not None:
            pass

    def shutdown(self):
        if self.rank == 0:
            self._round(torch.empty((0, 2), dtype=torch.int64, device=self.device))
