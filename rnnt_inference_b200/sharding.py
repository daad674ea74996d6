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
