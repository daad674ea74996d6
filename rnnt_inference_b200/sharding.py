"""Data-parallel sharding of an Offline query over the GPUs of one box.

Utterances are independent, so the path shards without any collective on the data path: every rank sorts the whole
query by length (descending, stable — `PytorchSUT.issue_queries`, models/pytorch_sut.py:58-65), keeps the samples
`rank, rank + world, ...` of that order — so every rank sees the same length distribution and the same amount of
work, which dealing whole batches of 1024 would not give — and cuts its share into batches (the reference's C++
harness deals sorted work to its instances likewise, csrc/torch_sut.cpp:88-122).  The only exchange is the gather of (sample id, length, tokens) at the end of a query,
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


def gather_transcripts(ids, lens, tokens, world=1, group=None):
    """ids int64 {n}, lens int32 {n}, tokens int32 {n, L} of this rank (any n, L; same device on every rank).
    Returns (ids, lens, tokens) of ALL ranks concatenated in rank order, tokens padded with -1 to the widest rank.
    One all_gather of the shapes and one per tensor; nothing else crosses the interconnect."""
    if world == 1:
        return ids, lens, tokens
    dev = ids.device
    shape = torch.tensor([tokens.shape[0], tokens.shape[1]], dtype=torch.int64, device=dev)
    shapes = [torch.empty_like(shape) for _ in range(world)]
    dist.all_gather(shapes, shape, group=group)
    n_max = int(max(s[0].item() for s in shapes))
    l_max = int(max(s[1].item() for s in shapes))

    def padded(t, fill, cols=None):
        out_shape = (n_max,) if cols is None else (n_max, cols)
        out = torch.full(out_shape, fill, dtype=t.dtype, device=dev)
        if cols is None:
            out[:t.shape[0]] = t
        else:
            out[:t.shape[0], :t.shape[1]] = t
        return out

    parts = []
    for t, fill, cols in ((ids, -1, None), (lens, 0, None), (tokens, -1, l_max)):
        mine = padded(t, fill, cols)
        bufs = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(bufs, mine, group=group)
        parts.append(torch.cat([b[:int(s[0].item())] for b, s in zip(bufs, shapes)], 0))
    return tuple(parts)
