"""N > 1 host logic on CPU (gloo, world_size 2): the query is partitioned without overlap, every rank's batches are sorted
by length, and the transcript gather returns every sample exactly once with its own tokens."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from rnnt_inference_b200 import sharding


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, lengths, batch, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = sharding.shard_batches(lengths, batch, world, rank)
        ids = torch.tensor([i for b in mine for i, _ in b], dtype=torch.int64)
        lens = torch.tensor([n % 7 for b in mine for _, n in b], dtype=torch.int32)       # fake transcript lengths
        width = int(lens.max().item()) if len(lens) else 1
        tokens = torch.full((len(ids), width), -1, dtype=torch.int32)
        for r, (i, n) in enumerate(zip(ids.tolist(), lens.tolist())):
            tokens[r, :n] = torch.arange(n, dtype=torch.int32) + i                        # tokens depend on the sample id
        g_ids, g_lens, g_tok = sharding.gather_transcripts(ids, lens, tokens, world)
        # the single-collective form: every rank knows the shard sizes and the token width from the query itself
        n_max = (len(lengths) + world - 1) // world
        h_ids, h_lens, h_tok = sharding.gather_transcripts(ids, lens, tokens, world, n_max=n_max, l_max=6)
        assert torch.equal(g_ids, h_ids) and torch.equal(g_lens, h_lens) and torch.equal(g_tok[:, :6], h_tok[:, :g_tok.shape[1]])
        torch.save({"mine": mine, "ids": g_ids, "lens": g_lens, "tok": g_tok}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_shard_and_gather_world2(tmp_path):
    rng = np.random.default_rng(5)
    lengths = rng.integers(40, 500, 203)
    batch, world = 32, 2
    mp.spawn(_worker, args=(world, _free_port(), lengths, batch, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(os.path.join(tmp_path, f"r{r}.pt")) for r in range(world)]
    seen = sorted(i for o in outs for b in o["mine"] for i, _ in b)
    assert seen == list(range(len(lengths)))                                              # disjoint cover of the query
    for o in outs:
        for b in o["mine"]:
            ls = [n for _, n in b]
            assert ls == sorted(ls, reverse=True) and len(b) <= batch
    for o in outs:                                                                        # every rank holds the full gather
        assert sorted(o["ids"].tolist()) == list(range(len(lengths)))
        for i, n, row in zip(o["ids"].tolist(), o["lens"].tolist(), o["tok"]):
            assert n == int(lengths[i]) % 7
            assert row[:n].tolist() == list(range(i, i + n)) and bool((row[n:] == -1).all())


def test_shard_single_rank_matches_issue_queries_order():
    lengths = [5, 9, 9, 2, 7]
    b = sharding.shard_batches(lengths, 2)
    assert [[i for i, _ in x] for x in b] == [[1, 2], [4, 0], [3]]                        # stable descending sort


class _Sample:
    def __init__(self, id, index):
        self.id, self.index = id, index


def _front_worker(rank, world, port, lengths, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ran = []

        def run_shard(samples):            # stands in for the GPU pipeline: tokens derive from the sample index
            ran.append([s.index for s in samples])
            ids = torch.tensor([s.id for s in samples], dtype=torch.int64)
            lens = torch.tensor([s.index % 5 for s in samples], dtype=torch.int32)
            tok = torch.full((len(samples), 4), -1, dtype=torch.int32)
            for r, s in enumerate(samples):
                tok[r, :s.index % 5] = torch.arange(s.index % 5, dtype=torch.int32) + 10 * s.index
            return ids, lens, tok

        answered = {}

        def complete(responses):
            import ctypes
            for rid, ptr, size in responses:
                answered[rid] = np.frombuffer(ctypes.string_at(ptr, size), np.int32).tolist() if size else []

        front = sharding.DistributedOfflineSUT(run_shard, lengths, torch.device("cpu"), l_max=4, complete=complete,
                                               make_sample=_Sample, make_response=lambda i, p, n: (i, p, n))
        if rank == 0:
            for q in range(2):          # two queries through the same front
                samples = [_Sample(1000 * q + k, (7 * k + q) % len(lengths)) for k in range(37)]
                front.issue_queries(samples)
                for s in samples:
                    n = s.index % 5
                    assert answered[s.id] == [10 * s.index + j for j in range(n)], (s.id, s.index, answered[s.id])
            front.shutdown()
        else:
            front.serve()
        torch.save({"ran": ran}, os.path.join(out_dir, f"f{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_distributed_offline_front_world2(tmp_path):
    """One LoadGen-facing front on rank 0: the query is broadcast, every rank decodes its share of the length-sorted
    order, one gather brings the transcripts back and rank 0 answers every sample with its own tokens."""
    lengths = [int(v) for v in np.random.default_rng(3).integers(40, 500, 50)]
    world = 2
    mp.spawn(_front_worker, args=(world, _free_port(), lengths, str(tmp_path)), nprocs=world, join=True)
    outs = [torch.load(os.path.join(tmp_path, f"f{r}.pt")) for r in range(world)]
    for q in range(2):
        shares = [o["ran"][q] for o in outs]
        assert sorted(shares[0] + shares[1]) == sorted((7 * k + q) % 50 for k in range(37))      # disjoint cover of the query
        for sh in shares:
            ls = [lengths[i] for i in sh]
            assert ls == sorted(ls, reverse=True)                                                   # each share longest first
        assert abs(len(shares[0]) - len(shares[1])) <= 1


def test_plan_chunks_covers_the_query_and_respects_the_batch_size():
    """The pipelined SUT's chunk plan: every sample exactly once, no chunk above batch_size, shortest chunk first, plain
    chunks when the short end is too small to matter, head of the short end = the requested share of its frames."""
    rng = np.random.default_rng(5)
    for n, bs in ((1, 4), (2, 4), (3, 2), (10, 4), (2513, 2048), (4096, 2048), (24576, 2048), (5000, 1024)):
        lens = np.sort(rng.integers(1, 500, n))[::-1]
        for hf, mf in ((0.0, 0), (0.2, 0), (0.2, 65536), (0.5, 0)):
            plan = sharding.plan_chunks(list(lens), bs, hf, mf)
            seen = np.zeros(n, np.int32)
            for lo, hi in plan:
                assert 0 <= lo < hi <= n and hi - lo <= bs
                seen[lo:hi] += 1
            assert (seen == 1).all()
            assert plan[0][1] == n                                   # the shortest samples run first
            if hf == 0.0:
                assert plan == [(i, min(i + bs, n)) for i in range(0, n, bs)][::-1]
    lens = np.sort(rng.integers(50, 500, 2513))[::-1]
    plan = sharding.plan_chunks(list(lens), 2048, 0.2, 0)
    assert len(plan) == 2
    head = lens[plan[0][0]:plan[0][1]].sum() / lens.sum()
    assert 0.2 <= head < 0.21
    assert sharding.plan_chunks([30, 20, 10], 4, 0.2, 65536) == [(0, 3)]      # tiny query: plain
