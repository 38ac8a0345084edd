"""CPU (gloo, world_size 2): the N>1 host-side plumbing — row sharding, global<->local slot
mapping, the all-gather layout [world, m, 8] of 32-byte candidate records and the merge
comparator (sim desc, lru_seq asc) — with the oracle standing in for each rank's GPU scan."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import qre_oracle as O
from qre_b200 import _cabi, sharded, synth


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _local_candidates(w, rank, world):
    """What qre_scan_dev produces on one rank: its rows' exactly scored top-8 per query."""
    idx = np.arange(rank, w.rows.shape[0], world)
    img = O.CacheImage(w.rows[idx], w.row_ctx[idx], w.row_ts_us[idx], w.row_lru[idx])
    _, _, top_rows, top_sims = O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, 0.85, topk=_cabi.TOPK)
    cand = np.zeros((w.queries.shape[0], _cabi.TOPK), _cabi.CAND_DTYPE)
    cand["slot"], cand["sim"] = top_rows, top_sims
    cand["lru_seq"] = np.where(top_rows >= 0, w.row_lru[idx][np.maximum(top_rows, 0)], np.iinfo(np.uint64).max)
    cand["shard"] = rank
    return cand


def _worker(rank, world, port, ret):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    w = synth.make_workload(3000, 96, 384, n_centroids=2, seed=33, frac_dup=0.05, frac_near=0.05)
    # duplicate some rows ACROSS shards so the tie-break must survive the merge
    w.rows[1::2][:40] = w.rows[0::2][:40]
    w.row_ctx[1::2][:40] = w.row_ctx[0::2][:40]
    mine = torch.from_numpy(_local_candidates(w, rank, world).view(np.uint8).reshape(-1))
    gathered = torch.empty(world * mine.numel(), dtype=torch.uint8)
    dist.all_gather_into_tensor(gathered, mine)
    cand_all = gathered.numpy().view(_cabi.CAND_DTYPE).reshape(world, w.queries.shape[0], _cabi.TOPK)
    merged = sharded.merge_candidates_host(cand_all)
    hit = merged["slot"][:, 0] >= 0
    glob = np.where(hit, merged["slot"][:, 0] * world + merged["shard"][:, 0], -1)
    img = O.CacheImage(w.rows, w.row_ctx, w.row_ts_us, w.row_lru)
    exp_row, exp_sim = O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, 0.85)
    ok = bool(np.array_equal(glob, exp_row)) and bool(np.allclose(merged["sim"][hit, 0], exp_sim[hit], rtol=1e-6))
    ret[rank] = (ok, int(hit.sum()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_gather_merge_equals_unsharded():
    world, port = 2, _free_port()
    with mp.Manager() as man:
        ret = man.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert all(ret[r][0] for r in range(world)), dict(ret)
        assert ret[0][1] > 20


def test_merge_comparator_ties_and_empties():
    c = np.zeros((2, 1, _cabi.TOPK), _cabi.CAND_DTYPE)
    c["slot"] = -1
    c["sim"] = -np.inf
    c[0, 0, 0] = (0.9, 0, 5, 70, 1.0)
    c[1, 0, 0] = (0.9, 1, 9, 30, 1.0)        # same score, earlier in LRU order -> wins
    c[1, 0, 1] = (0.95, 1, 2, 99, 1.0)
    m = sharded.merge_candidates_host(c)
    assert m["slot"][0, :3].tolist() == [2, 9, 5] and m["shard"][0, :3].tolist() == [1, 1, 0]
    assert m["slot"][0, 3] == -1


def test_split_sequences_covers_the_batch_in_equal_slices():
    """Partition used by the data-parallel embedder in front of the sharded lookup (ShardedLookup.embed_ids)."""
    from qre_b200.sharded import ShardedLookup
    for m in (0, 1, 3, 61, 4096, 4097):
        for world in (1, 2, 3, 4, 8):
            ranges, per = ShardedLookup.split_sequences(m, world)
            assert len(ranges) == world and per * world >= m
            assert ranges[0][0] == 0 and ranges[-1][1] == m
            assert all(a1 == b0 for (_, a1), (b0, _) in zip(ranges, ranges[1:]))
            assert all(0 <= i1 - i0 <= per for i0, i1 in ranges)
            assert all(i0 == min(m, r * per) for r, (i0, _) in enumerate(ranges))     # slice r starts at r * per
