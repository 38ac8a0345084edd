"""GPU x2 (NCCL): row-sharded scan -> all-gather -> merge+route equals the unsharded oracle.
Skipped on a single-GPU box."""
import os
import socket

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    from oracle import qre_oracle as O
    from qre_b200 import _cabi, synth
    from qre_b200.sharded import ShardedLookup
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    w = synth.make_workload(30000, 300, 384, n_centroids=8, seed=44, frac_dup=0.03, frac_near=0.05)
    w.rows[1::2][:60] = w.rows[0::2][:60]                 # cross-shard duplicates: tie-break via lru
    w.row_ctx[1::2][:60] = w.row_ctx[0::2][:60]
    idx = np.arange(rank, w.rows.shape[0], world)
    sh = _cabi.Shard(dim=384, capacity=idx.size, max_batch=300, device=rank)
    sh.append(w.rows[idx], w.row_ctx[idx], w.row_ts_us[idx], w.row_lru[idx], w.row_pred_dev[idx],
              w.row_pred_conf[idx])
    sh.set_centroids(w.centroids, w.owner)
    cfg = _cabi.RouteConfig(ttl_us=w.ttl_us)
    look = ShardedLookup(sh, cfg, rank, world)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)   # noqa: E731
    out = look.step(t(w.queries), t(w.q_ctx.view(np.int64)), w.now_us, t(w.tok), t(w.heur_dev), t(w.heur_conf),
                    check_overflow=True)
    torch.cuda.synchronize()
    slot, shard = out["hit_slot"].cpu().numpy(), out["hit_shard"].cpu().numpy()
    glob = np.where(slot >= 0, slot * world + shard, -1)
    img = O.CacheImage(w.rows, w.row_ctx, w.row_ts_us, w.row_lru)
    exp_row, exp_sim = O.scan_batch(w.queries, w.q_ctx, img, w.now_us, w.ttl_us, 0.85)
    hit = exp_row >= 0
    r = O.route_batch(w.queries, w.tok, w.heur_dev, w.heur_conf, w.centroids, w.owner, O.RouteParams())
    ok = (np.array_equal(glob, exp_row)
          and np.allclose(out["hit_sim"].cpu().numpy()[hit], exp_sim[hit], rtol=1e-5)
          and np.array_equal(out["hit_pred_dev"].cpu().numpy()[hit], w.row_pred_dev[exp_row[hit]])
          and float(np.mean(out["dev"].cpu().numpy() == r["dev"])) > 0.995)
    # pipelined: three different batches in flight back to back must each equal the serial result
    perms = [np.random.Generator(np.random.PCG64(s_)).permutation(300) for s_ in range(3)]
    serial = []
    for pm in perms:
        o = look.step(t(w.queries[pm]), t(w.q_ctx[pm].view(np.int64)), w.now_us, t(w.tok[pm]), t(w.heur_dev[pm]),
                      t(w.heur_conf[pm]))
        torch.cuda.synchronize()
        serial.append({k: v.cpu().numpy().copy() for k, v in o.items()})
    ins = [(t(w.queries[pm]), t(w.q_ctx[pm].view(np.int64)), t(w.tok[pm]), t(w.heur_dev[pm]), t(w.heur_conf[pm]))
           for pm in perms]
    torch.cuda.synchronize()
    handles = [look.step_pipelined(a, b, w.now_us, c, d, e) for a, b, c, d, e in ins]
    # with 3 batches, batch 0's buffers were reused by batch 2; check batches 1 and 2
    for i in (1, 2):
        o, ev = handles[i]
        ev.synchronize()
        for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf", "score", "sem_method"):
            ok = ok and np.array_equal(o[k].cpu().numpy(), serial[i][k])
    # peer-memory exchange instead of the NCCL all-gather: same answers, several batches in a row
    look_p = ShardedLookup(sh, cfg, rank, world, exchange="p2p")
    for i, pm in enumerate(perms + perms):
        a, b, c, d, e = ins[i % 3]
        o = look_p.step(a, b, w.now_us, c, d, e)
        torch.cuda.synchronize()
        for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf", "score", "sem_method", "hit_pred_dev"):
            ok = ok and np.array_equal(o[k].cpu().numpy(), serial[i % 3][k])
    # back to back without host synchronisation (the two-slot protocol must hold up)
    for i in range(12):
        a, b, c, d, e = ins[i % 3]
        o = look_p.step(a, b, w.now_us, c, d, e)
    torch.cuda.synchronize()
    for k in ("hit_slot", "hit_sim", "dev"):
        ok = ok and np.array_equal(o[k].cpu().numpy(), serial[11 % 3][k])
    # pipelined + peer exchange: publish and merge on the side stream under the next scan
    handles = [look_p.step_pipelined(a, b, w.now_us, c, d, e) for a, b, c, d, e in ins]
    for i in (1, 2):
        o, ev = handles[i]
        ev.synchronize()
        for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf"):
            ok = ok and np.array_equal(o[k].cpu().numpy(), serial[i][k])
    torch.cuda.synchronize()
    # sliced results: each rank merges + routes only its own half of the batch; that half must equal
    # the replicated answer, the other half of the output arrays stays untouched
    assert look_p.enable_slices(True)
    q0, q1 = look_p.slice_range(300)
    ok = ok and (q0, q1) == ((0, 150) if rank == 0 else (150, 300))
    handles = [look_p.step_pipelined(a, b, w.now_us, c, d, e) for a, b, c, d, e in ins + ins]
    for i in (4, 5):
        o, ev = handles[i]
        ev.synchronize()
        for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf", "score", "sem_method", "hit_pred_dev"):
            ok = ok and np.array_equal(o[k][q0:q1].cpu().numpy(), serial[i % 3][k][q0:q1])
    torch.cuda.synchronize()
    # the host-buffer call on the sharded cache: every rank passes the whole batch, uploads only its
    # slice (the ranks exchange the query slices over NVLink) and gets its slice of the results back
    import copy
    cfg_s = copy.copy(cfg)
    cfg_s.slice_results = 1
    for rep in range(2):
        for sl_, pm in enumerate(perms[:2]):
            sh.lookup_route_begin(sl_, w.queries[pm], w.q_ctx[pm], cfg_s, w.now_us, tok=w.tok[pm],
                                  heur_dev=w.heur_dev[pm], heur_conf=w.heur_conf[pm])
        for sl_ in range(2):
            r_ = sh.lookup_route_end(sl_)
            for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf", "score", "sem_method"):
                ok = ok and np.array_equal(r_[k][q0:q1], serial[sl_][k][q0:q1])
    # ... and with replicated results (every rank uploads and returns everything)
    sh.lookup_route_begin(0, w.queries[perms[2]], w.q_ctx[perms[2]], cfg, w.now_us, tok=w.tok[perms[2]],
                          heur_dev=w.heur_dev[perms[2]], heur_conf=w.heur_conf[perms[2]])
    r_ = sh.lookup_route_end(0)
    for k in ("hit_slot", "hit_sim", "hit_shard", "dev", "conf"):
        ok = ok and np.array_equal(r_[k], serial[2][k])
    look_p.enable_slices(False)
    dist.barrier()
    # data-parallel embedding in front of the lookup: each rank embeds half of the sequences, one all-gather,
    # and the [m, hidden] matrix equals what a single encoder call over the whole batch produces, bit for bit
    from oracle import minilm_oracle as MO
    eo = MO.EncoderConfig(vocab_size=500, hidden=384, layers=1, heads=12, intermediate=384, max_positions=64,
                          max_seq_len=48)
    enc = _cabi.Encoder(_cabi.EncoderConfig(500, 384, 1, 12, 384, 64, 2, 1e-12, 48), MO.random_weights(eo, seed=9),
                        max_tokens=4096, max_seqs=64, device=rank)
    rs = np.random.Generator(np.random.PCG64(12))
    lens = rs.integers(1, 48, 61)                                    # 61: world does not divide it
    seqs = [rs.integers(0, 500, n).astype(np.int32) for n in lens]
    off = np.zeros(62, np.int32)
    off[1:] = np.cumsum(lens)
    whole = enc.encode_ids(seqs)
    for _ in range(2):
        q_all = look.embed_ids(enc, np.concatenate(seqs), off)
        torch.cuda.synchronize()
        ok = ok and np.array_equal(q_all.cpu().numpy(), whole)
    want = MO.encode_ids(MO.random_weights(eo, seed=9), seqs[:4], eo)
    ok = ok and float((whole[:4] * want).sum(1).min()) > 0.999
    enc.close()
    ret[rank] = (bool(ok), int(hit.sum()))
    dist.barrier()
    dist.destroy_process_group()
    sh.close()


def test_two_gpu_sharded_lookup_equals_oracle():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import torch.multiprocessing as mp
    world, port = 2, _free_port()
    with mp.Manager() as man:
        ret = man.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert all(ret[r][0] for r in range(world)), dict(ret)
        assert ret[0][1] > 50
