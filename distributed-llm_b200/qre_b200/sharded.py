"""Row-sharded lookup+route over the GPUs of one box.

One process per GPU.  Each rank owns a `Shard` holding rows {r : r % world == rank} or any
other disjoint subset; queries and per-query features are replicated.  Per batch:

    scan_dev (local tcgen05 scan + exact re-score)  ->  cand[m, 8]  (32-byte records)
    all_gather_into_tensor over NCCL / NVLink       ->  cand_all[world, m, 8]
    merge_route_dev (comparator (sim desc, lru_seq asc) + routing epilogue) on every rank

torch is plumbing here: device buffers, streams and the NCCL collective.  With world == 1 the
collective is skipped.  CPU tests drive `merge_candidates_host` over gloo instead.

`step_pipelined` overlaps everything that is not the scan with the scans of the neighbouring
batches.  With the peer exchange (or one GPU) that is the library's own pipeline (`qre_submit_dev`):
query preparation of batch i+1 and re-score / publish / merge + route of batch i-1 run on the
handle's auxiliary streams beside the scan of batch i.  With the NCCL exchange the all-gather +
merge of batch i run on a side stream here.  The scan kernel is persistent (one CTA per SM, all
shared memory), so it leaves `reserve_sms` SMs free for those kernels; candidate buffers, outputs
and the handle's per-batch state are double-buffered (`pipeline_slot`).
"""
from __future__ import annotations

from typing import Optional

import numpy as np

from . import _cabi

RESULT_SPECS = {            # field -> (torch dtype name, trailing shape symbol)
    "hit_slot": ("int64", ()), "hit_sim": ("float32", ()), "hit_pred_dev": ("uint8", ()),
    "hit_pred_conf": ("float64", ()), "hit_shard": ("int32", ()), "centroid_sim": ("float32", ("K",)),
    "tok_dev": ("uint8", ()), "tok_conf": ("float64", ()), "sem_dev": ("uint8", ()),
    "sem_conf": ("float64", ()), "sem_method": ("uint8", ()), "dev": ("uint8", ()),
    "conf": ("float64", ()), "score": ("float64", ("D",)), "flags": ("uint8", ()),
}


def merge_candidates_host(cand_all: np.ndarray) -> np.ndarray:
    """Host restatement of the merge comparator, for CPU (gloo) tests of the N>1 plumbing.
    cand_all: structured [world, m, 8] -> merged [m, 8], best first; (sim desc, lru_seq asc)."""
    world, m, k = cand_all.shape
    out = np.empty((m, k), _cabi.CAND_DTYPE)
    flat = np.transpose(cand_all, (1, 0, 2)).reshape(m, world * k)
    for q in range(m):
        c = flat[q]
        valid = c["slot"] >= 0
        order = np.lexsort((np.arange(c.size), c["lru_seq"], -c["sim"].astype(np.float64), ~valid))
        out[q] = c[order[:k]]
    return out


class _SlotDone:
    """Completion of one pipelined batch (the handle's post-stream event for that slot)."""

    def __init__(self, shard, slot):
        self.shard, self.slot = shard, slot

    def synchronize(self):
        self.shard.wait_dev(self.slot, 0)

    def wait(self, stream):
        self.shard.wait_dev(self.slot, stream.cuda_stream or 1)


class _EventDone:
    def __init__(self, ev):
        self.ev = ev

    def synchronize(self):
        self.ev.synchronize()

    def wait(self, stream):
        stream.wait_event(self.ev)


class ShardedLookup:
    def __init__(self, shard: _cabi.Shard, cfg: _cabi.RouteConfig, rank: int = 0, world: int = 1,
                 group=None, reserve_sms: int = 8, exchange: str = "nccl", slice_results: bool = False):
        import copy
        import torch
        self.torch = torch
        self.shard, self.cfg, self.rank, self.world, self.group = shard, cfg, rank, world, group
        self.dev = torch.device("cuda", shard.device)
        mb = shard.max_batch
        nb = mb * _cabi.TOPK * 32
        self.cand2 = [torch.empty(nb, dtype=torch.uint8, device=self.dev) for _ in range(2)]
        self.cand_all2 = ([torch.empty(world * nb, dtype=torch.uint8, device=self.dev) for _ in range(2)]
                          if world > 1 else self.cand2)
        self.cand, self.cand_all = self.cand2[0], self.cand_all2[0]
        self.ovf = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.out = None
        # pipelined mode
        self.out2 = [None, None]
        self.cfg2 = []
        for b in range(2):
            c = copy.copy(cfg)
            c.pipeline_slot, c.reserve_sms = b, reserve_sms
            self.cfg2.append(c)
        self.side = None
        self._prepared = [None, None]
        self._ptrs2 = [None, None]
        self._done_obj = [None, None]
        self.done = [None, None]
        self.iter = 0
        # exchange = "p2p": candidates are stored straight into every peer's buffer by a publish
        # kernel (CUDA IPC over NVLink) instead of an NCCL all-gather; handles travel once, here.
        self.exchange = exchange if world > 1 else "nccl"
        if self.exchange == "p2p":
            import torch.distributed as dist
            ok, mine = 1, b""
            try:
                mine = shard.xchg_init(world, rank)
            except _cabi.QreError:
                ok = 0
            handles = [None] * world
            dist.all_gather_object(handles, (ok, mine), group=group)
            if all(o for o, _ in handles):
                try:
                    for peer, (_, hnd) in enumerate(handles):
                        if peer != rank:
                            shard.xchg_connect(peer, hnd)
                except _cabi.QreError:
                    ok = 0
            else:
                ok = 0
            flag = torch.tensor([ok], device=self.dev, dtype=torch.int32)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)     # every rank takes the same path
            if int(flag.item()) == 0:
                self.exchange = "nccl"
        # sliced results (pipelined peer-exchange path only): rank r merges + routes queries
        # slice_range(m) and leaves the rest of the output arrays untouched
        self.slice_results = False
        self.enable_slices(slice_results)

    def enable_slices(self, on: bool) -> bool:
        """Switch sliced results on / off for the following `step_pipelined` calls (every rank must
        switch at the same batch).  Returns whether slicing is now active."""
        self.slice_results = bool(on) and self.exchange == "p2p" and self.world > 1
        for c in self.cfg2:
            c.slice_results = 1 if self.slice_results else 0
        return self.slice_results

    def slice_range(self, m: int):
        """Queries [q0, q1) whose results this rank produces in `step_pipelined` (all of them unless
        sliced results are enabled on a peer-exchange world)."""
        if not self.slice_results:
            return 0, m
        per = (m + self.world - 1) // self.world
        return min(m, self.rank * per), min(m, (self.rank + 1) * per)

    def _new_out(self, m: int):
        """Result tensors as views of ONE device buffer (`out["_blob"]`), so a caller that wants
        everything on the host needs a single device-to-host copy."""
        torch = self.torch
        K, D = max(1, self.shard.n_centroids), self.cfg.n_devices
        specs, off = [], 0
        for f, (dt, tail) in RESULT_SPECS.items():
            shape = (m,) + tuple({"K": K, "D": D}[t] for t in tail)
            dtype = getattr(torch, dt)
            nbytes = int(np.prod(shape)) * torch.empty((), dtype=dtype).element_size()
            specs.append((f, dtype, shape, off, nbytes))
            off = (off + nbytes + 255) // 256 * 256
        blob = torch.empty(off, dtype=torch.uint8, device=self.dev)
        out = {f: blob[o:o + nb].view(dtype).view(shape) for f, dtype, shape, o, nb in specs}
        out["_blob"] = blob
        self.layout = specs          # (field, dtype, shape, offset, nbytes) within the blob
        return out

    def _alloc_out(self, m: int):
        self.out = self._new_out(m)

    def step_pipelined(self, q, q_ctx, now_us: int, tok=None, heur_dev=None, heur_conf=None):
        """Like `step`, but only the scan of this batch occupies the handle's main stream; the rest
        is queued beside the neighbouring batches' scans.  Inputs must be valid on the current torch
        stream.  Returns (outputs, done): the outputs are valid once `done.synchronize()` returns
        or after `done.wait(stream)` on that stream.  Alternate calls use alternate slots, so a
        batch's outputs are overwritten by the call after the next one."""
        torch = self.torch
        m = q.shape[0]
        b = self.iter & 1
        self.iter += 1
        if self.out2[b] is None or self.out2[b]["hit_slot"].shape[0] != m:
            self.out2[b] = self._new_out(m)
        out = self.out2[b]
        ptrs = self._ptrs2[b]
        if ptrs is None or ptrs[0] is not out:
            ptrs = (out, {f: t.data_ptr() for f, t in out.items() if not f.startswith("_")})
            self._ptrs2[b] = ptrs
        ptrs = ptrs[1]
        cfg = self.cfg2[b]
        main = torch.cuda.current_stream(self.dev)
        if self.exchange == "p2p" or self.world == 1:
            # the marshalled call is cached per (slot, buffers, settings): only `now_us` and the stream change
            key = (m, q.data_ptr(), q_ctx.data_ptr(), 0 if tok is None else tok.data_ptr(),
                   0 if heur_dev is None else heur_dev.data_ptr(), 0 if heur_conf is None else heur_conf.data_ptr(),
                   out["_blob"].data_ptr(), cfg.slice_results, cfg.reserve_sms)
            ent = self._prepared[b]
            if ent is None or ent[0] != key:
                ent = (key, self.shard.submit_dev_prepared(b, m, key[1], key[2], cfg, ptrs, tok_ptr=key[3],
                                                           heur_dev_ptr=key[4], heur_conf_ptr=key[5]))
                self._prepared[b] = ent
                self._done_obj[b] = _SlotDone(self.shard, b)
            ent[1](now_us, main.cuda_stream or 1)
            self.done[b] = self._done_obj[b]
            return out, self.done[b]
        # NCCL exchange: scan chain on the current stream, all-gather + merge on a side stream
        if self.side is None:
            self.side = torch.cuda.Stream(device=self.dev)
        if self.done[b] is not None:
            self.done[b].wait(main)                  # slot b (batch i-2) fully consumed
        nbytes = m * _cabi.TOPK * 32
        self.shard.scan_dev(m, q.data_ptr(), q_ctx.data_ptr(), cfg, now_us, self.rank,
                            self.cand2[b].data_ptr(), 0, main.cuda_stream or 1)
        scanned = torch.cuda.Event()
        scanned.record(main)
        self.side.wait_event(scanned)
        with torch.cuda.stream(self.side):
            import torch.distributed as dist
            dist.all_gather_into_tensor(self.cand_all2[b][: self.world * nbytes], self.cand2[b][:nbytes],
                                        group=self.group)
            self.shard.merge_route_dev(m, self.world, self.cand_all2[b].data_ptr(), cfg, now_us, ptrs,
                                       tok_ptr=0 if tok is None else tok.data_ptr(),
                                       heur_dev_ptr=0 if heur_dev is None else heur_dev.data_ptr(),
                                       heur_conf_ptr=0 if heur_conf is None else heur_conf.data_ptr(),
                                       stream=self.side.cuda_stream)
            ev = torch.cuda.Event()
            ev.record(self.side)
        self.done[b] = _EventDone(ev)
        return out, self.done[b]

    @staticmethod
    def split_sequences(m: int, world: int):
        """Sequence ranges [i0, i1) per rank for data-parallel embedding: equal COUNTS (the all-gather
        needs equal slices), the last ranks padded with empty sequences when world does not divide m."""
        per = (m + world - 1) // world
        return [(min(m, r * per), min(m, (r + 1) * per)) for r in range(world)], per

    def embed_ids(self, enc, ids: np.ndarray, offsets: np.ndarray, normalize: bool = True):
        """Data-parallel embedding in front of the row-sharded lookup (SURVEY §8f-1 + §8e): the batch's
        token ids are replicated on the hosts (like the queries of `step` were); rank r embeds the
        sequences of its slice with the on-GPU encoder `enc` (a `_cabi.Encoder`), and one all-gather of
        the fp32 rows (m x hidden x 4 bytes, 6 MB at 4096 x 384) gives every rank the full [m, hidden]
        query matrix `step` takes.  Embeddings do not depend on batch composition, so the result equals
        embedding everything on one GPU — bit for bit when both take the same FFN-down kernel variant
        (it switches on the token count of the call), to ~1e-3 per component otherwise.
        Returns the device tensor (no sync)."""
        torch = self.torch
        m = offsets.shape[0] - 1
        ranges, per = self.split_sequences(m, self.world)
        i0, i1 = ranges[self.rank]
        H = enc.cfg.hidden
        if getattr(self, "_q_all", None) is None or self._q_all.shape[0] != per * self.world:
            self._q_all = torch.zeros((per * self.world, H), dtype=torch.float32, device=self.dev)
            self._ids_pin = torch.empty(enc.max_tokens, dtype=torch.int32).pin_memory()
            self._off_pin = torch.empty(per + 1, dtype=torch.int32).pin_memory()
            self._ids_dev = torch.empty(enc.max_tokens, dtype=torch.int32, device=self.dev)
            self._off_dev = torch.empty(per + 1, dtype=torch.int32, device=self.dev)
        n = i1 - i0
        T = int(offsets[i1] - offsets[i0])
        if n > enc.max_seqs or T > enc.max_tokens:
            raise _cabi.QreError(f"rank slice of {n} sequences / {T} tokens exceeds the encoder's capacity")
        st = torch.cuda.current_stream(self.dev)
        mine = self._q_all[self.rank * per:(self.rank + 1) * per]
        if n > 0:
            if getattr(self, "_h2d_done", None) is not None:
                self._h2d_done.synchronize()        # the previous batch's ids have left the pinned buffers
            self._ids_pin[:T].copy_(torch.from_numpy(np.ascontiguousarray(ids[offsets[i0]:offsets[i1]], np.int32)))
            self._off_pin[:n + 1].copy_(torch.from_numpy((offsets[i0:i1 + 1] - offsets[i0]).astype(np.int32)))
            self._ids_dev[:T].copy_(self._ids_pin[:T], non_blocking=True)
            self._off_dev[:n + 1].copy_(self._off_pin[:n + 1], non_blocking=True)
            self._h2d_done = torch.cuda.Event()
            self._h2d_done.record(st)
            max_len = int(np.diff(offsets[i0:i1 + 1]).max())
            enc.encode_dev(n, self._ids_dev.data_ptr(), self._off_dev.data_ptr(), T, max_len, mine.data_ptr(),
                           normalize=normalize, stream=st.cuda_stream or 1)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_gather_into_tensor(self._q_all.view(-1), mine.reshape(-1), group=self.group)
        return self._q_all[:m]

    def drain(self):
        """Make the current stream wait for every pipelined batch issued so far."""
        main = self.torch.cuda.current_stream(self.dev)
        for d in self.done:
            if d is not None:
                d.wait(main)

    def step(self, q, q_ctx, now_us: int, tok=None, heur_dev=None, heur_conf=None, check_overflow=False):
        """q: float32 [m, dim] cuda tensor; q_ctx: int64 view of the u64 ids.  Enqueues on the
        current torch stream and returns the dict of result tensors (no synchronisation)."""
        torch = self.torch
        m = q.shape[0]
        if self.out is None or self.out["hit_slot"].shape[0] != m:
            self._alloc_out(m)
        # torch's default stream has handle 0, which the C ABI reads as "the handle's own
        # stream": name the legacy default stream explicitly (cudaStreamLegacy == 0x1).
        st = torch.cuda.current_stream(self.dev).cuda_stream or 1
        nbytes = m * _cabi.TOPK * 32
        if self.exchange == "p2p":
            self.shard.scan_publish_dev(m, q.data_ptr(), q_ctx.data_ptr(), self.cfg, now_us, self.ovf.data_ptr(), st)
            self.shard.merge_route_xchg_dev(
                m, self.cfg, now_us, {f: t.data_ptr() for f, t in self.out.items() if not f.startswith("_")},
                tok_ptr=0 if tok is None else tok.data_ptr(),
                heur_dev_ptr=0 if heur_dev is None else heur_dev.data_ptr(),
                heur_conf_ptr=0 if heur_conf is None else heur_conf.data_ptr(), stream=st)
            return self.out
        self.shard.scan_dev(m, q.data_ptr(), q_ctx.data_ptr(), self.cfg, now_us, self.rank,
                            self.cand.data_ptr(), self.ovf.data_ptr(), st)
        if check_overflow and int(self.ovf.item()) > 0:
            self.shard.resolve_overflow_dev(m, self.cfg, now_us, self.rank, self.cand.data_ptr(), st)
        if self.world > 1:
            import torch.distributed as dist
            dist.all_gather_into_tensor(self.cand_all[: self.world * nbytes], self.cand[:nbytes],
                                        group=self.group)
        ptrs = {f: t.data_ptr() for f, t in self.out.items() if not f.startswith("_")}
        self.shard.merge_route_dev(
            m, self.world, self.cand_all.data_ptr(), self.cfg, now_us, ptrs,
            tok_ptr=0 if tok is None else tok.data_ptr(),
            heur_dev_ptr=0 if heur_dev is None else heur_dev.data_ptr(),
            heur_conf_ptr=0 if heur_conf is None else heur_conf.data_ptr(), stream=st)
        return self.out
