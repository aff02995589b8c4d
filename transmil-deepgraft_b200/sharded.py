"""Token-sharded forward of ONE giant bag across the ranks of a process group (BASELINE config 4).

The reference runs a bag on one GPU (code/models/TransMIL.py:167-211); Nystrom attention makes the bag splittable along
the token axis with a handful of small exchanges per layer, so a 200k-token slide can use 2/4/8 B200s for latency:

  * every rank owns a contiguous range of the PADDED token axis [row0, row1), a multiple of 256 rows, and runs the ordinary
    per-kernel entry points of libtransmil_b200 on it (B = 1, Sp = rows; rank 0 also owns the P front-pad rows and cls);
  * landmarks: per-group partial sums -> allreduce(sum) of [2,8,256,64] fp32 (1 MB) -> means (groups may straddle ranks);
  * sim2 / pseudo-inverse / Z.W3: replicated (identical inputs on every rank -> identical bits, no exchange);
  * attn3.v: per-rank (max, sum, numerator) partials -> allgather (540 KB per rank) -> log-sum-exp merge in rank order;
  * res_conv (33 taps): 16-row halo of v to both neighbours (16 KB each way), added by a boundary-fix kernel;
  * PPEG between the layers: a 3-image-row halo of the layer-1 output to both neighbours (<= 4G rows each way);
  * fc1 is row-parallel; the duplicate-pad tail (TransMIL.py:177-180) is recomputed from the first `dup` input rows by the
    rank(s) that own it, so it needs no communication;  * the cls row lives on rank 0: logits there, broadcast at the end.

`ShardPlan` is pure integer arithmetic (testable anywhere; `tm_make_shard_plan` holds the same integers in C).

Two ways to run the schedule (`TokenShardedTransMIL(transport=...)`):
  * "peer" (default): ONE C call per rank and bag, `tm_transmil_fwd_sharded` (csrc/sharded_fwd.cu): the library issues every
    kernel and every exchange itself -- flag-guarded pulls over NVLink peer memory between communication regions that the
    ranks map through CUDA IPC (`PeerComm`); nothing is launched from Python, so the call is captured into a CUDA graph at any
    world size.  torch.distributed is used once, to pass the 64-byte IPC handles around.
  * "nccl": `sharded_forward` below issues the schedule from Python kernel by kernel with torch.distributed collectives in
    between (NCCL over NVLink on the GPUs).  All arithmetic is done by an `ops` object; the product one is `CudaOps`
    (ctypes -> libtransmil_b200.so, no fallback).  The `ops` seam exists so that the world_size-2 gloo test on CPU can drive
    the very same host logic with a torch restatement that lives under tests/ (never imported from here).  On the device
    it is the cross-check of the peer path.
"""
import ctypes as C
import math

import numpy as np
import torch
import torch.distributed as dist

from . import capi

LM, HEADS, DHEAD, DIM = 256, 8, 64, 512


def bag_plan(n):
    """Index contract of one bag, same integers as tm_make_plan (TransMIL.py:177-179,184-186; nystrom-attention pad)."""
    G = int(np.ceil(np.sqrt(n)))
    N = G * G
    S = N + 1
    rem = S % LM
    P = LM - rem if rem else 0
    return dict(n=n, G=G, N=N, dup=N - n, S=S, l=math.ceil(S / LM), Sp=S + P, P=P, padding=P)


class RankShard:
    """What one rank owns.  rows: padded rows [row0,row1); seq positions [s0,s1) (0 = cls); tokens [tok0,tok1)."""

    def __init__(self, plan, row0, row1, rank, world):
        p = plan
        self.rank, self.world = rank, world
        self.row0, self.row1, self.rows = row0, row1, row1 - row0
        self.P_local = p["P"] if rank == 0 else 0
        self.s0, self.s1 = max(row0 - p["P"], 0), row1 - p["P"]
        self.S_local = self.s1 - self.s0
        self.has_cls = rank == 0
        self.tok0, self.tok1 = max(self.s0 - 1, 0), self.s1 - 1
        self.ntok = self.tok1 - self.tok0
        G, n = p["G"], p["n"]
        # rows of x this rank projects: its tokens t < n are x[t]; tokens t >= n are the duplicate pad x[t - n]
        self.x_ranges = []
        if self.tok0 < n:
            self.x_ranges.append((self.tok0, min(self.tok1, n)))
        if self.tok1 > n:
            self.x_ranges.append((max(self.tok0, n) - n, self.tok1 - n))
        # PPEG: whole image rows [y_in0, y_in1) = own rows + 3-row halo
        self.y_in0 = max(0, self.tok0 // G - 3)
        self.y_in1 = min(G, (self.tok1 - 1) // G + 4)
        self.halo_lo = self.tok0 - self.y_in0 * G        # tokens needed from lower ranks
        self.halo_hi = self.y_in1 * G - self.tok1        # tokens needed from higher ranks


class ShardPlan:
    def __init__(self, n, world):
        self.p = p = bag_plan(n)
        self.world = world
        l = p["l"]
        if l < world:
            raise ValueError(f"a bag of {n} patches has only {l} blocks of 256 padded rows: cannot shard over {world} ranks")
        cuts = [256 * ((l * r) // world) for r in range(world + 1)]
        self.shards = [RankShard(p, cuts[r], cuts[r + 1], r, world) for r in range(world)]
        for r, s in enumerate(self.shards):
            lo_ok = s.halo_lo == 0 or (r > 0 and s.halo_lo <= self.shards[r - 1].ntok)
            hi_ok = s.halo_hi == 0 or (r + 1 < world and s.halo_hi <= self.shards[r + 1].ntok)
            if not (lo_ok and hi_ok and s.rows >= 32):
                raise ValueError(f"a bag of {n} patches is too small to shard over {world} ranks "
                                 f"(the PPEG halo of rank {r} would span more than one neighbour)")

    def take(self, x, rank):
        """The rows of the full bag x [n,D] (or [1,n,D]) that `rank` has to hold: concatenation of its x_ranges."""
        if x.dim() == 3:
            x = x[0]
        return torch.cat([x[a:b] for a, b in self.shards[rank].x_ranges], dim=0)

    def take_from_file(self, path, rank, pin=None, dtype=None):
        """`take` for a bag that lives in one of the reference's feature files (code/datasets/feature_dataloader.py:252-254):
        only this rank's rows are read (h5bag.read_bag_rows), straight into pinned host memory."""
        from . import h5bag
        return h5bag.read_bag_rows(path, self.shards[rank].x_ranges, pin=pin, dtype=dtype)


# ---------------------------------------------------------------------------------------------------------------------
class CudaOps:
    """Arithmetic of the sharded path on the current CUDA device: every method is one (or a few) C-ABI calls."""

    def __init__(self, model):
        self.m = model
        self.L = capi.lib()
        self.dev = model.cls_token.device
        if self.dev.type != "cuda":
            raise capi.TransmilError("token-sharded TransMIL needs the model on a CUDA device; there is no CPU path")
        self.flags = capi.TM_FLAG_TF32_OPERANDS | capi.TM_FLAG_F16_ACT
        self._st = lambda: torch.cuda.current_stream(self.dev).cuda_stream
        self._keep = []
        self._side = None
        self._arena_key, self._arena, self._arena_pos = None, [], 0
        self._weights_key = None
        self.refresh_weights()
        self.packed = self.f32(50 * DIM)
        self.pinv_scratch = self.f32(int(self.L.tm_pinv_scratch_floats(1)))

    def refresh_weights(self):
        """Operand copies of the GEMM weights (tf32-rounded fc1, fp16 to_qkv / to_out: what tm_transmil_fwd derives into its
        workspace on every call), re-derived whenever a parameter changed since they were made -- load_state_dict, an optimizer
        step or a replaced tensor (keyed on data_ptr and the tensors' version counters), so the sharded path never mixes old
        operand copies with new norm / bias / conv tensors."""
        model = self.m
        key = tuple((p.data_ptr(), p._version) for p in model.parameters())
        if key == self._weights_key:
            return
        self._weights_key = key
        self._keep = []
        self.w = model._weights()
        f = model._fc1
        self.w.fc1_w0 = self._tf32(f[0].weight)
        if model._fc1_kind == capi.TM_FC1_2048:
            self.w.fc1_w1 = self._tf32(f[3].weight)
        elif model._fc1_kind == capi.TM_FC1_768:
            self.w.fc1_w1 = self._tf32(f[4].weight)
        self.layers = {}
        for name in ("layer1", "layer2"):
            layer = getattr(model, name)
            self.layers[name] = dict(
                norm_w=layer.norm.weight, norm_b=layer.norm.bias, qkv16=self._f16(layer.attn.to_qkv.weight),
                out16=self._f16(layer.attn.to_out[0].weight), out_b=layer.attn.to_out[0].bias, conv_w=layer.attn.res_conv.weight)

    # -- small helpers
    # Buffers: a forward asks for the same sequence of buffers every time it runs the same plan, so they are handed out
    # from a replayed list (`begin_forward(key)` rewinds it) instead of ~60 allocator calls per bag.  Reuse is safe: every
    # consumer of a buffer is ordered before its next producer on the same stream (NCCL ops are ordered against it by
    # torch.distributed).
    def begin_forward(self, key):
        self.refresh_weights()
        if key != self._arena_key:
            self._arena_key, self._arena = key, []
        self._arena_pos = 0

    def _buf(self, shape, dtype):
        shape = tuple(int(v) for v in (shape[0] if len(shape) == 1 and isinstance(shape[0], (tuple, list)) else shape))
        if self._arena_key is None:
            return torch.empty(shape, dtype=dtype, device=self.dev)
        i = self._arena_pos
        self._arena_pos += 1
        if i < len(self._arena):
            t = self._arena[i]
            if t.shape == shape and t.dtype == dtype:
                return t
            del self._arena[i:]
        t = torch.empty(shape, dtype=dtype, device=self.dev)
        self._arena.append(t)
        return t

    def f32(self, *shape):
        return self._buf(shape, torch.float32)

    def f16(self, *shape):
        return self._buf(shape, torch.float16)

    def _tf32(self, wt):
        out = torch.empty_like(wt)
        capi.check(self.L.tm_round_tf32(wt.data_ptr(), out.data_ptr(), wt.numel(), self._st()))
        self._keep.append(out)
        return out.data_ptr()

    def _f16(self, wt):
        out = torch.empty(wt.shape, dtype=torch.float16, device=self.dev)
        capi.check(self.L.tm_cvt_f16(wt.data_ptr(), out.data_ptr(), wt.numel(), self._st()))
        return out

    def to_device(self, x):
        return x.to(device=self.dev, dtype=torch.float32, non_blocking=True).contiguous()

    def comm_buffer(self, *shape, dtype=torch.float32):
        return self._buf(shape, dtype)

    def side_stream(self):
        """Context: kernels launched inside run on a high-priority side stream forked from the current stream."""
        if self._side is None:
            self._side = torch.cuda.Stream(self.dev, priority=-1)
        self._side.wait_stream(torch.cuda.current_stream(self.dev))
        return torch.cuda.stream(self._side)

    def join_side_stream(self):
        torch.cuda.current_stream(self.dev).wait_stream(self._side)

    # -- kernels
    def fc1_rows(self, x_rows, out):
        rows = x_rows.shape[0]
        f0 = {capi.TM_FC1_2048: 1024, capi.TM_FC1_768: 768}.get(self.m._fc1_kind, 0)
        scratch = self.f32(rows * f0) if f0 else None
        capi.check(self.L.tm_fc1_rows_fwd(C.byref(self.w), x_rows.data_ptr(), rows, out.data_ptr(),
                                          scratch.data_ptr() if f0 else 0, self.flags, self._st()))

    def cls_row(self, out_row):
        out_row.copy_(self.m.cls_token.detach().view(DIM))

    def ln_pad(self, h, name, S_local, P_local):
        lw = self.layers[name]
        xn = self.f16(S_local + P_local, DIM)
        capi.check(self.L.tm_ln_pad_fwd(h.data_ptr(), lw["norm_w"].data_ptr(), lw["norm_b"].data_ptr(), 1, S_local, P_local,
                                        xn.data_ptr(), self.flags, self._st()))
        return xn

    def qkv(self, xn, name):
        rows = xn.shape[0]
        q, k, v = self.f16(HEADS, rows, DHEAD), self.f16(HEADS, rows, DHEAD), self.f16(HEADS, rows, DHEAD)
        capi.check(self.L.tm_qkv16_fwd(xn.data_ptr(), self.layers[name]["qkv16"].data_ptr(), 1, rows, q.data_ptr(), k.data_ptr(),
                                       v.data_ptr(), self.flags, self._st()))
        return q, k, v

    def v_edges(self, v):
        return v[:, :16].contiguous(), v[:, -16:].contiguous()

    def landmark_sums(self, q, k, row0, l, sums):
        capi.check(self.L.tm_landmark_sums16_fwd(q.data_ptr(), k.data_ptr(), 1, q.shape[1], row0, l, sums.data_ptr(), self._st()))

    def landmark_finish(self, sums, l):
        lm = dict(q_lm=self.f32(HEADS, LM, DHEAD), k_lm=self.f32(HEADS, LM, DHEAD), qlm16=self.f16(HEADS, LM, DHEAD),
                  klm16=self.f16(HEADS, LM, DHEAD))
        capi.check(self.L.tm_landmark_finish_fwd(sums.data_ptr(), 1, l, lm["q_lm"].data_ptr(), lm["k_lm"].data_ptr(),
                                                 lm["qlm16"].data_ptr(), lm["klm16"].data_ptr(), self._st()))
        return lm

    def pinv(self, lm):
        attn2, Z = self.f32(HEADS, LM, LM), self.f32(HEADS, LM, LM)
        capi.check(self.L.tm_pinv_fwd(lm["q_lm"].data_ptr(), lm["k_lm"].data_ptr(), 1, attn2.data_ptr(), Z.data_ptr(),
                                      self.pinv_scratch.data_ptr(), self.flags, self._st()))
        return Z

    def partial_floats(self):
        return int(self.L.tm_attn3v_partial_floats(1))

    def attn3v_partial(self, lm, k, v, partial):
        rows = k.shape[1]
        scratch = self.f32(int(self.L.tm_attn3v_tc_scratch_floats(1, rows)))
        capi.check(self.L.tm_attn3v_tc_partial_fwd(lm["qlm16"].data_ptr(), k.data_ptr(), v.data_ptr(), 1, rows, partial.data_ptr(),
                                                   scratch.data_ptr(), self._st()))

    def attn3v_merge(self, gathered, world):
        W3 = self.f32(HEADS, LM, DHEAD)
        capi.check(self.L.tm_attn3v_merge_fwd(gathered.data_ptr(), world, 1, W3.data_ptr(), self._st()))
        return W3

    def zw(self, Z, W3):
        M, mt16 = self.f32(HEADS, LM, DHEAD), self.f16(HEADS, DHEAD, LM)
        capi.check(self.L.tm_zw16_fwd(Z.data_ptr(), W3.data_ptr(), 1, M.data_ptr(), mt16.data_ptr(), self._st()))
        return mt16

    def nystrom_out(self, q, lm, mt16, v, name):
        rows = q.shape[1]
        merged = self.f16(rows, DIM)
        capi.check(self.L.tm_nystrom_out_tc_fwd(q.data_ptr(), lm["klm16"].data_ptr(), mt16.data_ptr(), v.data_ptr(),
                                                self.layers[name]["conv_w"].data_ptr(), 1, rows, merged.data_ptr(), self.flags,
                                                self._st()))
        return merged

    def conv_halo_fix(self, merged, v_prev, v_next, name):
        capi.check(self.L.tm_conv_halo_fix_fwd(v_prev.data_ptr() if v_prev is not None else 0,
                                               v_next.data_ptr() if v_next is not None else 0,
                                               self.layers[name]["conv_w"].data_ptr(), 1, merged.shape[0], merged.data_ptr(),
                                               self._st()))

    def to_out(self, merged, name, h_in, S_local, P_local, h_out):
        lw = self.layers[name]
        capi.check(self.L.tm_to_out_residual_fwd(merged.data_ptr(), lw["out16"].data_ptr(), lw["out_b"].data_ptr(), h_in.data_ptr(),
                                                 1, S_local, P_local, h_out.data_ptr(), self.flags, self._st()))

    def ppeg_rows(self, in_rows, y_in0, y_in1, G, tok0, tok1, out):
        capi.check(self.L.tm_ppeg_rows_fwd(C.byref(self.w), in_rows.data_ptr(), y_in0, y_in1, G, tok0, tok1, out.data_ptr(),
                                           self.packed.data_ptr(), self._st()))

    def head(self, h, S_local):
        logits = self.f32(1, self.m.n_classes)
        capi.check(self.L.tm_head_fwd(C.byref(self.w), h.data_ptr(), 1, S_local, logits.data_ptr(), self._st()))
        return logits

    def empty_logits(self):
        return self.f32(1, self.m.n_classes)


# ---------------------------------------------------------------------------------------------------------------------
def _exchange(rank, world, send_lo=None, send_hi=None, recv_lo=None, recv_hi=None):
    """Neighbour exchange along the token axis: send_lo -> rank-1, send_hi -> rank+1, recv_lo <- rank-1, recv_hi <- rank+1.
    Returns the outstanding requests (wait on them before touching the recv buffers)."""
    p2p = []
    if rank > 0:
        if send_lo is not None:
            p2p.append(dist.P2POp(dist.isend, send_lo, rank - 1))
        if recv_lo is not None:
            p2p.append(dist.P2POp(dist.irecv, recv_lo, rank - 1))
    if rank + 1 < world:
        if send_hi is not None:
            p2p.append(dist.P2POp(dist.isend, send_hi, rank + 1))
        if recv_hi is not None:
            p2p.append(dist.P2POp(dist.irecv, recv_hi, rank + 1))
    return dist.batch_isend_irecv(p2p) if p2p else []


def _all_gather(gathered, mine):
    if dist.get_backend() == "nccl":
        dist.all_gather_into_tensor(gathered, mine)
    else:
        dist.all_gather(list(gathered.unbind(0)), mine)


def _layer(ops, plan, sh, name, h_in, h_out):
    """One TransLayer (TransMIL.py:45-57) on this rank's rows; 1 allreduce + 1 allgather + 1 neighbour exchange."""
    rank, world, l = sh.rank, sh.world, plan.p["l"]
    xn = ops.ln_pad(h_in, name, sh.S_local, sh.P_local)
    q, k, v = ops.qkv(xn, name)
    first, last = ops.v_edges(v)
    v_prev = ops.comm_buffer(*first.shape, dtype=first.dtype) if rank > 0 else None
    v_next = ops.comm_buffer(*first.shape, dtype=first.dtype) if rank + 1 < world else None
    reqs = _exchange(rank, world, send_lo=first, send_hi=last, recv_lo=v_prev, recv_hi=v_next)
    sums = ops.comm_buffer(2, HEADS, LM, DHEAD)
    ops.landmark_sums(q, k, sh.row0, l, sums)
    dist.all_reduce(sums)
    lm = ops.landmark_finish(sums, l)
    with ops.side_stream():                            # the latency-bound pinv chain runs beside attn3.v (own SMs)
        Z = ops.pinv(lm)                               # replicated: same inputs, same bits on every rank
    pf = ops.partial_floats()
    partial = ops.comm_buffer(pf)
    ops.attn3v_partial(lm, k, v, partial)
    gathered = ops.comm_buffer(world, pf)
    _all_gather(gathered, partial)
    W3 = ops.attn3v_merge(gathered, world)
    ops.join_side_stream()
    M = ops.zw(Z, W3)
    merged = ops.nystrom_out(q, lm, M, v, name)
    for r in reqs:
        r.wait()
    ops.conv_halo_fix(merged, v_prev, v_next, name)
    ops.to_out(merged, name, h_in, sh.S_local, sh.P_local, h_out)


def sharded_forward(ops, plan, x_local):
    """TransMIL.forward (TransMIL.py:167-211) of ONE bag over the default process group.

    x_local: this rank's rows of the bag, `plan.take(x, rank)` ([ntok, D], host or device).  Returns logits [1, n_classes]
    on every rank."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if world != plan.world:
        raise ValueError(f"plan was made for {plan.world} ranks, the process group has {world}")
    sh, p = plan.shards[rank], plan.p
    off = 1 if sh.has_cls else 0
    x_local = ops.to_device(x_local)
    if tuple(x_local.shape[:1]) != (sh.ntok,):
        raise ValueError(f"rank {rank} expects {sh.ntok} rows of the bag (plan.take), got {tuple(x_local.shape)}")
    # fc1 + (implicit) duplicate pad + cls  (TransMIL.py:175-186)
    h0 = ops.f32(sh.S_local, DIM)
    ops.fc1_rows(x_local, h0[off:])
    if sh.has_cls:
        ops.cls_row(h0[0])
    # layer 1 writes into a buffer with room for the PPEG halos on both sides
    lo, hi = sh.halo_lo, sh.halo_hi
    ext = ops.f32(lo + sh.S_local + hi, DIM)
    h1 = ext[lo:lo + sh.S_local]
    _layer(ops, plan, sh, "layer1", h0, h1)
    # PPEG halo: my last tokens go up, my first tokens go down (TransMIL.py:60-75 on the G x G token image)
    n_up = plan.shards[rank + 1].halo_lo if rank + 1 < world else 0
    n_dn = plan.shards[rank - 1].halo_hi if rank > 0 else 0
    toks = h1[off:]
    reqs = _exchange(rank, world,
                     send_lo=toks[:n_dn] if n_dn else None, send_hi=toks[sh.ntok - n_up:] if n_up else None,
                     recv_lo=ext[:lo] if lo else None, recv_hi=ext[lo + sh.S_local:] if hi else None)
    for r in reqs:
        r.wait()
    h2 = ops.f32(sh.S_local, DIM)
    ops.ppeg_rows(ext[off:], sh.y_in0, sh.y_in1, p["G"], sh.tok0, sh.tok1, h2[off:])
    if sh.has_cls:
        h2[0].copy_(h1[0])
    h3 = ops.f32(sh.S_local, DIM)
    _layer(ops, plan, sh, "layer2", h2, h3)
    # final norm + head on the cls row (TransMIL.py:202-204), rank 0; everyone gets the logits
    logits = ops.head(h3, sh.S_local) if sh.has_cls else ops.empty_logits()
    dist.broadcast(logits, 0)
    return logits


class PeerComm:
    """This rank's communication region for tm_transmil_fwd_sharded plus the mappings of the peers' regions.  The 64-byte
    CUDA IPC handles are exchanged once through the process group (any transport would do: they are plain bytes)."""

    def __init__(self, max_n, dev):
        L = self.L = capi.lib()
        self.rank, self.world, self.max_n = dist.get_rank(), dist.get_world_size(), int(max_n)
        if self.world > capi.TM_SHARD_MAX_WORLD:
            raise ValueError(f"token sharding supports up to {capi.TM_SHARD_MAX_WORLD} ranks")
        self.handle = C.c_void_p()
        with torch.cuda.device(dev):
            # every step is followed by an exchange of "did it work here", so a failure on one rank (no peer access, IPC refused
            # in a container) raises on EVERY rank instead of leaving the others waiting in a collective
            err, mine = None, (C.c_ubyte * capi.TM_SHARD_HANDLE_BYTES)()
            try:
                capi.check(L.tm_shard_comm_create(self.rank, self.world, self.max_n, C.byref(self.handle)))
                if self.world > 1:
                    capi.check(L.tm_shard_comm_handle(self.handle, mine))
            except Exception as e:      # noqa: BLE001
                err = f"rank {self.rank}: {e}"
            if self.world > 1:
                blobs = [None] * self.world
                dist.all_gather_object(blobs, (err, bytes(mine)))
                errs = [b[0] for b in blobs if b[0]]
                if not errs:
                    try:
                        capi.check(L.tm_shard_comm_attach(self.handle, b"".join(b[1] for b in blobs)))
                    except Exception as e:      # noqa: BLE001
                        err = f"rank {self.rank}: {e}"
                    errs2 = [None] * self.world
                    dist.all_gather_object(errs2, err)     # also the barrier: nobody starts a forward before every region is mapped
                    errs = [e for e in errs2 if e]
                err = "; ".join(errs) if errs else None
            if err:
                self.close()
                raise capi.TransmilError(f"token-sharded TransMIL: peer-memory transport unavailable ({err}); use transport='nccl'")

    def status(self):
        """0, or 1 + the index of the exchange whose producer timed out since the last call (synchronises the device)."""
        code = C.c_int32()
        capi.check(self.L.tm_shard_comm_status(self.handle, C.byref(code)))
        return int(code.value)

    def close(self):
        """Collective in spirit: call it on every rank after a barrier (a peer may still be reading this rank's region)."""
        if self.handle:
            self.L.tm_shard_comm_destroy(self.handle)
            self.handle = C.c_void_p()


def shard_plan_c(n, world, rank):
    """tm_make_shard_plan as a capi.ShardPlanC (same integers as ShardPlan(n, world).shards[rank])."""
    sp = capi.ShardPlanC()
    capi.check(capi.lib().tm_make_shard_plan(int(n), int(world), int(rank), C.byref(sp)))
    return sp


class TokenShardedTransMIL:
    """`TokenShardedTransMIL(model)(x)`: the drop-in module's forward for ONE bag, token-sharded over the default process
    group (one process per GPU).  `x` is the full bag [n,D] / [1,n,D] on every rank (host or device) or, with `n=...`, this
    rank's rows only (`plan(n).take(x, rank)`) -- the form a sharded loader would hand over.

    transport="peer" (default): ONE library call per bag, tm_transmil_fwd_sharded -- the whole per-rank schedule and the
    exchanges (flag-guarded pulls over NVLink peer memory) are issued by the library on the current stream; with
    cuda_graph=True the call is captured once per bag size and replayed (any world size); a host bag is copied straight into
    the graph's input buffer (one H2D), a device bag is copied into it too unless static_input=True and it is the very
    tensor the graph was captured on (a caller that refills one device buffer per bag: no copy at all).
    transport="nccl": the schedule is issued from Python (`sharded_forward`) kernel by kernel with torch.distributed
    collectives in between -- the cross-check of the peer path, and the form the gloo CPU tests drive."""

    def __init__(self, model, cuda_graph=False, transport="peer", max_n=1 << 20, static_input=False):
        if model.training:
            raise NotImplementedError("token-sharded path: inference only (model.eval())")
        if transport not in ("peer", "nccl"):
            raise ValueError("transport must be 'peer' or 'nccl'")
        self.model = model
        self.transport = transport
        self._plans = {}
        self._graphs = {}      # n -> dict(graph, x, logits) | "eager" (first call done, capture next) | "failed"
        if transport == "peer":
            self.dev = model.cls_token.device
            if self.dev.type != "cuda":
                raise capi.TransmilError("token-sharded TransMIL needs the model on a CUDA device; there is no CPU path")
            self.cuda_graph = bool(cuda_graph)
            self.static_input = bool(static_input)
            self.comm = PeerComm(max_n, self.dev)
            self._cplans, self._ws, self._w = {}, None, None
            return
        self.ops = CudaOps(model)
        # cuda_graph=True (single-rank groups only): after an eager first call per bag size the whole schedule -- kernels,
        # side-stream fork/join -- is captured into a CUDA graph and replayed, one launch per bag.  With more than one rank
        # the capture would have to include the NCCL point-to-point exchanges; that hung on a 2 x B200 box (r1), so
        # multi-rank groups always issue the schedule eagerly.
        self.cuda_graph = bool(cuda_graph) and dist.get_world_size() == 1

    def plan(self, n):
        key = (int(n), dist.get_world_size())
        if key not in self._plans:
            self._plans[key] = ShardPlan(*key)
        return self._plans[key]

    def _run_graph(self, plan, x):
        n = plan.p["n"]
        state = self._graphs.get(n)
        if state is None or state == "failed":
            if state is None:
                self._graphs[n] = "eager"
            return sharded_forward(self.ops, plan, x)
        if state == "eager":
            static_x = self.ops.to_device(x).clone()
            try:
                torch.cuda.synchronize(self.ops.dev)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g):
                    logits = sharded_forward(self.ops, plan, static_x)
                state = self._graphs[n] = dict(graph=g, x=static_x, logits=logits)
            except Exception as e:   # same kernels either way: fall back to issuing them one by one
                import warnings
                warnings.warn(f"token-sharded forward: CUDA graph capture failed ({e}); running eagerly")
                self._graphs[n] = "failed"
                torch.cuda.synchronize(self.ops.dev)
                return sharded_forward(self.ops, plan, x)
        state["x"].copy_(x, non_blocking=True)
        state["graph"].replay()
        return state["logits"].clone()

    # ---- transport="peer" -------------------------------------------------------------------------------------------
    def _peer_call(self, cplan, x_rows, logits):
        L, m = capi.lib(), self.model
        w = self._w = m._weights()
        nbytes = C.c_size_t()
        capi.check(L.tm_sharded_workspace_bytes(C.byref(w), C.byref(cplan), C.byref(nbytes)))
        if self._ws is None or self._ws.numel() < nbytes.value:
            if self._graphs:
                self._graphs = {}          # captured graphs point into the old workspace
            self._ws = torch.empty(int(nbytes.value), dtype=torch.uint8, device=self.dev)
        capi.check(L.tm_transmil_fwd_sharded(C.byref(w), x_rows.data_ptr(), C.byref(cplan), self.comm.handle, logits.data_ptr(),
                                             self._ws.data_ptr(), self._ws.numel(), 0, torch.cuda.current_stream(self.dev).cuda_stream))

    def _peer_forward(self, plan, x):
        n, rank = plan.p["n"], dist.get_rank()
        if n > self.comm.max_n:            # every rank sees the same n, so every rank rebuilds (a collective) together
            torch.cuda.synchronize(self.dev)
            dist.barrier()
            self.comm.close()
            self.comm = PeerComm(2 * n, self.dev)
            self._graphs = {}
        cplan = self._cplans.get(n)
        if cplan is None:
            cplan = self._cplans[n] = shard_plan_c(n, plan.world, rank)
        if tuple(x.shape) != (int(cplan.ntok), self.model.in_features):
            raise ValueError(f"rank {rank} expects {int(cplan.ntok)} rows of the bag (plan.take), got {tuple(x.shape)}")
        state = self._graphs.get(n) if self.cuda_graph else None
        if isinstance(state, dict):
            # replay: the graph reads a fixed device buffer.  A host bag is copied straight into it (one H2D, nothing else); a
            # device bag that IS that buffer (the graph was captured on the caller's tensor) costs nothing
            key = tuple(p.data_ptr() for p in self.model.parameters())
            if state["key"] == key:
                if not (x.is_cuda and x.data_ptr() == state["x"].data_ptr()):
                    state["x"].copy_(x, non_blocking=True)
                state["graph"].replay()
                return state["logits"].clone()
            state = self._graphs[n] = None       # a parameter tensor was replaced: the graph holds the old pointers
        given = x
        x = x.to(device=self.dev, dtype=torch.float32, non_blocking=True).contiguous()
        if not self.cuda_graph:
            logits = torch.empty(1, self.model.n_classes, device=self.dev)
            self._peer_call(cplan, x, logits)
            return logits
        # the captured call reads the parameters through their pointers at replay time (operand copies are re-derived inside
        # the graph), so in-place updates need nothing; a REPLACED parameter tensor needs a new capture
        key = tuple(p.data_ptr() for p in self.model.parameters())
        if state is None:
            # first call of a bag size: eager (loads the kernels, sizes the workspace) -- then capture.  All ranks take the
            # same branch because they see the same sequence of bag sizes.
            logits = torch.empty(1, self.model.n_classes, device=self.dev)
            self._peer_call(cplan, x, logits)
            self._graphs[n] = "eager"
            return logits
        # second call of a bag size: capture, on a private copy of the input -- or, with static_input=True, ON the caller's device
        # tensor (kept alive here): a caller that refills one buffer per bag then replays with no copy at all, and any OTHER
        # input passed later is copied into that tensor
        mine = not (given.is_cuda and given.data_ptr() == x.data_ptr())
        sx = x if (mine or self.static_input) else x.clone()
        sl = torch.empty(1, self.model.n_classes, device=self.dev)
        torch.cuda.synchronize(self.dev)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            self._peer_call(cplan, sx, sl)
        state = self._graphs[n] = dict(graph=g, x=sx, logits=sl, key=key, w=self._w)
        state["graph"].replay()
        return state["logits"].clone()

    @torch.no_grad()
    def __call__(self, x, n=None):
        if n is None:
            if x.dim() == 3:
                if x.shape[0] != 1:
                    raise ValueError("token sharding is for a single bag (B = 1); batch whole bags across ranks instead")
                x = x[0]
            plan = self.plan(x.shape[0])
            x = plan.take(x, dist.get_rank())
        else:
            plan = self.plan(n)
        if self.transport == "peer":
            with torch.cuda.device(self.dev):
                return self._peer_forward(plan, x)
        with torch.cuda.device(self.ops.dev):
            if self.cuda_graph:
                before = self.ops._weights_key
                self.ops.refresh_weights()
                if self.ops._weights_key != before:
                    self._graphs = {}                     # captured graphs hold the old operand copies
                return self._run_graph(plan, x)
            self.ops.begin_forward(plan.p["n"])
            return sharded_forward(self.ops, plan, x).clone()      # the buffer list is reused by the next call
