"""Training path of the drop-in TransMIL: forward with every stage kept + the backward of SURVEY 8 row A9.

The reference trains through torch.autograd over code/models/TransMIL.py:167-211 and nystrom-attention 0.0.11
(`training_step`, code/models/model_interface.py:333-383).  Here the forward is the fp32 per-kernel entry points of
libtransmil_b200 (include/transmil_b200.h) with the intermediates the backward needs kept, and the backward is the chain of
contractions written out in oracle/transmil_backward.py, executed by the `tm_*_bwd` / `tm_bgemm` kernels of
csrc/backward.cu.  `TransMILFunction` ties both into autograd so `loss.backward()` / optimizers / DDP see ordinary `.grad`s.

The contractions of the backward run on the tensor cores (csrc/bgemm_tc.cu: tcgen05, tf32 operands, fp32 accumulate; every
transposition form) and the forward of a training step is the tcgen05 engine of the inference path with its stages kept; the
fp32 CUDA-core engine of both stays selectable as the on-device cross-check (tm_bgemm_engine, FORWARD_ENGINE).
There is no CPU or PyTorch fallback: every arithmetic step is a library call.  (`Kernels(lib=...)` lets the CPU tests inject
the host build of the SAME kernel source, tests/cuda_emu -- test infrastructure, never loaded here.)
"""
import ctypes as C

import torch

from . import capi

BWD_SIGNATURES = capi.BWD_SIGNATURES   # C ABI of csrc/backward.cu (declared in include/transmil_b200.h)

H, DH, DIM, LM = 8, 64, 512, 256
TO_OUT_DROPOUT = 0.7    # TransLayer passes dropout=0.7 to NystromAttention (TransMIL.py:33)
STEM768_DROPOUT = 0.6   # the two Dropout layers inside the 768-d _fc1 (TransMIL.py:123-126), active in train() only


def bind(lib):
    for name, (res, args) in BWD_SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


class Kernels:
    """Thin typed front of the backward entry points on one device / stream."""

    def __init__(self, lib=None, stream=0):
        self.L = lib if lib is not None else capi.lib()
        self.st = stream

    def check(self, status):
        if status != 0:
            msg = self.L.tm_last_error().decode("utf-8", "replace")
            raise (ValueError if status in (-1, -2, -6) else capi.TransmilError)(f"libtransmil_b200 (backward): {msg}")

    @staticmethod
    def new(ref, *shape):
        return torch.empty(shape, dtype=torch.float32, device=ref.device)

    def gemm(self, A, a_str, B, b_str, C_, M, N, K, ldc, batch=1, strides=(0, 0, 0), alpha=1.0, beta=0.0, L=None, ldl=None,
             strideL=None):
        """C[z][m,n] = alpha * sum_k A[z][m*sam + k*sak] * B[z][k*sbk + n*sbn] + beta * L[z][m*ldl + n]."""
        nsc = int(self.L.tm_bgemm_scratch_floats(M, N, K, batch))    # split-K partials of the tensor-core engine (0 = none)
        sc = self.new(A, nsc) if nsc else None
        self.check(self.L.tm_bgemm_ex(A.data_ptr(), B.data_ptr(), C_.data_ptr(), L.data_ptr() if L is not None else None, M, N, K,
                                      a_str[0], a_str[1], b_str[0], b_str[1], ldc, ldc if ldl is None else ldl, batch, strides[0],
                                      strides[1], strides[2], strides[2] if strideL is None else strideL, alpha, beta,
                                      sc.data_ptr() if nsc else None, nsc, self.st))
        return C_

    def colsum(self, X, rows, cols, ld, out=None):
        out = self.new(X, cols) if out is None else out
        sc = self.new(X, int(self.L.tm_colsum_scratch_floats(rows, cols)))
        self.check(self.L.tm_colsum(X.data_ptr(), rows, cols, ld, out.data_ptr(), sc.data_ptr(), self.st))
        return out

    def ln_bwd(self, x, gamma, dy, rows, Cc, rows_per_bag, dy_bag_rows, dy_row_off, add):
        dx = torch.empty_like(x)
        gg, gb = self.new(x, Cc), self.new(x, Cc)
        sc = self.new(x, int(self.L.tm_ln_bwd_scratch_floats(rows, Cc)))
        self.check(self.L.tm_ln_rows_bwd(x.data_ptr(), gamma.data_ptr(), dy.data_ptr(), rows, Cc, rows_per_bag, dy_bag_rows,
                                         dy_row_off, add.data_ptr() if add is not None else None, dx.data_ptr(), gg.data_ptr(),
                                         gb.data_ptr(), sc.data_ptr(), self.st))
        return dx, gg, gb


# ---- backward ---------------------------------------------------------------------------------------------------------------
def layer_backward(K, p, name, h_in, sv, dy, plan, seed=0, drop=0.0, keep_dxn=False):
    """Backward of TransLayer.forward (TransMIL.py:45-57): h_in + attn(norm(h_in)).  sv: the layer's kept stages.
    Returns dh_in and the gradients of the layer's six parameter tensors (oracle/transmil_backward.py nystrom_backward)."""
    B, S, _ = h_in.shape
    P, Sp, l = plan["P"], plan["Sp"], plan["l"]
    nb = B * H
    new = lambda *s: K.new(h_in, *s)
    q, k, v, q_lm, k_lm, A2, Z, W3, M, merged, xn = (sv[n] for n in ("q", "k", "v", "q_lm", "k_lm", "attn2", "Z", "W3", "M", "merged", "xn"))
    Wo, Wc, Wqkv = p[name + ".attn.to_out.0.weight"], p[name + ".attn.res_conv.weight"], p[name + ".attn.to_qkv.weight"]
    g = {}
    R = B * Sp
    # to_out (+ Dropout in training) and the `[:, -S:]` slice
    dyp = new(B, Sp, DIM)
    K.check(K.L.tm_pad_front_bwd(dy.data_ptr(), B, S, P, DIM, seed, drop, dyp.data_ptr(), K.st))
    g[name + ".attn.to_out.0.weight"] = K.gemm(dyp, (1, DIM), merged, (DIM, 1), new(DIM, DIM), DIM, DIM, R, DIM)
    g[name + ".attn.to_out.0.bias"] = K.colsum(dyp, R, DIM, DIM)
    dOm = K.gemm(dyp, (DIM, 1), Wo, (DIM, 1), new(B, Sp, DIM), R, DIM, DIM, DIM)
    dO = new(B, H, Sp, DH)
    K.check(K.L.tm_split_heads(dOm.data_ptr(), B, Sp, dO.data_ptr(), K.st))
    del dOm, dyp
    # res_conv
    dv, g_Wc = new(B, H, Sp, DH), new(H, 1, 33, 1)
    sc = new(int(K.L.tm_conv33_bwd_scratch_floats(B, Sp)))
    K.check(K.L.tm_conv33_bwd(dO.data_ptr(), v.data_ptr(), Wc.data_ptr(), B, Sp, dv.data_ptr(), g_Wc.data_ptr(), sc.data_ptr(), K.st))
    del sc
    g[name + ".attn.res_conv.weight"] = g_Wc
    # attn1 = softmax(q k_lm^T) [nb,Sp,256], attn3 = softmax(q_lm k^T) [nb,256,Sp]: recomputed, not kept by the forward
    sQ, sL, s1, s3, s2 = Sp * DH, LM * DH, Sp * LM, LM * Sp, LM * LM
    A1 = K.gemm(q, (DH, 1), k_lm, (1, DH), new(nb, Sp, LM), Sp, LM, DH, LM, nb, (sQ, sL, s1))
    K.check(K.L.tm_softmax_rows(A1.data_ptr(), nb * Sp, LM, K.st))
    A3 = K.gemm(q_lm, (DH, 1), k, (1, DH), new(nb, LM, Sp), LM, Sp, DH, Sp, nb, (sL, sQ, s3))
    K.check(K.L.tm_softmax_rows(A3.data_ptr(), nb * LM, Sp, K.st))
    # out = attn1 (Z W3)
    dA1 = K.gemm(dO, (DH, 1), M, (1, DH), new(nb, Sp, LM), Sp, LM, DH, LM, nb, (sQ, sL, s1))
    dM = K.gemm(A1, (1, LM), dO, (DH, 1), new(nb, LM, DH), LM, DH, Sp, DH, nb, (s1, sQ, sL))
    dZ = K.gemm(dM, (DH, 1), W3, (1, DH), new(nb, LM, LM), LM, LM, DH, LM, nb, (sL, sL, s2))
    dW3 = K.gemm(Z, (1, LM), dM, (DH, 1), new(nb, LM, DH), LM, DH, LM, DH, nb, (s2, sL, sL))
    dA3 = K.gemm(dW3, (DH, 1), v, (1, DH), new(nb, LM, Sp), LM, Sp, DH, Sp, nb, (sL, sQ, s3))
    K.gemm(A3, (1, Sp), dW3, (DH, 1), dv, Sp, DH, LM, DH, nb, (s3, sL, sQ), beta=1.0, L=dv)
    # pinv
    dA2 = new(nb, LM, LM)
    sc = new(int(K.L.tm_pinv_bwd_scratch_floats(nb)))
    K.check(K.L.tm_pinv_bwd(A2.data_ptr(), dZ.data_ptr(), nb, dA2.data_ptr(), sc.data_ptr(), K.st))
    del sc, dZ
    # the three softmaxes (in place: dA* become dS*)
    K.check(K.L.tm_softmax_rows_bwd(A1.data_ptr(), dA1.data_ptr(), nb * Sp, LM, K.st))
    K.check(K.L.tm_softmax_rows_bwd(A2.data_ptr(), dA2.data_ptr(), nb * LM, LM, K.st))
    K.check(K.L.tm_softmax_rows_bwd(A3.data_ptr(), dA3.data_ptr(), nb * LM, Sp, K.st))
    dS1, dS2, dS3 = dA1, dA2, dA3
    dq = K.gemm(dS1, (LM, 1), k_lm, (DH, 1), new(nb, Sp, DH), Sp, DH, LM, DH, nb, (s1, sL, sQ))
    dk_lm = K.gemm(dS1, (1, LM), q, (DH, 1), new(nb, LM, DH), LM, DH, Sp, DH, nb, (s1, sQ, sL))
    K.gemm(dS2, (1, LM), q_lm, (DH, 1), dk_lm, LM, DH, LM, DH, nb, (s2, sL, sL), beta=1.0, L=dk_lm)
    dq_lm = K.gemm(dS2, (LM, 1), k_lm, (DH, 1), new(nb, LM, DH), LM, DH, LM, DH, nb, (s2, sL, sL))
    K.gemm(dS3, (Sp, 1), k, (DH, 1), dq_lm, LM, DH, Sp, DH, nb, (s3, sQ, sL), beta=1.0, L=dq_lm)
    dk = K.gemm(dS3, (1, Sp), q_lm, (DH, 1), new(nb, Sp, DH), Sp, DH, LM, DH, nb, (s3, sL, sQ))
    del A1, A3, dS1, dS3, dA1, dA3
    # landmark means, q scale, head merge; to_qkv
    dqkv = new(B, Sp, 3 * DIM)
    K.check(K.L.tm_qkv_grad_merge(dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), dq_lm.data_ptr(), dk_lm.data_ptr(), B, Sp, l,
                                  dqkv.data_ptr(), K.st))
    g[name + ".attn.to_qkv.weight"] = K.gemm(dqkv, (1, 3 * DIM), xn, (DIM, 1), new(3 * DIM, DIM), 3 * DIM, DIM, R, DIM)
    dxn = K.gemm(dqkv, (3 * DIM, 1), Wqkv, (DIM, 1), new(B, Sp, DIM), R, DIM, 3 * DIM, DIM)
    # LayerNorm + residual
    dh, gg, gb = K.ln_bwd(h_in, p[name + ".norm.weight"], dxn, B * S, DIM, S, Sp, P, dy)
    g[name + ".norm.weight"], g[name + ".norm.bias"] = gg, gb
    if keep_dxn:    # gradient of the LayerNorm output (front pad rows included): what a GradCAM hook on `<layer>.norm` asks for
        g[name + "._dxn"] = dxn
    return dh, g


def _linear_fwd(K, x, W, b, R, Din, Dout):
    return K.gemm(x, (Din, 1), W, (1, Din), K.new(x, R, Dout), R, Dout, Din, Dout, beta=1.0, L=b, ldl=0)


def _linear_wgrad(K, dz, a, g, key, R, Din, Dout):
    g[key + ".weight"] = K.gemm(dz, (1, Dout), a, (Din, 1), K.new(dz, Dout, Din), Dout, Din, R, Din)
    g[key + ".bias"] = K.colsum(dz, R, Dout, Dout)


def _drop(K, t, seed, p):
    if p > 0.0:
        K.check(K.L.tm_dropout_fwd(t.data_ptr(), t.data_ptr(), t.numel(), seed, p, K.st))
    return t


def stem768(K, p, xf, R, D, seeds=(0, 0), drop=0.0):
    """The 768-d stem up to its last LayerNorm (TransMIL.py:123-126): Linear GELU Dropout LN(768) Linear(768,512) GELU Dropout.
    Returns (z0, a0, a1, z4, a4) with a0 / a4 AFTER their Dropout (`drop` = 0.6 in train(), 0 in eval; the keep masks are
    counter-based functions of (seed, element index), so the backward regenerates them).  Used by the training-mode forward
    and, recomputing, by fc1_backward."""
    F0 = p["_fc1.0.weight"].shape[0]
    z0 = _linear_fwd(K, xf, p["_fc1.0.weight"], p["_fc1.0.bias"], R, D, F0)
    a0 = torch.empty_like(z0)
    K.check(K.L.tm_gelu_fwd(z0.data_ptr(), a0.data_ptr(), R * F0, K.st))
    _drop(K, a0, seeds[0], drop)
    a1 = torch.empty_like(z0)
    K.check(K.L.tm_ln_rows_fwd(a0.data_ptr(), p["_fc1.3.weight"].data_ptr(), p["_fc1.3.bias"].data_ptr(), R, F0, a1.data_ptr(), K.st))
    z4 = _linear_fwd(K, a1, p["_fc1.4.weight"], p["_fc1.4.bias"], R, F0, DIM)
    a4 = torch.empty_like(z4)
    K.check(K.L.tm_gelu_fwd(z4.data_ptr(), a4.data_ptr(), R * DIM, K.st))
    _drop(K, a4, seeds[1], drop)
    return z0, a0, a1, z4, a4


def stem768_train_forward(K, p, x, dup, seeds):
    """h0 [B,1+n+dup,512] of a train()-mode forward with the 768-d stem: tm_fc1_fwd is the eval stem (no Dropout inside), so the
    training forward strings the stem together from the same kernels the backward recomputes it with."""
    B, n, D = x.shape
    R = B * n
    a4 = stem768(K, p, x.reshape(R, D), R, D, seeds, STEM768_DROPOUT)[4]
    hf = torch.empty_like(a4)
    K.check(K.L.tm_ln_rows_fwd(a4.data_ptr(), p["_fc1.7.weight"].data_ptr(), p["_fc1.7.bias"].data_ptr(), R, DIM, hf.data_ptr(), K.st))
    h0 = K.new(x, B, 1 + n + dup, DIM)
    K.check(K.L.tm_dup_cls_fwd(hf.data_ptr(), p["cls_token"].data_ptr(), B, n, dup, h0.data_ptr(), K.st))
    return h0


def fc1_backward(K, p, kind, x, dhf, stem_seeds=(0, 0), stem_drop=0.0, want_dx=False):
    """Backward of TransMIL._fc1 (TransMIL.py:100-133) for the parameters and, when `want_dx` (the bag requires grad: the
    reference's autograd would propagate through _fc1 too), for the bag: g["_dx"] = dz0 . W0, one more contraction.  The
    stem's pre-activations are recomputed.  oracle/transmil_backward.py fc1_backward."""
    B, n, D = x.shape
    R = B * n
    g = {}
    xf = x.reshape(R, D)
    dhf = dhf.reshape(R, DIM)

    def dx_from(dz0, F0):
        if want_dx:
            g["_dx"] = K.gemm(dz0, (F0, 1), p["_fc1.0.weight"], (D, 1), K.new(dz0, B, n, D), R, D, F0, D)
    if kind == capi.TM_FC1_SINGLE:
        z0 = _linear_fwd(K, xf, p["_fc1.0.weight"], p["_fc1.0.bias"], R, D, DIM)
        K.check(K.L.tm_gelu_bwd(z0.data_ptr(), dhf.data_ptr(), z0.data_ptr(), R * DIM, K.st))
        _linear_wgrad(K, z0, xf, g, "_fc1.0", R, D, DIM)
        dx_from(z0, DIM)
        return g
    if kind == capi.TM_FC1_2048:
        F0 = p["_fc1.0.weight"].shape[0]
        z0 = _linear_fwd(K, xf, p["_fc1.0.weight"], p["_fc1.0.bias"], R, D, F0)
        a0 = torch.empty_like(z0)
        K.check(K.L.tm_gelu_fwd(z0.data_ptr(), a0.data_ptr(), R * F0, K.st))
        a1 = torch.empty_like(z0)
        K.check(K.L.tm_ln_rows_fwd(a0.data_ptr(), p["_fc1.2.weight"].data_ptr(), p["_fc1.2.bias"].data_ptr(), R, F0, a1.data_ptr(), K.st))
        z3 = _linear_fwd(K, a1, p["_fc1.3.weight"], p["_fc1.3.bias"], R, F0, DIM)
        K.check(K.L.tm_gelu_bwd(z3.data_ptr(), dhf.data_ptr(), z3.data_ptr(), R * DIM, K.st))          # z3 <- dz3
        _linear_wgrad(K, z3, a1, g, "_fc1.3", R, F0, DIM)
        da1 = K.gemm(z3, (DIM, 1), p["_fc1.3.weight"], (F0, 1), a1, R, F0, DIM, F0)                      # a1 <- da1
        da0, g["_fc1.2.weight"], g["_fc1.2.bias"] = K.ln_bwd(a0, p["_fc1.2.weight"], da1, R, F0, R, R, 0, None)
        K.check(K.L.tm_gelu_bwd(z0.data_ptr(), da0.data_ptr(), z0.data_ptr(), R * F0, K.st))            # z0 <- dz0
        _linear_wgrad(K, z0, xf, g, "_fc1.0", R, D, F0)
        dx_from(z0, F0)
        return g
    if kind == capi.TM_FC1_768:
        # Linear GELU [Dropout] LN(768) Linear(768,512) GELU [Dropout] LN(512) (TransMIL.py:123-126); stem_drop = 0.6 in train()
        # (the masks of the forward are regenerated from stem_seeds), 0 for the eval-mode graph (keep_graph)
        F0 = p["_fc1.0.weight"].shape[0]
        z0, a0, a1, z4, a4 = stem768(K, p, xf, R, D, stem_seeds, stem_drop)
        da4, g["_fc1.7.weight"], g["_fc1.7.bias"] = K.ln_bwd(a4, p["_fc1.7.weight"], dhf, R, DIM, R, R, 0, None)
        _drop(K, da4, stem_seeds[1], stem_drop)
        K.check(K.L.tm_gelu_bwd(z4.data_ptr(), da4.data_ptr(), z4.data_ptr(), R * DIM, K.st))           # z4 <- dz4
        _linear_wgrad(K, z4, a1, g, "_fc1.4", R, F0, DIM)
        da1 = K.gemm(z4, (DIM, 1), p["_fc1.4.weight"], (F0, 1), a1, R, F0, DIM, F0)                      # a1 <- da1
        da0, g["_fc1.3.weight"], g["_fc1.3.bias"] = K.ln_bwd(a0, p["_fc1.3.weight"], da1, R, F0, R, R, 0, None)
        _drop(K, da0, stem_seeds[0], stem_drop)
        K.check(K.L.tm_gelu_bwd(z0.data_ptr(), da0.data_ptr(), z0.data_ptr(), R * F0, K.st))            # z0 <- dz0
        _linear_wgrad(K, z0, xf, g, "_fc1.0", R, D, F0)
        dx_from(z0, F0)
        return g
    raise ValueError(f"unknown _fc1 kind {kind}")


def backward(K, p, w, kind, x, saved, dlogits, seeds=(0, 0, 0, 0), drop=0.0, keep_dxn=False, want_dx=False):
    """Gradients of sum(logits * dlogits) for every parameter (reference state_dict keys).  p: name -> parameter tensor,
    w: capi.Weights of the same tensors, saved: stages kept by the forward.  Stage order = reverse of TransMIL.forward
    (TransMIL.py:167-211); mirrors oracle/transmil_backward.py backward()."""
    B, n, _ = x.shape
    plan = saved["plan"]
    S, G, dup = plan["S"], plan["G"], plan["dup"]
    nc = dlogits.shape[1]
    h0, h1, h2, h3 = saved["h0"], saved["h1"], saved["h2"], saved["h3"]
    new = lambda *s: K.new(x, *s)
    g = {}
    dlogits = dlogits.contiguous().float()
    # head
    dcls = new(B, DIM)
    g["norm.weight"], g["norm.bias"], g["_fc.weight"], g["_fc.bias"] = new(DIM), new(DIM), new(nc, DIM), new(nc)
    K.check(K.L.tm_head_bwd(C.byref(w), h3.data_ptr(), dlogits.data_ptr(), B, S, dcls.data_ptr(), g["norm.weight"].data_ptr(),
                            g["norm.bias"].data_ptr(), g["_fc.weight"].data_ptr(), g["_fc.bias"].data_ptr(), K.st))
    dh3 = torch.zeros((B, S, DIM), dtype=torch.float32, device=x.device)
    dh3[:, 0] = dcls
    dh2, gl = layer_backward(K, p, "layer2", h2, saved["layer2"], dh3, plan, seeds[1], drop, keep_dxn)
    g.update(gl)
    del dh3
    # PPEG
    dh1 = new(B, S, DIM)
    gp = {k: torch.empty_like(p["pos_layer." + k]) for k in ("proj.weight", "proj.bias", "proj1.weight", "proj1.bias", "proj2.weight", "proj2.bias")}
    sc = new(int(K.L.tm_ppeg_bwd_scratch_floats(B, G)))
    K.check(K.L.tm_ppeg_bwd(C.byref(w), h1.data_ptr(), dh2.data_ptr(), B, G, dh1.data_ptr(), gp["proj.weight"].data_ptr(),
                            gp["proj.bias"].data_ptr(), gp["proj1.weight"].data_ptr(), gp["proj1.bias"].data_ptr(),
                            gp["proj2.weight"].data_ptr(), gp["proj2.bias"].data_ptr(), sc.data_ptr(), K.st))
    g.update({"pos_layer." + k: v for k, v in gp.items()})
    del dh2, sc
    dh0, gl = layer_backward(K, p, "layer1", h0, saved["layer1"], dh1, plan, seeds[0], drop, keep_dxn)
    g.update(gl)
    del dh1
    # cls token + duplicate pad
    g["cls_token"] = K.colsum(dh0, B, DIM, S * DIM).reshape(1, 1, DIM)
    dhf = new(B, n, DIM)
    K.check(K.L.tm_dup_cls_bwd(dh0.data_ptr(), B, n, dup, dhf.data_ptr(), K.st))
    # seeds = (layer1 to_out, layer2 to_out, stem Dropout 0, stem Dropout 1); the stem's Dropout exists in the 768-d form only
    if kind == capi.TM_FC1_TOKENS:   # the stem ran outside the library (CTMIL's conv stem): hand d(loss)/d(tokens) back to autograd
        g["_dx"] = dhf
        return g
    stem_drop = STEM768_DROPOUT if (drop > 0.0 and kind == capi.TM_FC1_768) else 0.0
    g.update(fc1_backward(K, p, kind, x, dhf, tuple(seeds[2:4]) if len(seeds) >= 4 else (0, 0), stem_drop, want_dx))
    return g


# ---- forward with kept stages ---------------------------------------------------------------------------------------------------
# "tc": the tcgen05 engine of the inference path (tf32 fc1, fp16 q/k/v/xn/merged, fp16 pinv chain) with every stage kept;
# "simt": the fp32 CUDA-core engine (the on-device cross-check).  tests/test_zz_backward_gpu.py runs both.
FORWARD_ENGINE = "tc"


def forward_keep(model, x, training, seeds):
    if FORWARD_ENGINE == "tc" and capi.lib().tm_device_supported() == 1 and not (int(model.flags) & capi.TM_FLAG_SIMT_ONLY):
        return forward_keep_tc(model, x, training, seeds)
    return forward_keep_simt(model, x, training, seeds)


def forward_keep_tc(model, x, training, seeds):
    """TransMIL.forward (TransMIL.py:167-211) through the per-kernel TENSOR-CORE entry points (the kernels tm_transmil_fwd
    strings together for inference), every stage the backward reads kept in its own tensor.  The fp16 stages (q, k, v, xn,
    merged) are handed to the backward as exact fp32 copies (tm_cvt_f32): the gradient is taken at the values the forward
    actually used.  Training mode applies the to_out Dropout(0.7) with the library's counter-based mask (seed per layer)."""
    L = capi.lib()
    dev = x.device
    st = torch.cuda.current_stream(dev).cuda_stream
    ck = capi.check
    w = model._weights()
    B, n, D = x.shape
    plan = capi.make_plan(n).as_dict()
    S, Sp, P, l, G = plan["S"], plan["Sp"], plan["P"], plan["l"], plan["G"]
    f32 = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
    f16 = lambda *s: torch.empty(s, dtype=torch.float16, device=dev)
    flags = capi.TM_FLAG_TF32_OPERANDS | capi.TM_FLAG_F16_ACT
    keep = []   # operand copies of the GEMM weights must outlive the launches that read them

    def tf32_copy(t):
        o = torch.empty_like(t)
        ck(L.tm_round_tf32(t.data_ptr(), o.data_ptr(), t.numel(), st))
        keep.append(o)
        return o.data_ptr()

    def f16_copy(t):
        o = torch.empty(t.shape, dtype=torch.float16, device=dev)
        ck(L.tm_cvt_f16(t.data_ptr(), o.data_ptr(), t.numel(), st))
        keep.append(o)
        return o.data_ptr()

    def up(t):
        o = torch.empty(t.shape, dtype=torch.float32, device=dev)
        ck(L.tm_cvt_f32(t.data_ptr(), o.data_ptr(), t.numel(), st))
        return o

    wt = capi.Weights.from_buffer_copy(w)
    F0 = {capi.TM_FC1_SINGLE: 0, capi.TM_FC1_2048: 1024, capi.TM_FC1_768: 768, capi.TM_FC1_TOKENS: 0}[model._fc1_kind]
    if model._fc1_kind != capi.TM_FC1_TOKENS:
        f = model._fc1
        wt.fc1_w0 = tf32_copy(f[0].weight)
        if model._fc1_kind == capi.TM_FC1_2048:
            wt.fc1_w1 = tf32_copy(f[3].weight)
        elif model._fc1_kind == capi.TM_FC1_768:
            wt.fc1_w1 = tf32_copy(f[4].weight)
    if training and model._fc1_kind == capi.TM_FC1_768:   # Dropout(0.6) inside the stem (TransMIL.py:123-126)
        h0 = stem768_train_forward(Kernels(stream=st), {k: t.detach() for k, t in model.named_parameters()}, x, plan["dup"], seeds[2:4])
    else:
        h0 = f32(B, S, DIM)
        scratch = f32(max(1, B * n * F0))
        ck(L.tm_fc1_fwd(C.byref(wt), x.data_ptr(), B, n, h0.data_ptr(), scratch.data_ptr(), flags, st))
        del scratch
    saved = {"plan": plan, "h0": h0}

    def layer(name, lw, mod, h_in, seed):
        qkv16, out16 = f16_copy(mod.attn.to_qkv.weight), f16_copy(mod.attn.to_out[0].weight)
        xn = f16(B, Sp, DIM)
        ck(L.tm_ln_pad_fwd(h_in.data_ptr(), lw.norm_w, lw.norm_b, B, S, P, xn.data_ptr(), flags, st))
        q, k, v = f16(B, H, Sp, DH), f16(B, H, Sp, DH), f16(B, H, Sp, DH)
        ck(L.tm_qkv16_fwd(xn.data_ptr(), qkv16, B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), flags, st))
        q_lm, k_lm, qlm16, klm16 = f32(B, H, LM, DH), f32(B, H, LM, DH), f16(B, H, LM, DH), f16(B, H, LM, DH)
        ck(L.tm_landmarks16_fwd(q.data_ptr(), k.data_ptr(), B, Sp, l, q_lm.data_ptr(), k_lm.data_ptr(), qlm16.data_ptr(),
                                klm16.data_ptr(), st))
        attn2, Z = f32(B, H, LM, LM), f32(B, H, LM, LM)
        sc = f32(int(L.tm_pinv_scratch_floats(B)))
        ck(L.tm_pinv_fwd(q_lm.data_ptr(), k_lm.data_ptr(), B, attn2.data_ptr(), Z.data_ptr(), sc.data_ptr(), flags, st))
        W3 = f32(B, H, LM, DH)
        sc = f32(int(L.tm_attn3v_tc_scratch_floats(B, Sp)))
        ck(L.tm_attn3v_tc_fwd(qlm16.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, W3.data_ptr(), sc.data_ptr(), flags, st))
        M, mt16 = f32(B, H, LM, DH), f16(B, H, DH, LM)
        ck(L.tm_zw16_fwd(Z.data_ptr(), W3.data_ptr(), B, M.data_ptr(), mt16.data_ptr(), st))
        merged = f16(B, Sp, DIM)
        ck(L.tm_nystrom_out_tc_fwd(q.data_ptr(), klm16.data_ptr(), mt16.data_ptr(), v.data_ptr(), lw.conv_w, B, Sp,
                                   merged.data_ptr(), flags, st))
        h_out = f32(B, S, DIM)
        ck(L.tm_to_out_residual_fwd(merged.data_ptr(), out16, lw.out_b, h_in.data_ptr(), B, S, P, h_out.data_ptr(), flags, st))
        if training:
            ck(L.tm_dropout_resid_fwd(h_out.data_ptr(), h_in.data_ptr(), B * S * DIM, seed, TO_OUT_DROPOUT, st))
        saved[name] = dict(xn=up(xn), q=up(q), k=up(k), v=up(v), q_lm=q_lm, k_lm=k_lm, attn2=attn2, Z=Z, W3=W3, M=M, merged=up(merged))
        return h_out

    h1 = layer("layer1", w.layer1, model.layer1, h0, seeds[0])
    h2 = f32(B, S, DIM)
    packed = f32(DIM * 50)
    ck(L.tm_ppeg_fwd(C.byref(w), h1.data_ptr(), B, G, h2.data_ptr(), packed.data_ptr(), st))
    h3 = layer("layer2", w.layer2, model.layer2, h2, seeds[1])
    logits = f32(B, model.n_classes)
    ck(L.tm_head_fwd(C.byref(w), h3.data_ptr(), B, S, logits.data_ptr(), st))
    saved.update(h1=h1, h2=h2, h3=h3)
    del keep
    return logits, saved, w


def forward_keep_simt(model, x, training, seeds):
    """TransMIL.forward (TransMIL.py:167-211) through the per-kernel fp32 entry points, keeping what the backward reads.
    Training mode applies the to_out Dropout(0.7) with the library's counter-based mask (seed per layer)."""
    L = capi.lib()
    dev = x.device
    st = torch.cuda.current_stream(dev).cuda_stream
    ck = capi.check
    w = model._weights()
    B, n, D = x.shape
    plan = capi.make_plan(n).as_dict()
    S, Sp, P, l, G = plan["S"], plan["Sp"], plan["P"], plan["l"], plan["G"]
    new = lambda *s: torch.empty(s, dtype=torch.float32, device=dev)
    flags = capi.TM_FLAG_SIMT_ONLY
    F0 = {capi.TM_FC1_SINGLE: 0, capi.TM_FC1_2048: 1024, capi.TM_FC1_768: 768, capi.TM_FC1_TOKENS: 0}[model._fc1_kind]
    if training and model._fc1_kind == capi.TM_FC1_768:   # Dropout(0.6) inside the stem (TransMIL.py:123-126)
        h0 = stem768_train_forward(Kernels(stream=st), {k: t.detach() for k, t in model.named_parameters()}, x, plan["dup"], seeds[2:4])
    else:
        h0 = new(B, S, DIM)
        scratch = new(max(1, B * n * F0))
        ck(L.tm_fc1_fwd(C.byref(w), x.data_ptr(), B, n, h0.data_ptr(), scratch.data_ptr(), flags, st))
        del scratch
    saved = {"plan": plan, "h0": h0}

    def layer(name, lw, h_in, seed):
        sv = {}
        xn = new(B, Sp, DIM)
        ck(L.tm_ln_pad_fwd(h_in.data_ptr(), lw.norm_w, lw.norm_b, B, S, P, xn.data_ptr(), flags, st))
        q, k, v = new(B, H, Sp, DH), new(B, H, Sp, DH), new(B, H, Sp, DH)
        ck(L.tm_qkv_fwd(xn.data_ptr(), lw.qkv_w, B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), flags, st))
        q_lm, k_lm = new(B, H, LM, DH), new(B, H, LM, DH)
        ck(L.tm_landmarks_fwd(q.data_ptr(), k.data_ptr(), B, Sp, l, q_lm.data_ptr(), k_lm.data_ptr(), st))
        attn2, Z = new(B, H, LM, LM), new(B, H, LM, LM)
        sc = new(int(L.tm_pinv_scratch_floats(B)))
        ck(L.tm_pinv_fwd(q_lm.data_ptr(), k_lm.data_ptr(), B, attn2.data_ptr(), Z.data_ptr(), sc.data_ptr(), flags, st))
        W3 = new(B, H, LM, DH)
        sc = new(int(L.tm_attn3v_scratch_floats(B, Sp)))
        ck(L.tm_attn3v_fwd(q_lm.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, W3.data_ptr(), sc.data_ptr(), flags, st))
        M = new(B, H, LM, DH)
        ck(L.tm_zw_fwd(Z.data_ptr(), W3.data_ptr(), B, M.data_ptr(), st))
        merged = new(B, Sp, DIM)
        ck(L.tm_nystrom_out_fwd(q.data_ptr(), k_lm.data_ptr(), M.data_ptr(), v.data_ptr(), lw.conv_w, B, Sp, merged.data_ptr(), flags, st))
        h_out = new(B, S, DIM)
        ck(L.tm_to_out_residual_fwd(merged.data_ptr(), lw.out_w, lw.out_b, h_in.data_ptr(), B, S, P, h_out.data_ptr(), flags, st))
        if training:
            ck(L.tm_dropout_resid_fwd(h_out.data_ptr(), h_in.data_ptr(), B * S * DIM, seed, TO_OUT_DROPOUT, st))
        sv.update(xn=xn, q=q, k=k, v=v, q_lm=q_lm, k_lm=k_lm, attn2=attn2, Z=Z, W3=W3, M=M, merged=merged)
        saved[name] = sv
        return h_out

    h1 = layer("layer1", w.layer1, h0, seeds[0])
    h2 = new(B, S, DIM)
    packed = new(DIM * 50)
    ck(L.tm_ppeg_fwd(C.byref(w), h1.data_ptr(), B, G, h2.data_ptr(), packed.data_ptr(), st))
    h3 = layer("layer2", w.layer2, h2, seeds[1])
    logits = new(B, model.n_classes)
    ck(L.tm_head_fwd(C.byref(w), h3.data_ptr(), B, S, logits.data_ptr(), st))
    saved.update(h1=h1, h2=h2, h3=h3)
    return logits, saved, w


class TransMILFunction(torch.autograd.Function):
    """logits = TransMIL(x) with the library's forward and backward; parameters enter as autograd inputs so that
    `.grad`, optimizers and DistributedDataParallel (gradient allreduce over NCCL) work unchanged."""

    @staticmethod
    def forward(ctx, model, x, *params):
        bind(capi.lib())
        training = model.training
        seeds = tuple(int(s) for s in torch.randint(0, 2 ** 62, (4,))) if training else (0, 0, 0, 0)   # to_out x 2, 768-d stem x 2
        with _kernels_for(x)[0]:
            logits, saved, w = forward_keep(model, x, training, seeds)
        ctx.model, ctx.x, ctx.saved, ctx.seeds, ctx.training = model, x.detach(), saved, seeds, training
        ctx.names = [n for n, _ in model._path_named_parameters()]
        return logits

    @staticmethod
    def backward(ctx, dlogits):
        p, g = _param_grads(ctx, dlogits, want_dx=ctx.needs_input_grad[1])
        dx = g["_dx"].reshape(ctx.x.shape) if ("_dx" in g and ctx.needs_input_grad[1]) else None   # a bag / token tensor that requires grad
        return (None, dx) + tuple(g[n].reshape(p[n].shape) for n in ctx.names)


def _kernels_for(x):
    """(device context, Kernels) of the tensor's device and current stream; the host harness of the tests replaces this."""
    return torch.cuda.device(x.device), Kernels(stream=torch.cuda.current_stream(x.device).cuda_stream)


def _param_grads(ctx, dlogits, keep_dxn=False, want_dx=False):
    """The library backward for what a forward kept on `ctx`: (parameters by name, gradients by name)."""
    model, x = ctx.model, ctx.x
    if ctx.saved is None:
        raise RuntimeError("libtransmil_b200: the stages kept by this forward were released by an earlier backward (the library "
                           "backward runs once per forward; call the model again instead of backward(retain_graph=True))")
    p = {n: t.detach() for n, t in model._path_named_parameters()}
    ppeg = getattr(model, "PPEG_NAME", "pos_layer")
    # the backward speaks TransMIL's names; sibling models call the PPEG block pos_layer_0 (CTMIL) and the classifier _fc2 (MDMIL)
    canon = {(("pos_layer" + n[len(ppeg):]) if n.startswith(ppeg + ".") else n): t for n, t in p.items()}
    dev, K = _kernels_for(x)
    with dev:
        g = backward(K, canon, model._weights(), model._fc1_kind, x, ctx.saved, dlogits, ctx.seeds,
                     TO_OUT_DROPOUT if ctx.training else 0.0, keep_dxn, want_dx)
    ctx.saved = None
    if model.HEAD != "_fc":
        g[model.HEAD + ".weight"], g[model.HEAD + ".bias"] = g.pop("_fc.weight"), g.pop("_fc.bias")
    if ppeg != "pos_layer":
        for k in [k for k in g if k.startswith("pos_layer.")]:
            g[ppeg + k[len("pos_layer"):]] = g.pop(k)
    return p, g


# ---- GradCAM-style hooks on `<layer>.norm` (consumers: code/train_loop.py:87, code/test_visualize.py:121-127) -----
class _Holder:
    """Links the two autograd nodes of a hooked forward."""

    def __init__(self):
        self.ctx, self.inputs, self.grads, self.params = None, {}, None, None


class _HookedFunction(torch.autograd.Function):
    """TransMILFunction that also returns the LayerNorm outputs of the hooked layers (rows of `xn` behind the front pad,
    [B,S,512]) as autograd outputs, so that a consumer can keep them and register gradient hooks on them."""

    @staticmethod
    def forward(ctx, model, x, holder, hooked, *params):
        bind(capi.lib())
        training = model.training
        seeds = tuple(int(s) for s in torch.randint(0, 2 ** 62, (4,))) if training else (0, 0, 0, 0)   # to_out x 2, 768-d stem x 2
        with _kernels_for(x)[0]:
            logits, saved, _w = forward_keep(model, x, training, seeds)
        ctx.model, ctx.x, ctx.saved, ctx.seeds, ctx.training = model, x, saved, seeds, training
        ctx.names = [n for n, _ in model._path_named_parameters()]
        ctx.hooked, ctx.P, ctx.holder = hooked, int(saved["plan"]["P"]), holder
        holder.ctx = ctx
        holder.inputs = {"layer1": saved["h0"], "layer2": saved["h2"], "norm": saved["h3"]}
        K = _kernels_for(x)[1]
        with _kernels_for(x)[0]:
            acts = tuple(_final_norm(K, model, saved["h3"]) if name == "norm" else saved[name]["xn"][:, ctx.P:].clone() for name in hooked)
        return (logits,) + acts

    @staticmethod
    def backward(ctx, dlogits, *dacts):
        holder = ctx.holder
        if holder.grads is not None:                    # _AttachGrad ran the library backward already
            p, g = holder.params, holder.grads
            holder.grads = holder.params = None
        else:
            p, g = _param_grads(ctx, dlogits)
        return (None, None, None, None) + tuple(g[n].reshape(p[n].shape) for n in ctx.names)


class _AttachGrad(torch.autograd.Function):
    """Identity on logits that makes the hooked activations part of the graph: its backward runs the library backward once,
    hands d(loss)/d(activation) to autograd (tensor hooks on the activations fire with the true gradient) and leaves the
    parameter gradients for _HookedFunction.backward, which runs right after it."""

    @staticmethod
    def forward(ctx, logits, holder, *acts):
        ctx.holder = holder
        return logits.clone()

    @staticmethod
    def backward(ctx, dlogits):
        holder = ctx.holder
        main = holder.ctx
        holder.params, holder.grads = _param_grads(main, dlogits.contiguous(), keep_dxn=True)
        dacts = tuple(_final_norm_grad(main, dlogits.contiguous()) if name == "norm" else holder.grads[name + "._dxn"][:, main.P:].contiguous()
                      for name in main.hooked)
        return (dlogits, None) + dacts


def _layer_norm_rows(K, h, weight, bias):
    """nn.LayerNorm(512) of every row of h [..., 512] (fp32, tm_ln_rows_fwd)."""
    out = torch.empty_like(h)
    K.check(K.L.tm_ln_rows_fwd(h.data_ptr(), weight.data_ptr(), bias.data_ptr(), h.numel() // DIM, DIM, out.data_ptr(), K.st))
    return out


def _final_norm(K, model, h3):
    """Output of `self.norm(h)` (TransMIL.py:202) on every row: what a forward hook on the final LayerNorm sees in the reference
    (the head kernel itself normalises the cls row only)."""
    return _layer_norm_rows(K, h3.contiguous(), model.norm.weight.detach(), model.norm.bias.detach())


def _final_norm_grad(main, dlogits):
    """d(loss)/d(norm output) [B,S,512]: only the cls row reaches the classifier (TransMIL.py:202-204), row 0 = dlogits . W_fc."""
    model, h3 = main.model, main.holder.inputs["norm"]
    dev, K = _kernels_for(h3)
    W = getattr(model, model.HEAD).weight.detach()
    nc = W.shape[0]
    B = h3.shape[0]
    d = torch.zeros_like(h3)
    prev = K.L.tm_bgemm_engine(1)   # a [B,nc] x [nc,512] product: the fp32 CUDA-core kernel, not tensor-core work
    try:
        with dev:
            row0 = K.gemm(dlogits.float().contiguous(), (nc, 1), W, (DIM, 1), K.new(h3, B, DIM), B, DIM, nc, DIM)
    finally:
        K.L.tm_bgemm_engine(prev)
    d[:, 0] = row0
    return d


def fire_hooks_no_graph(model, hooked, inputs):
    """Forward hooks on `layer{1,2}.norm` / `norm` on a call without a graph (no_grad, train() under no_grad, or the eval-mode
    `return_attn=True` call of code/visualize_mil.py:580): each hook is called as `hook(norm_module, (its input [B,S,512],),
    LayerNorm(input) [B,S,512])`, the way torch calls it in the reference (TransMIL.py:47,202).  The activation is computed in fp32
    from the kept layer input; the input tensor is a view of the forward's workspace, valid until the next forward."""
    for name in hooked:
        norm = model._norm_module(name)
        h_in = inputs[name]
        dev, K = _kernels_for(h_in)
        with dev:
            act = _layer_norm_rows(K, h_in.contiguous(), norm.weight.detach(), norm.bias.detach())
        for hook in list(norm._forward_hooks.values()):
            if hook(norm, (h_in,), act) is not None:
                raise NotImplementedError("libtransmil_b200: a forward hook on %s returned a replacement output" % name)


def forward_with_norm_hooks(model, x, hooked):
    """Forward of the autograd path with the forward hooks of `<layer>.norm` honoured for the layers in `hooked`: each hook
    is called as `hook(norm_module, (layer input [B,S,512],), LayerNorm output [B,S,512])`, the way torch calls it in the
    reference, where TransLayer runs `self.norm(x)` (TransMIL.py:47).  Hooks that return a replacement output are refused."""
    holder = _Holder()
    outs = _HookedFunction.apply(model, x, holder, tuple(hooked), *model._path_parameters())
    logits, acts = outs[0], outs[1:]
    for name, act in zip(hooked, acts):
        norm = model._norm_module(name)
        for hook in list(norm._forward_hooks.values()):
            if hook(norm, (holder.inputs[name],), act) is not None:
                raise NotImplementedError("libtransmil_b200: a forward hook on %s returned a replacement output" % name)
    return _AttachGrad.apply(logits, holder, *acts)
