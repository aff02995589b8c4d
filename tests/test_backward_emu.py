"""Backward kernels (csrc/backward.cu) and their host orchestration (training.py) against the backward oracle, on the CPU.

The build container has no GPU, so the kernel SOURCE is compiled with g++ against tests/cuda_emu/cuda_emu.h (grid / block /
__syncthreads executed by fibers on the host) and driven through the same C ABI and the same Python orchestration the GPU
path uses.  This checks indexing, strides, reductions and the order of contractions; the GPU suite
(tests/test_zz_backward_gpu.py) checks the nvcc build of the same source on the device.
"""
import ctypes as C
import importlib.util
import os
import subprocess

import pytest
import torch
import torch.nn.functional as F

from common import ROOT, build_model, make_input
from oracle import transmil_backward as OB
from oracle import transmil_oracle as O

EMU_DIR = os.path.join(ROOT, "tests", "cuda_emu")
EMU_LIB = os.path.join(EMU_DIR, "_build", "libtransmil_b200_emu.so")
SRC = os.path.join(ROOT, "transmil-deepgraft_b200", "csrc", "backward.cu")


@pytest.fixture(scope="module")
def T():
    import transmil_b200 as T
    return T


@pytest.fixture(scope="module")
def K(T):
    os.makedirs(os.path.dirname(EMU_LIB), exist_ok=True)
    deps = [SRC, os.path.join(EMU_DIR, "cuda_emu.h"), os.path.join(ROOT, "include", "transmil_b200.h")]
    if not os.path.isfile(EMU_LIB) or os.path.getmtime(EMU_LIB) < max(os.path.getmtime(d) for d in deps):
        subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-DTM_HOST_EMU", "-x", "c++", "-I", os.path.join(ROOT, "include"),
                        "-I", EMU_DIR, SRC, "-o", EMU_LIB, "-lpthread"], check=True)
    lib = C.CDLL(EMU_LIB)
    lib.tm_last_error.restype = C.c_char_p
    from transmil_b200 import training
    training.bind(lib)
    return training.Kernels(lib=lib, stream=None)


def _close(a, b, rtol=2e-4, what=""):
    a, b = a.double(), b.double()
    err = (a - b).abs().max().item()
    bound = rtol * max(b.abs().max().item(), 1e-6)
    assert err <= bound, f"{what}: max abs err {err:.3e} > {bound:.3e} (ref max {b.abs().max().item():.3e})"


def test_every_declared_backward_symbol_is_exported(K, T):
    from transmil_b200 import training
    for name in training.BWD_SIGNATURES:
        assert hasattr(K.L, name)
    hdr = open(os.path.join(ROOT, "include", "transmil_b200.h")).read()
    for name in training.BWD_SIGNATURES:
        assert name + "(" in hdr, f"{name} is not declared in include/transmil_b200.h"


@pytest.mark.parametrize("tA,tB", [(False, False), (False, True), (True, False), (True, True)])
def test_bgemm_every_orientation(K, tA, tB):
    g = torch.Generator().manual_seed(3)
    nb, M, N, Kd = 3, 70, 37, 45
    A = torch.randn(nb, *((Kd, M) if tA else (M, Kd)), generator=g)
    B = torch.randn(nb, *((N, Kd) if tB else (Kd, N)), generator=g)
    Lt = torch.randn(nb, M, N, generator=g)
    out = torch.empty(nb, M, N)
    K.gemm(A, (1, M) if tA else (Kd, 1), B, (1, Kd) if tB else (N, 1), out, M, N, Kd, N, nb, (M * Kd, N * Kd, M * N), alpha=0.5, beta=-2.0, L=Lt)
    ref = 0.5 * (A.transpose(1, 2) if tA else A).double() @ (B.transpose(1, 2) if tB else B).double() - 2.0 * Lt.double()
    _close(out, ref, 1e-5, "bgemm")


def test_bias_broadcast_through_ldl_zero(K):
    g = torch.Generator().manual_seed(4)
    x, W, b = torch.randn(50, 24, generator=g), torch.randn(12, 24, generator=g), torch.randn(12, generator=g)
    from transmil_b200 import training
    _close(training._linear_fwd(K, x, W, b, 50, 24, 12), F.linear(x.double(), W.double(), b.double()), 1e-5, "linear")


def test_softmax_rows_and_backward(K):
    g = torch.Generator().manual_seed(5)
    X = torch.randn(7, 300, generator=g) * 3
    P = X.clone()
    K.check(K.L.tm_softmax_rows(P.data_ptr(), 7, 300, None))
    _close(P, X.double().softmax(-1), 1e-5, "softmax")
    dP = torch.randn(7, 300, generator=g)
    ref = OB.softmax_backward(P.double(), dP.double())
    K.check(K.L.tm_softmax_rows_bwd(P.data_ptr(), dP.data_ptr(), 7, 300, None))
    _close(dP, ref, 1e-5, "softmax backward")


def test_layernorm_backward_with_row_map(K):
    g = torch.Generator().manual_seed(6)
    B, S, P, Cc = 2, 37, 5, 512
    x, gamma = torch.randn(B, S, Cc, generator=g), torch.rand(Cc, generator=g) + 0.5
    dyp = torch.randn(B, S + P, Cc, generator=g)
    add = torch.randn(B, S, Cc, generator=g)
    dx, gg, gb = K.ln_bwd(x, gamma, dyp, B * S, Cc, S, S + P, P, add)
    rdx, rgg, rgb = OB.layernorm_backward(x.double(), gamma.double(), dyp[:, P:].double())
    _close(dx, rdx + add.double(), 1e-5, "LN dx")
    _close(gg, rgg, 1e-4, "LN dgamma")
    _close(gb, rgb, 1e-4, "LN dbeta")


def test_layernorm_backward_many_rows_two_stage(K):
    g = torch.Generator().manual_seed(7)
    R, Cc = 700, 96
    x, gamma, dy = torch.randn(R, Cc, generator=g), torch.rand(Cc, generator=g) + 0.5, torch.randn(R, Cc, generator=g)
    dx, gg, gb = K.ln_bwd(x, gamma, dy, R, Cc, R, R, 0, None)
    rdx, rgg, rgb = OB.layernorm_backward(x.double(), gamma.double(), dy.double())
    _close(dx, rdx, 1e-5, "LN dx")
    _close(gg, rgg, 1e-4, "LN dgamma")
    _close(gb, rgb, 1e-4, "LN dbeta")
    _close(K.colsum(dy, R, Cc, Cc), dy.double().sum(0), 1e-4, "colsum")


def test_gelu_forward_backward(K):
    g = torch.Generator().manual_seed(8)
    z, dy = torch.randn(1000, generator=g) * 2, torch.randn(1000, generator=g)
    a, dz = torch.empty_like(z), torch.empty_like(z)
    K.check(K.L.tm_gelu_fwd(z.data_ptr(), a.data_ptr(), 1000, None))
    K.check(K.L.tm_gelu_bwd(z.data_ptr(), dy.data_ptr(), dz.data_ptr(), 1000, None))
    _close(a, F.gelu(z.double()), 1e-5, "gelu")
    _close(dz, OB.gelu_backward(z.double(), dy.double()), 1e-5, "gelu backward")


def test_pinv_backward(K):
    """dA of the 6-iteration pinv with the global init scalar; compared after the softmax backward, which annihilates the
    per-row constants that ties among the (all ~1) row sums distribute differently in fp32 and fp64."""
    g = torch.Generator().manual_seed(9)
    nb = 2
    A = (torch.randn(1, nb, 256, 256, generator=g) * 0.5).softmax(-1)
    dZ = torch.randn(1, nb, 256, 256, generator=g)
    ref = OB.softmax_backward(A.double(), OB.pinv_backward(A.double(), dZ.double()))
    dA = torch.empty(nb, 256, 256)
    sc = torch.empty(int(K.L.tm_pinv_bwd_scratch_floats(nb)))
    dZw = dZ.clone()    # clobbered by the call
    K.check(K.L.tm_pinv_bwd(A.data_ptr(), dZw.data_ptr(), nb, dA.data_ptr(), sc.data_ptr(), None))
    K.check(K.L.tm_softmax_rows_bwd(A.data_ptr(), dA.data_ptr(), nb * 256, 256, None))
    _close(dA.view(1, nb, 256, 256), ref, 2e-3, "pinv backward")


def _weights(T, sd, kind, D, nc):
    w = T.capi.Weights()
    w.fc1_kind, w.in_features, w.n_classes = kind, D, nc
    for f, k in (("ppeg_w7", "pos_layer.proj.weight"), ("ppeg_b7", "pos_layer.proj.bias"), ("ppeg_w5", "pos_layer.proj1.weight"),
                 ("ppeg_b5", "pos_layer.proj1.bias"), ("ppeg_w3", "pos_layer.proj2.weight"), ("ppeg_b3", "pos_layer.proj2.bias"),
                 ("norm_w", "norm.weight"), ("norm_b", "norm.bias"), ("fc_w", "_fc.weight"), ("fc_b", "_fc.bias")):
        setattr(w, f, sd[k].data_ptr())
    return w


def test_ppeg_backward(K, T):
    m = build_model(2, 512)
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    g = torch.Generator().manual_seed(10)
    B, G = 2, 9
    h, dy = torch.randn(B, G * G + 1, 512, generator=g), torch.randn(B, G * G + 1, 512, generator=g)
    rdh, rg = OB.ppeg_backward({k: v.double() for k, v in sd.items()}, h.double(), G, dy.double())
    w = _weights(T, sd, 0, 512, 2)
    dh = torch.empty_like(h)
    gp = {k: torch.empty_like(sd["pos_layer." + k]) for k in ("proj.weight", "proj.bias", "proj1.weight", "proj1.bias", "proj2.weight", "proj2.bias")}
    sc = torch.empty(int(K.L.tm_ppeg_bwd_scratch_floats(B, G)))
    K.check(K.L.tm_ppeg_bwd(C.byref(w), h.data_ptr(), dy.data_ptr(), B, G, dh.data_ptr(), gp["proj.weight"].data_ptr(), gp["proj.bias"].data_ptr(),
                            gp["proj1.weight"].data_ptr(), gp["proj1.bias"].data_ptr(), gp["proj2.weight"].data_ptr(), gp["proj2.bias"].data_ptr(),
                            sc.data_ptr(), None))
    _close(dh, rdh, 1e-5, "ppeg dh")
    for k, v in gp.items():
        _close(v, rg["pos_layer." + k], 1e-4, "ppeg " + k)


def test_conv33_backward(K):
    g = torch.Generator().manual_seed(11)
    B, Sp = 2, 256
    dO, v, w = torch.randn(B, 8, Sp, 64, generator=g), torch.randn(B, 8, Sp, 64, generator=g), torch.randn(8, 1, 33, 1, generator=g)
    vv = v.double().requires_grad_(True)
    ww = w.double().requires_grad_(True)
    F.conv2d(vv, ww, None, padding=(16, 0), groups=8).backward(dO.double())
    dv, gw = torch.empty_like(v), torch.empty_like(w)
    K.check(K.L.tm_conv33_bwd(dO.data_ptr(), v.data_ptr(), w.data_ptr(), B, Sp, dv.data_ptr(), gw.data_ptr(), None, None))
    _close(dv, vv.grad, 1e-5, "conv dv")
    _close(gw, ww.grad, 1e-4, "conv dw")


def _saved_from_oracle(sd, x):
    """What training.forward_keep keeps, taken from the forward oracle (fp32)."""
    logits, st = O.forward(sd, x, return_stages=True)
    p = O.plan(x.shape[1])
    saved = {"plan": p, "h0": st["h0"], "h1": st["h1"], "h2": st["h2"], "h3": st["h3"]}
    for name in ("layer1", "layer2"):
        s = st[name]
        sv = {k: s[k].contiguous() for k in ("q", "k", "v", "q_lm", "k_lm", "attn2", "Z", "W3", "merged")}
        sv["xn"] = F.pad(s["xn"], (0, 0, p["P"], 0)).contiguous()
        sv["M"] = (s["Z"] @ s["W3"]).contiguous()
        saved[name] = sv
    return logits, {k: (v.contiguous() if torch.is_tensor(v) else v) for k, v in saved.items()}


@pytest.mark.parametrize("D,nc,n,B", [(96, 3, 300, 2), (2048, 2, 130, 1)])
def test_whole_backward_against_oracle(K, T, D, nc, n, B):
    """Every parameter gradient of the path: kernels + orchestration vs the float64 backward oracle (itself pinned to autograd
    of the forward oracle by tests/test_oracle_backward.py)."""
    from transmil_b200 import training
    m = build_model(nc, D)
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    x = make_input("randn", B, n, D, 77)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(12))
    ref, _ = OB.backward(sd, x, dlogits.double())
    _, saved = _saved_from_oracle(sd, x)
    w = _weights(T, sd, m._fc1_kind, D, nc)
    g = training.backward(K, sd, w, m._fc1_kind, x.contiguous(), saved, dlogits)
    assert set(g) == set(ref) == set(sd)
    for k in sorted(ref):
        _close(g[k].reshape(ref[k].shape), ref[k], 2e-3, k)


@pytest.mark.parametrize("D,n,B", [(768, 90, 2), (2048, 33, 1), (1024, 50, 1)])
def test_stem_backward_of_every_fc1_form(K, T, D, n, B):
    """fc1 backward alone for the 768-d (two LayerNorms), 2048-d and single-Linear stems vs the float64 oracle."""
    from transmil_b200 import training
    m = build_model(2, D)
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    x = make_input("randn", B, n, D, 78)
    dh = torch.randn(B, n, 512, generator=torch.Generator().manual_seed(14))
    ref_dx, ref = OB.fc1_backward({k: v.double() for k, v in sd.items()}, x.double(), dh.double())
    g = training.fc1_backward(K, sd, m._fc1_kind, x.contiguous(), dh.clone())
    assert set(g) == set(ref) == {k for k in sd if k.startswith("_fc1.")}
    for k in sorted(ref):
        _close(g[k].reshape(ref[k].shape), ref[k], 1e-3, k)
    # a bag that requires grad: d/dx through the stem as well (the reference's autograd propagates it; TransMIL.py:175)
    g = training.fc1_backward(K, sd, m._fc1_kind, x.contiguous(), dh.clone(), want_dx=True)
    _close(g.pop("_dx"), ref_dx, 1e-3, "dx")
    assert set(g) == set(ref)


def test_gradcam_style_hook_on_layer2_norm(K, T, monkeypatch):
    """A forward hook on model.layer2.norm (code/train_loop.py:87) that keeps the activation and registers a gradient hook on it,
    as pytorch_grad_cam does: the hook sees LayerNorm(h2) and the true d(loss)/d(activation), and every parameter gradient is
    what the unhooked backward gives.  Kernels run through the host harness, kept stages come from the forward oracle."""
    import contextlib
    from transmil_b200 import training
    D, nc, n, B = 96, 3, 130, 1
    m = build_model(nc, D)
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    x = make_input("randn", B, n, D, 79)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(15))
    monkeypatch.setattr(training, "_kernels_for", lambda t: (contextlib.nullcontext(), K))
    monkeypatch.setattr(training, "forward_keep", lambda model, t, tr, seeds: (*_saved_from_oracle(sd, t), None))
    monkeypatch.setattr(training, "bind", lambda lib: lib)
    monkeypatch.setattr(T.capi, "lib", lambda: K.L)
    m._weights = lambda: _weights(T, sd, m._fc1_kind, D, nc)
    seen = {}

    def hook(mod, inp, out):
        seen["module"], seen["input"], seen["act"] = mod, inp[0], out.detach().clone()
        out.register_hook(lambda gr: seen.__setitem__("grad", gr.detach().clone()))

    m.layer2.norm.register_forward_hook(hook)
    final = {}

    def final_hook(mod, inp, out):      # visualize_mil.py:227 targets the FINAL LayerNorm (model.norm)
        final["module"], final["input"], final["act"] = mod, inp[0], out.detach().clone()
        out.register_hook(lambda gr: final.__setitem__("grad", gr.detach().clone()))

    m.norm.register_forward_hook(final_hook)
    assert m._hooked_norms() == ["layer2", "norm"]
    logits = training.forward_with_norm_hooks(m, x, ["layer2", "norm"])
    (logits * dlogits).sum().backward()
    # reference: float64 oracle, layer 2 re-run from its LayerNorm output under autograd
    sd64 = {k: v.double() for k, v in sd.items()}
    ref_logits, st = O.forward(sd64, x.double(), return_stages=True, dtype=torch.float64)
    h2 = st["h2"]
    xn = F.layer_norm(h2, (512,), sd64["layer2.norm.weight"], sd64["layer2.norm.bias"], 1e-5).requires_grad_(True)
    h3 = h2 + O.nystrom(sd64, "layer2.attn.", xn)
    hn = F.layer_norm(h3, (512,), sd64["norm.weight"], sd64["norm.bias"], 1e-5)
    hn.retain_grad()
    out = F.linear(hn[:, 0], sd64["_fc.weight"], sd64["_fc.bias"])
    _close(out, ref_logits, 1e-9, "oracle layer-2 replay")
    (out * dlogits.double()).sum().backward()
    assert final["module"] is m.norm and final["act"].shape == (B, n_tokens(n), 512)
    _close(final["input"], h3.detach(), 1e-5, "final-norm hook input (h3)")
    _close(final["act"], hn.detach(), 1e-5, "final-norm hook output (LayerNorm(h3), every row)")
    _close(final["grad"], hn.grad, 1e-5, "gradient delivered to the final-norm activation (cls row only)")
    assert final["grad"][:, 1:].abs().max().item() == 0.0
    assert seen["module"] is m.layer2.norm and seen["act"].shape == (B, n_tokens(n), 512)
    _close(seen["input"], h2, 1e-5, "hook input (layer input h2)")
    _close(seen["act"], xn.detach(), 1e-5, "hook output (LayerNorm(h2))")
    _close(seen["grad"], xn.grad, 2e-3, "gradient delivered to the hooked activation")
    ref, _ = OB.backward(sd, x, dlogits.double())
    for k, prm in m.named_parameters():
        _close(prm.grad, ref[k].reshape(prm.shape), 2e-3, k)


def n_tokens(n):
    return O.plan(n)["S"]


def test_dropout_mask_is_the_same_in_forward_and_backward(K):
    g = torch.Generator().manual_seed(13)
    B, S, P, Cc, p, seed = 2, 50, 6, 512, 0.7, 123456789
    h_in, t = torch.randn(B, S, Cc, generator=g), torch.randn(B, S, Cc, generator=g)
    y = (h_in + t).contiguous()
    K.check(K.L.tm_dropout_resid_fwd(y.data_ptr(), h_in.data_ptr(), B * S * Cc, seed, p, None))
    mask = ((y - h_in).abs() > 0).float()
    keep = mask.mean().item()
    assert abs(keep - (1 - p)) < 0.01
    _close(y, h_in + mask * t / (1 - p), 1e-5, "dropout forward")
    dy = torch.randn(B, S, Cc, generator=g)
    out = torch.empty(B, S + P, Cc)
    K.check(K.L.tm_pad_front_bwd(dy.data_ptr(), B, S, P, Cc, seed, p, out.data_ptr(), None))
    assert out[:, :P].abs().max().item() == 0.0
    _close(out[:, P:], dy * mask / (1 - p), 1e-6, "dropout backward")


def test_stem_dropout_and_dup_cls_forward_kernels(K):
    """tm_dropout_fwd draws the SAME counter-based mask as the to_out Dropout for a (seed, index), in place or not;
    tm_dup_cls_fwd == oracle.square_pad_cls (TransMIL.py:177-186)."""
    g = torch.Generator().manual_seed(21)
    n_el, p, seed = 3 * 70 * 512, 0.6, 424242
    x = torch.randn(n_el, generator=g)
    y = torch.empty_like(x)
    K.check(K.L.tm_dropout_fwd(x.data_ptr(), y.data_ptr(), n_el, seed, p, None))
    zeros, t = torch.zeros(n_el), x.clone()
    K.check(K.L.tm_dropout_resid_fwd(t.data_ptr(), zeros.data_ptr(), n_el, seed, p, None))
    assert torch.equal(y, t)
    mask = (y != 0).float()
    assert abs(mask.mean().item() - (1 - p)) < 0.01
    _close(y, x * mask / (1 - p), 1e-6, "dropout forward")
    K.check(K.L.tm_dropout_fwd(x.data_ptr(), x.data_ptr(), n_el, seed, p, None))
    assert torch.equal(x, y)
    for B, n in [(2, 10), (1, 16), (3, 1)]:
        hf = torch.randn(B, n, 512, generator=g)
        cls = torch.randn(1, 1, 512, generator=g)
        ref, pl = O.square_pad_cls({"cls_token": cls}, hf)
        h0 = torch.full((B, pl["S"], 512), float("nan"))
        K.check(K.L.tm_dup_cls_fwd(hf.data_ptr(), cls.data_ptr(), B, n, pl["dup"], h0.data_ptr(), None))
        assert torch.equal(h0, ref)


def test_768_stem_in_training_mode_forward_and_backward(K, T):
    """train() of the 768-d stem (TransMIL.py:123-126, Dropout(0.6) twice inside _fc1): the library's forward and parameter
    gradients against the oracle with the SAME masks (regenerated from the seeds through tm_dropout_fwd on ones)."""
    from transmil_b200 import training
    B, n, D = 2, 45, 768
    m = build_model(2, D)
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    x = make_input("randn", B, n, D, 80).contiguous()
    seeds = (1234567, 7654321)
    pl = O.plan(n)
    h0 = training.stem768_train_forward(K, sd, x, pl["dup"], seeds)
    masks = []
    for seed, width in zip(seeds, (768, 512)):
        ones = torch.ones(B, n, width)
        K.check(K.L.tm_dropout_fwd(ones.data_ptr(), ones.data_ptr(), ones.numel(), seed, training.STEM768_DROPOUT, None))
        assert abs((ones > 0).float().mean().item() - 0.4) < 0.02
        masks.append(ones)
    sd64 = {k: v.double() for k, v in sd.items()}
    m64 = tuple(t.double() for t in masks)
    ref_h0, _ = O.square_pad_cls(sd64, O.fc1(sd64, x.double(), masks=m64))
    _close(h0, ref_h0, 1e-4, "h0 of the training-mode 768 stem")
    assert not torch.allclose(h0.double(), O.square_pad_cls(sd64, O.fc1(sd64, x.double()))[0], atol=1e-2)
    dh = torch.randn(B, n, 512, generator=torch.Generator().manual_seed(16))
    _, ref = OB.fc1_backward(sd64, x.double(), dh.double(), masks=m64)
    g = training.fc1_backward(K, sd, m._fc1_kind, x, dh.clone(), seeds, training.STEM768_DROPOUT)
    assert set(g) == set(ref)
    for k in sorted(ref):
        _close(g[k].reshape(ref[k].shape), ref[k], 1e-3, k)


def test_ctmil_autograd_path_returns_token_gradient_and_renamed_parameter_gradients(K, T, monkeypatch):
    """CTMIL: tokens of a torch stem enter the library's autograd Function with requires_grad; the backward hands
    d(loss)/d(tokens) back (the conv stem is differentiated by torch) and names the PPEG / classifier gradients pos_layer_0.* /
    _fc2.* (CTMIL.py:82,122).  Kernels through the host harness, kept stages from the forward oracle; reference = autograd of
    the float64 oracle."""
    import contextlib
    from transmil_b200 import training
    nc, N, n = 3, 1, 25     # one tile: the batch-coupled pinv backward is covered by test_whole_backward_against_oracle
    torch.manual_seed(2021)
    m = T.CTMIL(n_classes=nc, in_features=64).eval()
    m.keep_graph = True
    sd = {k: v.detach().contiguous() for k, v in O.ctmil_state(m.state_dict()).items()}
    tok = torch.randn(N, n, 512, generator=torch.Generator().manual_seed(81)).requires_grad_(True)
    dlogits = torch.randn(N, nc, generator=torch.Generator().manual_seed(17))
    monkeypatch.setattr(training, "_kernels_for", lambda t: (contextlib.nullcontext(), K))
    monkeypatch.setattr(training, "forward_keep", lambda model, t, tr, seeds: (*_saved_from_oracle_tokens(sd, t), None))
    monkeypatch.setattr(training, "bind", lambda lib: lib)
    monkeypatch.setattr(T.capi, "lib", lambda: K.L)
    m._weights = lambda: _weights(T, {**sd, "_fc1.0.weight": sd["cls_token"], "_fc1.0.bias": sd["cls_token"]}, T.capi.TM_FC1_TOKENS, 512, nc)
    logits = training.TransMILFunction.apply(m, tok, *m._path_parameters())
    (logits * dlogits).sum().backward()
    prm = {k: v.double().clone().requires_grad_(True) for k, v in sd.items()}
    t64 = tok.detach().double().requires_grad_(True)
    h, pl = O.square_pad_cls(prm, t64)
    h = O.trans_layer(prm, "layer2", O.ppeg(prm, O.trans_layer(prm, "layer1", h), pl["G"]))
    ref_logits = F.linear(F.layer_norm(h, (512,), prm["norm.weight"], prm["norm.bias"], 1e-5)[:, 0], prm["_fc.weight"], prm["_fc.bias"])
    _close(logits, ref_logits.detach(), 1e-5, "logits")
    (ref_logits * dlogits.double()).sum().backward()
    _close(tok.grad, t64.grad, 2e-3, "d(loss)/d(tokens)")
    names = {n_ for n_, _ in m._path_named_parameters()}
    for k, p in m.named_parameters():
        if k not in names:
            assert p.grad is None, k          # conv stem / unused _fc1: not the library's to differentiate
            continue
        ref = prm[k.replace("pos_layer_0.", "pos_layer.").replace("_fc2.", "_fc.")].grad
        _close(p.grad, ref.reshape(p.shape), 2e-3, k)


def _saved_from_oracle_tokens(sd, tok):
    logits, st = O.forward(sd, tok, return_stages=True, tokens=True)
    p = O.plan(tok.shape[1])
    saved = {"plan": p, "h0": st["h0"], "h1": st["h1"], "h2": st["h2"], "h3": st["h3"]}
    for name in ("layer1", "layer2"):
        s = st[name]
        sv = {k: s[k].contiguous() for k in ("q", "k", "v", "q_lm", "k_lm", "attn2", "Z", "W3", "merged")}
        sv["xn"] = F.pad(s["xn"], (0, 0, p["P"], 0)).contiguous()
        sv["M"] = (s["Z"] @ s["W3"]).contiguous()
        saved[name] = sv
    return logits, {k: (v.contiguous() if torch.is_tensor(v) else v) for k, v in saved.items()}


def test_bag_that_requires_grad_gets_its_gradient_through_the_stem(K, T, monkeypatch):
    """x.requires_grad (saliency-style consumers): the reference's autograd reaches the bag through _fc1 (TransMIL.py:175); the
    library's autograd Function returns the same d(loss)/dx, and the parameter gradients are unchanged."""
    import contextlib
    from transmil_b200 import training
    D, nc, n, B = 96, 3, 70, 1
    m = build_model(nc, D)
    m.keep_graph = True
    sd = {k: v.detach().contiguous() for k, v in m.state_dict().items()}
    x = make_input("randn", B, n, D, 82).requires_grad_(True)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(18))
    monkeypatch.setattr(training, "_kernels_for", lambda t: (contextlib.nullcontext(), K))
    monkeypatch.setattr(training, "forward_keep", lambda model, t, tr, seeds: (*_saved_from_oracle(sd, t), None))
    monkeypatch.setattr(training, "bind", lambda lib: lib)
    monkeypatch.setattr(T.capi, "lib", lambda: K.L)
    m._weights = lambda: _weights(T, sd, m._fc1_kind, D, nc)
    logits = training.TransMILFunction.apply(m, x, *m._path_parameters())
    (logits * dlogits).sum().backward()
    ref, ref_dx = OB.backward(sd, x.detach(), dlogits.double())
    _close(x.grad, ref_dx, 2e-3, "d(loss)/dx")
    for k, prm in m.named_parameters():
        _close(prm.grad, ref[k].reshape(prm.shape), 2e-3, k)


def test_hooks_fire_without_a_graph(K, T, monkeypatch):
    """training.fire_hooks_no_graph: hook(norm_module, (input,), LayerNorm(input)) for layer norms and the final norm."""
    import contextlib
    from transmil_b200 import training
    m = build_model(2, 512)
    monkeypatch.setattr(training, "_kernels_for", lambda t: (contextlib.nullcontext(), K))
    g = torch.Generator().manual_seed(22)
    h = {nm: torch.randn(2, 17, 512, generator=g) for nm in ("layer1", "layer2", "norm")}
    with torch.no_grad():
        m.norm.weight.mul_(1.5); m.norm.bias.add_(0.25); m.layer1.norm.weight.mul_(0.5)
    seen = []
    handles = [m._norm_module(nm).register_forward_hook(lambda mod, inp, out: seen.append((mod, inp[0], out))) for nm in ("layer1", "norm")]
    assert m._hooked_norms() == ["layer1", "norm"]
    training.fire_hooks_no_graph(m, m._hooked_norms(), h)
    assert [s_[0] for s_ in seen] == [m.layer1.norm, m.norm]
    for (mod, inp, out), nm in zip(seen, ("layer1", "norm")):
        assert inp is h[nm] and not out.requires_grad
        _close(out, F.layer_norm(h[nm].double(), (512,), mod.weight.double(), mod.bias.double(), 1e-5), 1e-5, nm)
    for hd in handles:
        hd.remove()
    m.norm.register_forward_hook(lambda mod, inp, out: out * 2)
    with pytest.raises(NotImplementedError, match="replacement"):
        training.fire_hooks_no_graph(m, ["norm"], h)
