"""GPU suite: the device-tuned HBM-bound kernels of the backward (csrc/backward_fast.cu) against the plain kernels of
csrc/backward.cu (tm_bgemm_engine(1), the ones the CPU suite runs through tests/cuda_emu) and against float64 torch."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib(built_lib):
    import transmil_b200 as T
    return T, T.training.bind(T.capi.lib())


def _both(L, fn):
    out = []
    for engine in (1, 0):
        prev = L.tm_bgemm_engine(engine)
        try:
            out.append(fn())
            torch.cuda.synchronize()
        finally:
            L.tm_bgemm_engine(prev)
    return out


@pytest.mark.parametrize("B,Sp", [(1, 256), (2, 1280), (1, 50432), (3, 768)])
def test_conv33_bwd(lib, B, Sp):
    T, L = lib
    g = torch.Generator().manual_seed(B * 100 + Sp)
    dO = torch.randn(B, 8, Sp, 64, generator=g).cuda()
    v = torch.randn(B, 8, Sp, 64, generator=g).cuda()
    w = (torch.randn(8, 1, 33, 1, generator=g) * 0.2).cuda()
    st = torch.cuda.current_stream().cuda_stream

    def run():
        dv, gw = torch.full_like(v, float("nan")), torch.full_like(w, float("nan"))
        sc = torch.empty(int(L.tm_conv33_bwd_scratch_floats(B, Sp)), device="cuda")
        T.capi.check(L.tm_conv33_bwd(dO.data_ptr(), v.data_ptr(), w.data_ptr(), B, Sp, dv.data_ptr(), gw.data_ptr(), sc.data_ptr(), st))
        return dv, gw

    (dv0, gw0), (dv1, gw1) = _both(L, run)
    # float64 reference through autograd of the forward op (Conv2d(8,8,(33,1),groups=8,padding=(16,0)), cross-correlation)
    vd = v.double().cpu().requires_grad_(True)
    wd = w.double().cpu().requires_grad_(True)
    out = torch.nn.functional.conv2d(vd, wd, padding=(16, 0), groups=8)
    out.backward(dO.double().cpu())
    for got, name in ((dv0, "plain dv"), (dv1, "fast dv")):
        assert (got.double().cpu() - vd.grad).abs().max().item() <= 1e-5 * vd.grad.abs().max().item(), name
    for got, name in ((gw0, "plain g_w"), (gw1, "fast g_w")):
        assert (got.double().cpu() - wd.grad).abs().max().item() <= 2e-5 * wd.grad.abs().max().item(), name


@pytest.mark.parametrize("rows,len_", [(1000, 256), (777, 512), (300, 1024), (64, 1280), (50, 100)])
def test_softmax_rows_and_backward(lib, rows, len_):
    T, L = lib
    g = torch.Generator().manual_seed(rows + len_)
    X = (torch.randn(rows, len_, generator=g) * 3).cuda()
    dP = torch.randn(rows, len_, generator=g).cuda()
    st = torch.cuda.current_stream().cuda_stream

    def run():
        P = X.clone()
        T.capi.check(L.tm_softmax_rows(P.data_ptr(), rows, len_, st))
        d = dP.clone()
        T.capi.check(L.tm_softmax_rows_bwd(P.data_ptr(), d.data_ptr(), rows, len_, st))
        return P, d

    (P0, d0), (P1, d1) = _both(L, run)
    Pr = torch.softmax(X.double(), -1)
    dr = Pr * (dP.double() - (dP.double() * Pr).sum(-1, keepdim=True))
    for P, d, name in ((P0, d0, "plain"), (P1, d1, "fast")):
        assert (P.double() - Pr).abs().max().item() < 2e-6, name
        assert (d.double() - dr).abs().max().item() < 1e-5 * max(1.0, dr.abs().max().item()), name


@pytest.mark.parametrize("B,G", [(1, 32), (2, 45), (1, 142), (3, 7)])
def test_ppeg_backward(lib, B, G):
    T, L = lib
    from common import build_model
    m = build_model(2, 512).cuda()
    w = m._weights()
    S = G * G + 1
    g = torch.Generator().manual_seed(B * 10 + G)
    h_in = torch.randn(B, S, 512, generator=g).cuda()
    dy = torch.randn(B, S, 512, generator=g).cuda()
    st = torch.cuda.current_stream().cuda_stream
    shapes = [(512, 1, 7, 7), (512,), (512, 1, 5, 5), (512,), (512, 1, 3, 3), (512,)]

    def run():
        dh = torch.full_like(h_in, float("nan"))
        gs = [torch.full(s, float("nan"), device="cuda") for s in shapes]
        sc = torch.empty(int(L.tm_ppeg_bwd_scratch_floats(B, G)), device="cuda")
        T.capi.check(L.tm_ppeg_bwd(C.byref(w), h_in.data_ptr(), dy.data_ptr(), B, G, dh.data_ptr(), *[t.data_ptr() for t in gs],
                                   sc.data_ptr(), st))
        return [dh] + gs

    plain, fast = _both(L, run)
    # float64 reference: autograd through x + dw7(x) + dw5(x) + dw3(x) on the GxG image (TransMIL.py:60-75)
    p = m.pos_layer
    ws = [t.detach().double().cpu().requires_grad_(True) for t in (p.proj.weight, p.proj.bias, p.proj1.weight, p.proj1.bias,
                                                                   p.proj2.weight, p.proj2.bias)]
    x = h_in.double().cpu().requires_grad_(True)
    img = x[:, 1:].transpose(1, 2).reshape(B, 512, G, G)
    F = torch.nn.functional
    out = img + F.conv2d(img, ws[0], ws[1], padding=3, groups=512) + F.conv2d(img, ws[2], ws[3], padding=2, groups=512) + \
        F.conv2d(img, ws[4], ws[5], padding=1, groups=512)
    y = torch.cat([x[:, :1], out.flatten(2).transpose(1, 2)], 1)
    y.backward(dy.double().cpu())
    ref = [x.grad] + [t.grad for t in ws]
    for got_set, name in ((plain, "plain"), (fast, "fast")):
        for got, r, what in zip(got_set, ref, ["dh", "w7", "b7", "w5", "b5", "w3", "b3"]):
            err = (got.double().cpu() - r).abs().max().item()
            assert err <= 3e-5 * max(1.0, r.abs().max().item()), f"{name} {what}: {err:.3e}"
