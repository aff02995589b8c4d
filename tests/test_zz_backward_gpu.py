"""GPU suite of the backward, sorted last on purpose: the nvcc build of csrc/backward.cu and the autograd
path of the drop-in module against the float64 backward oracle on the same seeded bags.

Both engines of the backward run every test (fixture `engine`).  /root/reference is never read here.
"""
import ctypes as C

import pytest
import torch

from common import LOGIT_TOL, build_model, make_input
from oracle import transmil_backward as OB
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu
# max-abs error of a gradient tensor relative to its largest entry, vs the float64 backward oracle, per engine of the backward
# contractions (tm_bgemm_engine): the fp32 CUDA-core kernels, and the tcgen05 engine (tf32 operands -- 10-bit mantissas -- with
# fp32 accumulation, on top of the fp16/tf32 tensor-core forward whose kept stages xn, q, k, v, merged are fp16).  Measured on
# B200 with tools/grad_budget.py over the bags below (profiles/r2/grad_budget.txt): fp32 engine worst parameter 4.0e-6
# (median 1e-6), tensor-core engine worst 4.3e-3 (LayerNorm / bias gradients; median 2e-3).  Budgets = measured x 2.3 ... 5.
GRAD_RTOL = {"simt": 2e-5, "tc": 1e-2}
ENGINES = {"simt": 1, "tc": 0}


@pytest.fixture(params=["simt", "tc"])
def engine(request, built_lib):
    import transmil_b200 as T
    L = T.capi.lib()
    prev = L.tm_bgemm_engine(ENGINES[request.param])
    prev_fwd = T.training.FORWARD_ENGINE
    T.training.FORWARD_ENGINE = request.param
    yield request.param
    L.tm_bgemm_engine(prev)
    T.training.FORWARD_ENGINE = prev_fwd


@pytest.fixture(scope="module")
def T(built_lib):
    import transmil_b200 as T
    assert torch.cuda.is_available()
    assert T.capi.lib().tm_device_supported() == 1
    return T


def _close(a, b, rtol, what):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err = (a - b).abs().max().item()
    bound = rtol * max(b.abs().max().item(), 1e-6)
    assert err <= bound, f"{what}: max abs err {err:.3e} > {bound:.3e}"


@pytest.mark.parametrize("tA,tB", [(False, False), (False, True), (True, False), (True, True)])
def test_bgemm_every_orientation(T, tA, tB, engine):
    from transmil_b200 import training
    K = training.Kernels(stream=torch.cuda.current_stream().cuda_stream)
    g = torch.Generator().manual_seed(3)
    nb, M, N, Kd = 3, 300, 130, 77
    A = torch.randn(nb, *((Kd, M) if tA else (M, Kd)), generator=g).cuda()
    B = torch.randn(nb, *((N, Kd) if tB else (Kd, N)), generator=g).cuda()
    Lt = torch.randn(nb, M, N, generator=g).cuda()
    out = torch.empty(nb, M, N, device="cuda")
    K.gemm(A, (1, M) if tA else (Kd, 1), B, (1, Kd) if tB else (N, 1), out, M, N, Kd, N, nb, (M * Kd, N * Kd, M * N), alpha=0.5, beta=-2.0, L=Lt)
    ref = 0.5 * (A.transpose(1, 2) if tA else A).double() @ (B.transpose(1, 2) if tB else B).double() - 2.0 * Lt.double()
    _close(out, ref, 1e-5 if engine == "simt" else 2e-3, "bgemm")   # tc: operands truncated to tf32


@pytest.mark.parametrize("D,nc,n,B", [(512, 3, 300, 2), (2048, 2, 1000, 2), (1024, 2, 2500, 1)])
def test_parameter_gradients_match_the_backward_oracle(T, D, nc, n, B, engine):
    """eval-mode graph (keep_graph): logits.backward(dlogits) through the library == oracle backward, every parameter."""
    m = build_model(nc, D).cuda()
    m.keep_graph = True
    x = make_input("randn", B, n, D, 77)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(12))
    logits = m(x.cuda())
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    assert (logits.detach().cpu().double() - O.forward(sd, x, dtype=torch.float64)).abs().max().item() <= LOGIT_TOL
    logits.backward(dlogits.cuda())
    ref, _ = OB.backward(sd, x, dlogits.double())
    for k, p in m.named_parameters():
        assert p.grad is not None, k
        _close(p.grad, ref[k].reshape(p.shape), GRAD_RTOL[engine], k)


@pytest.mark.parametrize("D,n,B", [(512, 300, 2), (2048, 200, 1), (768, 100, 1)])
def test_bag_that_requires_grad_gets_its_gradient(T, D, n, B, engine):
    """x.requires_grad: d(loss)/dx through _fc1 (the reference's autograd reaches the bag, TransMIL.py:175) vs the oracle's dx."""
    m = build_model(2, D).cuda()
    m.keep_graph = True
    x = make_input("randn", B, n, D, 83)
    dlogits = torch.randn(B, 2, generator=torch.Generator().manual_seed(19))
    xd = x.cuda().requires_grad_(True)
    m(xd).backward(dlogits.cuda())
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    ref, ref_dx = OB.backward(sd, x, dlogits.double())
    _close(xd.grad, ref_dx, max(GRAD_RTOL[engine], 5e-5), "dx")
    _close(m._fc1[0].weight.grad, ref["_fc1.0.weight"], GRAD_RTOL[engine], "_fc1.0.weight")


def test_training_step_runs_and_learns(T):
    """train(): to_out Dropout(0.7) active, CE loss, a few SGD steps on one bag drive the loss down; gradients finite."""
    m = build_model(2, 1024).cuda()
    m.train()
    x = make_input("rand", 1, 1000, 1024, 5).cuda()
    y = torch.tensor([1], device="cuda")
    opt = torch.optim.SGD(m.parameters(), lr=1e-3)
    losses = []
    for _ in range(8):
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(m(x), y)
        loss.backward()
        assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in m.parameters())
        opt.step()
        losses.append(loss.item())
    assert min(losses[-3:]) < losses[0]


def test_dropout_mask_statistics_and_backward_consistency(T):
    from transmil_b200 import training
    L = T.capi.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(13)
    B, S, P, Cc, p, seed = 2, 500, 12, 512, 0.7, 987654321
    h_in, t = torch.randn(B, S, Cc, generator=g).cuda(), torch.randn(B, S, Cc, generator=g).cuda()
    y = (h_in + t).contiguous()
    T.capi.check(L.tm_dropout_resid_fwd(y.data_ptr(), h_in.data_ptr(), B * S * Cc, seed, p, st))
    mask = ((y - h_in).abs() > 0).float()
    assert abs(mask.mean().item() - (1 - p)) < 5e-3
    dy = torch.randn(B, S, Cc, generator=g).cuda()
    out = torch.empty(B, S + P, Cc, device="cuda")
    T.capi.check(L.tm_pad_front_bwd(dy.data_ptr(), B, S, P, Cc, seed, p, out.data_ptr(), st))
    assert out[:, :P].abs().max().item() == 0.0
    _close(out[:, P:], dy * mask / (1 - p), 1e-6, "dropout backward")


def test_fused_radam_and_lookahead_match_torch_on_device(T):
    """tm_radam_step / tm_lookahead_sync (one launch over every tensor) vs torch.optim.RAdam on the same device."""
    from types import SimpleNamespace
    from transmil_b200 import optim
    a, b = build_model(2, 1024, seed=3).cuda(), build_model(2, 1024, seed=3).cuda()
    ours = optim.create_optimizer(SimpleNamespace(opt="lookahead_radam", lr=1e-2, weight_decay=0.01), a)
    ref = torch.optim.RAdam(optim.add_weight_decay(b, 0.01), lr=1e-2, weight_decay=0.0, foreach=False)
    slow = {}
    for step in range(13):
        g = torch.Generator().manual_seed(100 + step)
        for pa, pb in zip(a.parameters(), b.parameters()):
            gr = (torch.randn(pa.shape, generator=g) * 0.1).cuda()
            pa.grad, pb.grad = gr.clone(), gr.clone()
        ours.step()
        ref.step()
        if (step + 1) % 6 == 0:
            with torch.no_grad():
                for p in b.parameters():
                    if id(p) not in slow:
                        slow[id(p)] = p.data.clone()
                    slow[id(p)].add_(p.data - slow[id(p)], alpha=0.5)
                    p.data.copy_(slow[id(p)])
        for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
            assert (pa - pb).abs().max().item() <= 5e-6 * max(1.0, pb.abs().max().item()), f"step {step + 1} {n}"
