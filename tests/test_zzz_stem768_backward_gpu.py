"""GPU suite: backward of the 768-d stem (TransMIL.py:122-126; two LayerNorms; eval-mode graph via keep_graph, and train()
with its two Dropout(0.6)) -- every parameter gradient vs the float64 backward oracle.  Sorted after the established suites: added when no device was at hand;
the same source passes on the host through tests/cuda_emu (tests/test_backward_emu.py)."""
import pytest
import torch

from common import LOGIT_TOL, build_model, make_input
from oracle import transmil_backward as OB
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu
GRAD_RTOL = 1e-2   # tensor-core engine: measured worst 4.3e-3 (profiles/r2/grad_budget.txt)


def test_768_stem_parameter_gradients(built_lib):
    D, nc, n, B = 768, 2, 400, 1
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
        a, b = p.grad.detach().double().cpu(), ref[k].reshape(p.shape).double()
        assert (a - b).abs().max().item() <= GRAD_RTOL * max(b.abs().max().item(), 1e-6), k


def test_768_stem_training_mode_on_device(built_lib):
    """train() of the 768-d stem: Dropout(0.6) twice inside _fc1 (TransMIL.py:123-126).  The stem's forward and parameter
    gradients against the oracle with the masks the kernels drew (regenerated from the seeds), then whole training steps:
    finite gradients for every parameter, two forwards of one bag differ (masks are redrawn), the loss goes down."""
    import transmil_b200 as T
    from transmil_b200 import training
    L = T.capi.lib()
    B, n, D = 2, 300, 768
    m = build_model(2, D).cuda()
    p = {k: t.detach() for k, t in m.named_parameters()}
    x = make_input("randn", B, n, D, 80).cuda()
    K = training.Kernels(stream=torch.cuda.current_stream().cuda_stream)
    seeds = (1234567, 7654321)
    pl = O.plan(n)
    h0 = training.stem768_train_forward(K, p, x, pl["dup"], seeds)
    masks = []
    for seed, width in zip(seeds, (768, 512)):
        ones = torch.ones(B, n, width, device="cuda")
        K.check(L.tm_dropout_fwd(ones.data_ptr(), ones.data_ptr(), ones.numel(), seed, training.STEM768_DROPOUT, K.st))
        assert abs((ones > 0).float().mean().item() - 0.4) < 0.01
        masks.append(ones.double().cpu())
    sd64 = {k: v.detach().double().cpu() for k, v in m.state_dict().items()}
    ref_h0, _ = O.square_pad_cls(sd64, O.fc1(sd64, x.double().cpu(), masks=tuple(masks)))
    assert (h0.double().cpu() - ref_h0).abs().max().item() <= 2.5e-3 * ref_h0.abs().max().item()   # tf32 operands (stage budget)
    dh = torch.randn(B, n, 512, generator=torch.Generator().manual_seed(16))
    _, ref = OB.fc1_backward(sd64, x.double().cpu(), dh.double(), masks=tuple(masks))
    g = training.fc1_backward(K, p, m._fc1_kind, x, dh.cuda(), seeds, training.STEM768_DROPOUT)
    for k in sorted(ref):
        a, b = g[k].double().cpu().reshape(ref[k].shape), ref[k]
        assert (a - b).abs().max().item() <= GRAD_RTOL * max(b.abs().max().item(), 1e-6), k
    # whole steps through autograd
    m.train()
    y = torch.tensor([1, 0], device="cuda")
    assert not torch.equal(m(x).detach(), m(x).detach())
    opt = torch.optim.SGD(m.parameters(), lr=2e-3)
    losses = []
    for _ in range(10):
        opt.zero_grad(set_to_none=True)
        loss = torch.nn.functional.cross_entropy(m(x), y)
        loss.backward()
        assert all(q.grad is not None and torch.isfinite(q.grad).all() for q in m.parameters())
        opt.step()
        losses.append(loss.item())
    assert min(losses[-4:]) < losses[0], losses
    # train() under no_grad: the reference's Dropout layers stay active there too; eval() is deterministic
    with torch.no_grad():
        a, b = m(x), m(x)
        assert not a.requires_grad and not torch.equal(a, b)
        m.eval()
        assert torch.equal(m(x), m(x))
