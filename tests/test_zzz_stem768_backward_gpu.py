"""GPU suite: backward of the 768-d stem (TransMIL.py:122-126; two LayerNorms, eval-mode graph via keep_graph) -- every
parameter gradient vs the float64 backward oracle.  Sorted after the established suites: added when no device was at hand;
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
    with pytest.raises(NotImplementedError):          # train mode of this stem needs Dropout(0.6) inside _fc1: refused, not faked
        m.train()(x.cuda())
