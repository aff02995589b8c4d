"""GPU suite, stress set of SURVEY 8c: `to_qkv.weight *= 8` ("sharp softmax"), rand and randn bags.

The path is ill-conditioned there: on the CPU, rounding only the operands to TF32 -- the reference's own GPU setting
(code/train.py:94) -- moves the logits by 3e-3 ... 1.8e-2, fp16 operands by 1e-3 ... 2.4e-2 (tools/precision_budget.py with the
scaled weights; table in DESIGN.md, Precision).  So the 1e-3 / 1e-4 gate of the default-init bags is asserted here for the fp32
cross-check engine (TM_FLAG_SIMT_ONLY), and the tensor-core engine is held to the budget TF32 itself would need (5e-2).
Sorted after the established suites: written when no device was at hand."""
import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, build_model, make_input
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D,n,gen", [(1024, 3000, "rand"), (2048, 1000, "randn")])
def test_sharp_softmax_stress(built_lib, D, n, gen):
    import transmil_b200 as T
    m = build_model(2, D)
    with torch.no_grad():
        for layer in (m.layer1, m.layer2):
            layer.attn.to_qkv.weight.mul_(8.0)
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = make_input(gen, 1, n, D, 7)
    ref = O.forward(sd, x, dtype=torch.float64)
    m = m.cuda()
    errs = {}
    for engine, flags in (("simt", T.capi.TM_FLAG_SIMT_ONLY), ("default", 0)):
        m.flags = flags
        with torch.no_grad():
            logits = m(x.cuda()).cpu().double()
        assert torch.isfinite(logits).all()
        errs[engine] = ((logits - ref).abs().max().item(), (logits.softmax(1) - ref.softmax(1)).abs().max().item())
    print("sharp-softmax stress", D, n, gen, errs)
    assert errs["simt"][0] <= LOGIT_TOL and errs["simt"][1] <= PROB_TOL, errs
    assert errs["default"][0] <= 5e-2, errs
