"""GPU suite, sibling model MDMIL: the same CUDA path behind `MDMIL(n_classes)` against the golden vectors of
the reference module (tests/golden/mdmil_golden.json) -- logits 1e-3, Y_prob 1e-4, probe row of the returned attn2."""
import json
import os

import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, ROOT, make_input

pytestmark = pytest.mark.gpu
with open(os.path.join(ROOT, "tests", "golden", "mdmil_golden.json")) as f:
    GOLD = json.load(f)


def _model(T, nc):
    torch.manual_seed(GOLD["weight_seed"])
    return T.MDMIL(n_classes=nc).eval().cuda()


def _err(a, b):
    return (a.detach().double().cpu() - b).abs().max().item()


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"n{c['n']}-B{c['B']}")
def test_mdmil_matches_reference_golden(built_lib, case):
    import transmil_b200 as T
    m = _model(T, case["n_classes"])
    x = make_input(case["gen"], case["B"], case["n"], 1024, case["input_seed"]).cuda()
    with torch.no_grad():
        logits, attn2 = m(x)
    ref = torch.tensor(case["logits"], dtype=torch.float64)
    assert _err(logits, ref) <= LOGIT_TOL
    assert _err(torch.softmax(logits, 1), torch.tensor(case["y_prob"], dtype=torch.float64)) <= PROB_TOL
    p = case["plan"]
    assert tuple(attn2.shape) == (case["B"], 8, p["Sp"], p["Sp"])
    pad, H = p["padding"], min(case["n"], 64)
    probe = attn2[0, :, pad + 1, pad + 1:pad + 1 + H]
    pref = torch.tensor(case["attn_probe"], dtype=torch.float64)
    assert _err(probe, pref) <= 2e-6 + 2e-2 * pref.abs().max().item()


def test_mdmil_training_step_fills_fc2_gradients(built_lib):
    import transmil_b200 as T
    m = _model(T, 2).train()
    x = make_input("randn", 1, 300, 1024, 11).cuda()
    logits, attn = m(x)
    assert attn is None and logits.requires_grad
    torch.nn.functional.cross_entropy(logits, torch.tensor([1], device="cuda")).backward()
    for name, prm in m.named_parameters():
        assert prm.grad is not None and torch.isfinite(prm.grad).all(), name
    assert m._fc2.weight.grad.abs().max() > 0
