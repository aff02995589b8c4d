"""CPU suite: pins the oracle (oracle/transmil_oracle.py) against the fixtures generated
from the unmodified reference module, and -- when /root/reference is present (build
container only) -- against that module executed live."""
import math

import numpy as np
import pytest
import torch

from common import checksum, load_golden, make_input
from oracle import ref_loader, transmil_oracle as O

GOLD = load_golden()
SMALL = [c for c in GOLD["cases"] if c["n"] * c["B"] <= 6000 or c["B"] == 64]


def _sd_for(case, cache={}):
    """Reference-layout weights re-created from the seed through the PRODUCT module's
    constructor (same sub-modules, same order as TransMIL.py:91-155)."""
    key = (case["in_features"], case["n_classes"])
    if key not in cache:
        import transmil_b200 as T
        torch.manual_seed(GOLD["weight_seed"])
        cache[key] = T.TransMIL(n_classes=key[1], in_features=key[0]).state_dict()
    return cache[key]


@pytest.mark.parametrize("case", SMALL, ids=lambda c: f"D{c['in_features']}-n{c['n']}-B{c['B']}")
def test_oracle_matches_golden(case):
    torch.set_float32_matmul_precision("highest")
    sd = _sd_for(case)
    for k, v in case["param_checksums"].items():
        assert checksum(sd[k]) == pytest.approx(v, rel=0, abs=0), f"seeded init drifted at {k}"
    x = make_input(case["gen"], case["B"], case["n"], case["in_features"], case["input_seed"])
    logits, stages = O.forward(sd, x, return_stages=True)
    # same host ISA => bit-identical; across CPU generations allow fp32 reduction-order noise
    assert torch.allclose(logits.double(), torch.tensor(case["logits"], dtype=torch.float64), atol=2e-5, rtol=0)
    assert torch.allclose(O.y_prob(logits).double(), torch.tensor(case["y_prob"], dtype=torch.float64), atol=1e-5, rtol=0)
    assert O.plan(case["n"]) == case["plan"]
    got = checksum(stages["layer1"]["Z"])
    assert got[1] == pytest.approx(case["stage_checksums"]["Z1"][1], rel=1e-4)


def test_attn_probe_and_padding_contract():
    """return_attn tuple, FRONT pad indexing attn[0,:,padding+1,padding+1:...] (TransMIL.py:200,221)."""
    case = next(c for c in GOLD["cases"] if c["n"] == 257)
    sd = _sd_for(case)
    x = make_input(case["gen"], case["B"], case["n"], case["in_features"], case["input_seed"])
    logits, (attn, padding) = O.forward(sd, x, return_attn=True)
    p = O.plan(case["n"])
    assert padding == case["padding"] == p["padding"] == p["P"]
    assert tuple(attn.shape) == (1, 8, p["Sp"], p["Sp"])
    probe = attn[0, :, padding + 1, padding + 1:padding + 1 + min(case["n"], 64)].double()
    assert torch.allclose(probe, torch.tensor(case["attn_probe"], dtype=torch.float64), atol=1e-6, rtol=1e-4)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("D,nc,n,B", [(2048, 2, 300, 2), (1024, 3, 257, 1), (768, 2, 100, 1), (512, 6, 5, 3)])
def test_oracle_bit_identical_to_reference_module(D, nc, n, B):
    torch.set_float32_matmul_precision("highest")
    model = ref_loader.build_reference(nc, D)
    x = torch.randn(B, n, D, generator=torch.Generator().manual_seed(7))
    ref, (attn, padding) = ref_loader.reference_forward(model, x, return_attn=True)
    mine, (attn2, padding2) = O.forward(model.state_dict(), x, return_attn=True)
    assert torch.equal(ref, mine)
    assert torch.equal(attn, attn2)
    assert padding == padding2
    # the dense attention does not influence the logits
    assert torch.equal(O.forward(model.state_dict(), x), ref)


def test_plan_formulas():
    for n in list(range(1, 3000)) + [5999, 6000, 20000, 50000, 100000, 200000, 299999]:
        p = O.plan(n)
        G = int(np.ceil(np.sqrt(n)))            # TransMIL.py:178
        assert p["G"] == G and p["N"] == G * G and p["dup"] == G * G - n and p["S"] == G * G + 1
        assert p["Sp"] % 256 == 0 and 0 <= p["P"] < 256 and p["Sp"] == p["S"] + p["P"]
        assert p["l"] == math.ceil(p["S"] / 256) and p["l"] * 256 == p["Sp"]
        assert p["dup"] <= n


def test_pinv_is_not_converged_and_uses_global_scalar():
    """6 iterations leave a visible residual, and the init scalar couples the batch."""
    g = torch.Generator().manual_seed(3)
    A = torch.softmax(torch.randn(2, 8, 256, 256, generator=g) * 0.3, -1)
    Z = O.pinv(A)
    res = (A @ Z @ A - A).norm() / A.norm()
    assert 1e-4 < float(res) < 0.5
    Z0 = O.pinv(A[:1])
    assert not torch.equal(Z0, Z[:1])


def test_front_pad_rows_count_in_attn3_softmax():
    """Zero-score pad rows take part in the attn3 denominator (no mask is passed)."""
    sd = _sd_for(GOLD["cases"][4])
    x = make_input("randn", 1, 255, 512, 5)
    _, st = O.forward(sd, x, return_stages=True)
    l1 = st["layer1"]
    P = l1["P"]
    assert P > 0 and float(l1["k"][:, :, :P].abs().max()) == 0.0 and float(l1["v"][:, :, :P].abs().max()) == 0.0
    sim3 = l1["q_lm"] @ l1["k"].transpose(-1, -2)
    w_with = torch.softmax(sim3, -1) @ l1["v"]
    w_without = torch.softmax(sim3[..., P:], -1) @ l1["v"][:, :, P:]
    assert torch.allclose(w_with, l1["W3"], atol=1e-6)
    assert not torch.allclose(w_without, l1["W3"], atol=1e-4)


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference only exists in the build container")
def test_768_stem_train_mode_masks_sit_where_the_reference_drops():
    """Training mode of the 768-d stem (TransMIL.py:123-126): the reference's `_fc1` in train() with a seeded CPU RNG ==
    oracle.fc1 with the two keep masks drawn from the same RNG stream in the same order (F.dropout of ones = mask / (1-p))."""
    import torch.nn.functional as F
    torch.set_float32_matmul_precision("highest")
    model = ref_loader.build_reference(2, 768)
    x = torch.randn(2, 40, 768, generator=torch.Generator().manual_seed(8))
    model._fc1.train()
    torch.manual_seed(99)
    with torch.no_grad():
        ref = model._fc1(x)
    model._fc1.eval()
    torch.manual_seed(99)
    m0 = F.dropout(torch.ones(2, 40, 768), 0.6, True)
    m4 = F.dropout(torch.ones(2, 40, 512), 0.6, True)
    assert 0.3 < (m0 > 0).float().mean().item() < 0.5 and m0.max().item() == pytest.approx(2.5)
    mine = O.fc1(model.state_dict(), x, masks=(m0, m4))
    assert torch.allclose(ref, mine, atol=1e-6, rtol=0)
    assert not torch.allclose(ref, O.fc1(model.state_dict(), x), atol=1e-2)
