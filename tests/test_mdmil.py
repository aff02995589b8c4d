"""CPU suite, sibling model of the same path: `MDMIL(n_classes)` keeps the reference's constructor, names,
state_dict layout and seeded init (code/models/MDMIL.py:60-113), is selected by `Model.name: MDMIL`, and the oracle restated
for TransMIL reproduces the reference MDMIL module (golden fixture made by tests/golden/make_mdmil_golden.py)."""
import json
import os

import pytest
import torch

from common import ROOT, checksum, make_input
from oracle import ref_loader, transmil_oracle as O

with open(os.path.join(ROOT, "tests", "golden", "mdmil_golden.json")) as f:
    GOLD = json.load(f)


def build_mdmil(nc, seed=2021):
    import transmil_b200 as T
    torch.manual_seed(seed)
    return T.MDMIL(n_classes=nc).eval()


def oracle_state(sd):
    return {k.replace("_fc2.", "_fc."): v for k, v in sd.items()}


def test_state_dict_layout_and_seeded_init():
    for case in GOLD["cases"]:
        sd = build_mdmil(case["n_classes"], GOLD["weight_seed"]).state_dict()
        assert list(sd) == list(case["param_checksums"])                      # same keys in the same order
        for k, v in case["param_checksums"].items():
            assert checksum(sd[k]) == v, k
    sd = build_mdmil(2).state_dict()
    assert tuple(sd["_fc1.0.weight"].shape) == (512, 1024) and tuple(sd["_fc2.weight"].shape) == (2, 512) and "_fc.weight" not in sd


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference only exists in the build container")
def test_init_bit_identical_to_live_reference_module():
    import importlib.util
    import sys
    sys.path.insert(0, os.path.dirname(ref_loader.__file__))
    spec = importlib.util.spec_from_file_location("_reference_MDMIL", "/root/reference/code/models/MDMIL.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    torch.manual_seed(7)
    ref = mod.MDMIL(n_classes=6).state_dict()
    own = build_mdmil(6, seed=7).state_dict()
    assert list(ref) == list(own)
    for k in ref:
        assert torch.equal(ref[k], own[k]), k


def test_plugin_loader_selects_mdmil():
    from transmil_b200 import plugin
    import transmil_b200 as T
    m = plugin.load_model({"name": "MDMIL", "n_classes": 3, "in_features": 1024, "backbone": "x"})   # only n_classes is a ctor arg
    assert isinstance(m, T.MDMIL) and m.n_classes == 3 and m.in_features == 1024
    if not torch.cuda.is_available():
        with pytest.raises(T.capi.TransmilError, match="no CPU path"):
            m.eval()(torch.zeros(1, 10, 1024))


@pytest.mark.parametrize("case", GOLD["cases"], ids=lambda c: f"n{c['n']}-B{c['B']}")
def test_oracle_reproduces_reference_mdmil(case):
    sd = oracle_state(build_mdmil(case["n_classes"], GOLD["weight_seed"]).state_dict())
    x = make_input(case["gen"], case["B"], case["n"], 1024, case["input_seed"])
    logits = O.forward(sd, x)
    assert (logits.double() - torch.tensor(case["logits"], dtype=torch.float64)).abs().max() <= 2e-6
    assert O.plan(case["n"]) == case["plan"]
