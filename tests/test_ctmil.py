"""CPU suite, sibling models of the same path: `CTMIL` (code/models/CTMIL.py:76-165) and the Nystrom forward of
`TransformerMIL` (the commented-out one, code/models/TransformerMIL.py:155-183) keep the reference's constructors, names,
state_dict layouts and seeded init, are selected by `Model.name`, and the oracle (conv stem restated, tokens into the TransMIL
oracle) reproduces the reference CTMIL module (golden fixture made by tests/golden/make_ctmil_golden.py)."""
import json
import os

import pytest
import torch

from common import ROOT, checksum
from oracle import ref_loader, transmil_oracle as O

with open(os.path.join(ROOT, "tests", "golden", "ctmil_golden.json")) as f:
    GOLD = json.load(f)
IDS = lambda c: f"D{c['in_features']}-N{c['N']}-H{c['H']}"   # noqa: E731


def build(cls, nc, D, seed=2021):
    import transmil_b200 as T
    torch.manual_seed(seed)
    return getattr(T, cls)(n_classes=nc, in_features=D).eval()


def tiles(case):
    g = torch.Generator().manual_seed(case["input_seed"])
    return (torch.rand if case["gen"] == "rand" else torch.randn)(1, case["N"], case["in_features"], case["H"], case["H"], generator=g)


@pytest.mark.parametrize("case", GOLD["cases"], ids=IDS)
def test_state_dict_layout_and_seeded_init(case):
    sd = build("CTMIL", case["n_classes"], case["in_features"], GOLD["weight_seed"]).state_dict()
    assert list(sd) == list(case["param_checksums"])                      # same keys in the same order
    for k, v in case["param_checksums"].items():
        assert checksum(sd[k]) == v, k


@pytest.mark.skipif(not ref_loader.available(), reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("name,D", [("CTMIL", 768), ("CTMIL", 1024), ("CTMIL", 96), ("TransformerMIL", 768), ("TransformerMIL", 384),
                                    ("TransformerMIL", 1024)])
def test_init_bit_identical_to_live_reference_module(name, D):
    mod = ref_loader.load_reference_sibling(name)
    torch.manual_seed(7)
    ref = getattr(mod, name)(n_classes=6, in_features=D).state_dict()
    own = build(name, 6, D, seed=7).state_dict()
    assert list(ref) == list(own)
    for k in ref:
        assert torch.equal(ref[k], own[k]), k


@pytest.mark.parametrize("case", GOLD["cases"], ids=IDS)
def test_oracle_reproduces_reference_ctmil(case):
    torch.set_float32_matmul_precision("highest")
    sd = build("CTMIL", case["n_classes"], case["in_features"], GOLD["weight_seed"]).state_dict()
    x = tiles(case)
    tok = O.ctmil_tokens(sd, x)
    assert tuple(tok.shape) == (case["N"], case["n_tokens"], 512) and O.plan(case["n_tokens"]) == case["plan"]
    assert checksum(tok)[1] == pytest.approx(case["tokens_checksum"][1], rel=1e-5)
    logits = O.ctmil_forward(sd, x)
    assert (logits.double() - torch.tensor(case["logits"], dtype=torch.float64)).abs().max() <= 2e-5
    # the commented Nystrom forward of TransformerMIL.py:155-183 is the same code on TransformerMIL's parameters
    t = build("TransformerMIL", case["n_classes"], case["in_features"])
    t.load_state_dict({k: v for k, v in sd.items() if not k.startswith("_fc1.")}, strict=False)
    assert torch.equal(O.ctmil_forward(t.state_dict(), x), logits)


def test_tokens_are_a_view_not_a_transpose():
    """CTMIL.py:142 reinterprets [N,512,h,h] as [N,h*h,512]: token t of a tile is 512 CONSECUTIVE floats of the NCHW buffer."""
    case = GOLD["cases"][0]
    sd = build("CTMIL", case["n_classes"], case["in_features"], GOLD["weight_seed"]).state_dict()
    tok = O.ctmil_tokens(sd, tiles(case))
    assert [float(v) for v in tok[0, 0, :16].double()] == pytest.approx(case["tokens_probe"], abs=1e-6)
    N, n, _ = tok.shape
    h = int(round(n ** 0.5))
    nchw = tok.reshape(N, 512, h, h)
    assert torch.equal(nchw.reshape(N, -1)[:, 512:1024], tok[:, 1])


@pytest.mark.parametrize("name", ["CTMIL", "TransformerMIL"])
def test_plugin_loader_selects_the_sibling(name):
    from transmil_b200 import plugin
    import transmil_b200 as T
    m = plugin.load_model({"name": name, "n_classes": 3, "in_features": 768, "backbone": "x"})
    assert isinstance(m, getattr(T, name)) and m.n_classes == 3 and m.in_features == 512 and m._fc1_kind == T.capi.TM_FC1_TOKENS
    path = {n for n, _ in m._path_named_parameters()}
    assert len(path) == 23 and not any(n.startswith(("conv", "_fc1", "fc1", "transformer")) for n in path)
    if not torch.cuda.is_available():
        with pytest.raises(T.capi.TransmilError, match="no CPU path"):
            m.eval()(torch.zeros(1, 2, 768, 8, 8))
    if name == "TransformerMIL":
        with pytest.raises(NotImplementedError, match="Nystrom"):
            m.forward_plain(torch.zeros(1, 4, 768))


def test_tokens_stem_is_declared_and_validated():
    """TM_FC1_TOKENS in the C ABI: workspace query accepts it without fc1 pointers, refuses in_features != 512."""
    import ctypes as C
    import transmil_b200 as T
    L = T.capi.lib()
    w = T.capi.Weights()
    w.fc1_kind, w.in_features, w.n_classes = T.capi.TM_FC1_TOKENS, 512, 2
    for f in ("cls_token", "norm_w", "norm_b", "fc_w", "fc_b"):
        setattr(w, f, 256)
    nbytes = C.c_size_t()
    T.capi.check(L.tm_workspace_bytes(4, 169, C.byref(w), 0, C.byref(nbytes)))
    assert nbytes.value > 0
    w.in_features = 768
    with pytest.raises(ValueError, match="512-d tokens"):
        T.capi.check(L.tm_workspace_bytes(4, 169, C.byref(w), 0, C.byref(nbytes)))
    assert "TM_FC1_TOKENS = 3" in open(os.path.join(ROOT, "include", "transmil_b200.h")).read()
