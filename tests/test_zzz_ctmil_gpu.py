"""GPU suite, sibling models CTMIL / TransformerMIL-Nystrom: the same CUDA path fed with the tokens of their conv
stem (TM_FC1_TOKENS) against the golden vectors of the reference CTMIL module (tests/golden/ctmil_golden.json) -- logits 1e-3,
Y_prob 1e-4 -- first with the oracle's tokens (isolates the library's path), then through the whole module (torch conv stem
with TF32 off, so that cuDNN computes what the fp32 golden did)."""
import json
import os

import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, ROOT
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu
with open(os.path.join(ROOT, "tests", "golden", "ctmil_golden.json")) as f:
    GOLD = json.load(f)
IDS = lambda c: f"D{c['in_features']}-N{c['N']}-H{c['H']}"   # noqa: E731


def _model(T, case, cls="CTMIL"):
    torch.manual_seed(GOLD["weight_seed"])
    return getattr(T, cls)(n_classes=case["n_classes"], in_features=case["in_features"]).eval().cuda()


def _tiles(case):
    g = torch.Generator().manual_seed(case["input_seed"])
    return (torch.rand if case["gen"] == "rand" else torch.randn)(1, case["N"], case["in_features"], case["H"], case["H"], generator=g)


def _err(a, b):
    return (a.detach().double().cpu() - b).abs().max().item()


@pytest.fixture()
def fp32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32 = old


@pytest.mark.parametrize("engine", ["tensor-core", "fp32"])
@pytest.mark.parametrize("case", GOLD["cases"], ids=IDS)
def test_ctmil_matches_reference_golden(built_lib, fp32_convs, case, engine):
    import transmil_b200 as T
    m = _model(T, case)
    m.engine = engine
    x = _tiles(case)
    ref = torch.tensor(case["logits"], dtype=torch.float64)
    pref = torch.tensor(case["y_prob"], dtype=torch.float64)
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    tok = O.ctmil_tokens(sd, x)                                  # the oracle's tokens: the library's path alone
    with torch.no_grad():
        logits = T.TransMIL.forward(m, tok.cuda())
    assert tuple(logits.shape) == (case["N"], case["n_classes"])
    assert _err(logits, ref) <= LOGIT_TOL and _err(torch.softmax(logits, 1), pref) <= PROB_TOL
    with torch.no_grad():
        dev_tok = m.tokens(x.cuda())
        logits2 = m(x.cuda())                                    # whole module: torch conv stem + the library
    assert _err(dev_tok, tok.double()) <= 1e-4 * tok.abs().max().item()
    assert _err(logits2, ref) <= LOGIT_TOL and _err(torch.softmax(logits2, 1), pref) <= PROB_TOL


def test_transformermil_nystrom_forward_is_ctmil_on_its_own_parameters(built_lib, fp32_convs):
    import transmil_b200 as T
    case = GOLD["cases"][0]
    c = _model(T, case)
    t = _model(T, case, "TransformerMIL")
    t.load_state_dict({k: v for k, v in c.state_dict().items() if not k.startswith("_fc1.")}, strict=False)
    x = _tiles(case).cuda()
    with torch.no_grad():
        assert (t(x) - c(x)).abs().max().item() <= 1e-6
    assert _err(t(x), torch.tensor(case["logits"], dtype=torch.float64)) <= LOGIT_TOL


def test_tokens_stem_entry_point(built_lib):
    """tm_fc1_fwd with TM_FC1_TOKENS = duplicate pad + cls of the caller's tokens, bit-exact (CTMIL.py:144-152)."""
    import ctypes as C
    import transmil_b200 as T
    L = T.capi.lib()
    case = GOLD["cases"][0]
    m = _model(T, case)
    for B, n in [(3, 25), (2, 10), (1, 300)]:
        tok = torch.randn(B, n, 512, generator=torch.Generator().manual_seed(n)).cuda()
        p = O.plan(n)
        h0 = torch.full((B, p["S"], 512), float("nan"), device="cuda")
        T.capi.check(L.tm_fc1_fwd(C.byref(m._weights()), tok.data_ptr(), B, n, h0.data_ptr(), None, 0, torch.cuda.current_stream().cuda_stream))
        ref, _ = O.square_pad_cls({"cls_token": m.cls_token.detach()}, tok)
        assert torch.equal(h0, ref)


def test_ctmil_training_step_differentiates_stem_and_path(built_lib):
    """train(): torch differentiates the conv stem from the d(loss)/d(tokens) the library's backward returns; every parameter
    the forward reads gets a finite gradient, the never-called _fc1 gets none, and a few SGD steps drive the loss down."""
    import transmil_b200 as T
    case = GOLD["cases"][0]
    m = _model(T, case).train()
    x = _tiles(case).cuda()
    y = torch.tensor([1, 0, 1], device="cuda")
    opt = torch.optim.SGD(m.parameters(), lr=1e-3)
    losses = []
    for _ in range(6):
        opt.zero_grad(set_to_none=True)
        logits = m(x)
        assert logits.requires_grad and tuple(logits.shape) == (3, 2)
        loss = torch.nn.functional.cross_entropy(logits, y)
        loss.backward()
        for name, prm in m.named_parameters():
            if name.startswith("_fc1."):
                assert prm.grad is None, name
            else:
                assert prm.grad is not None and torch.isfinite(prm.grad).all(), name
        assert m.conv1[0].weight.grad.abs().max() > 0 and m._fc2.weight.grad.abs().max() > 0
        opt.step()
        losses.append(loss.item())
    assert min(losses[-3:]) < losses[0], losses
    # eval-mode graph: the token gradient against autograd of the float64 oracle on the same tokens
    m.eval()
    m.keep_graph = True
    sd = {k: v.detach().double().cpu() for k, v in O.ctmil_state(m.state_dict()).items()}
    tok = torch.randn(2, 25, 512, generator=torch.Generator().manual_seed(3)).cuda().requires_grad_(True)
    dl = torch.randn(2, 2, generator=torch.Generator().manual_seed(4))
    T.TransMIL.forward(m, tok).backward(dl.cuda())
    t64 = tok.detach().double().cpu().requires_grad_(True)
    (O.forward(sd, t64, dtype=torch.float64, tokens=True) * dl.double()).sum().backward()
    assert _err(tok.grad, t64.grad) <= 1e-2 * t64.grad.abs().max().item()
