"""CPU suite: the explicit backward restatement (oracle/transmil_backward.py) against torch.autograd of
the forward oracle, float64.  The forward oracle is pinned to the reference module (tests/test_oracle.py); the reference's
own gradients are autograd of that module, so this pins the backward oracle a CUDA backward will be tested against."""
import pytest
import torch

from common import build_model, make_input
from oracle import transmil_backward as Bk
from oracle import transmil_oracle as O


def _sd64(m, sharpen=1.0):
    sd = {k: v.detach().double().clone() for k, v in m.state_dict().items()}
    for k in sd:
        if k.endswith("to_qkv.weight"):
            sd[k] *= sharpen
    return sd


def _autograd(sd, x, dlogits):
    prm = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    xr = x.clone().requires_grad_(True)
    h = O.fc1(prm, xr)
    h, p = O.square_pad_cls(prm, h)
    h = O.trans_layer(prm, "layer1", h)
    h = O.ppeg(prm, h, p["G"])
    h = O.trans_layer(prm, "layer2", h)
    hn = torch.nn.functional.layer_norm(h, (512,), prm["norm.weight"], prm["norm.bias"], 1e-5)[:, 0]
    logits = torch.nn.functional.linear(hn, prm["_fc.weight"], prm["_fc.bias"])
    (logits * dlogits).sum().backward()
    return {k: v.grad for k, v in prm.items()}, xr.grad


@pytest.mark.parametrize("sharp", [1.0, 8.0])
def test_pinv_backward_matches_autograd(sharp):
    torch.manual_seed(3)
    A = (torch.randn(2, 8, 256, 256, dtype=torch.float64) * sharp).softmax(-1).requires_grad_(True)
    dZ = torch.randn(2, 8, 256, 256, dtype=torch.float64)
    (O.pinv(A) * dZ).sum().backward()
    got = Bk.pinv_backward(A.detach(), dZ)
    assert (got - A.grad).abs().max().item() <= 1e-9 * A.grad.abs().max().item()


@pytest.mark.parametrize("D,nc,n,B,sharp", [(512, 2, 300, 2, 1.0), (1024, 3, 70, 1, 4.0), (2048, 2, 257, 1, 1.0), (768, 2, 50, 2, 2.0)])
def test_full_backward_matches_autograd(D, nc, n, B, sharp):
    m = build_model(nc, D)
    sd = _sd64(m, sharp)
    x = make_input("randn", B, n, D, 11).double()
    torch.manual_seed(5)
    dlogits = torch.randn(B, nc, dtype=torch.float64)
    ref, ref_dx = _autograd(sd, x, dlogits)
    got, got_dx = Bk.backward(sd, x, dlogits)
    assert set(got) == set(sd), sorted(set(sd) ^ set(got))
    for k, g in ref.items():
        scale = max(g.abs().max().item(), 1e-30)
        assert (got[k].reshape(g.shape) - g).abs().max().item() <= 1e-8 * scale, k
    assert (got_dx - ref_dx).abs().max().item() <= 1e-8 * ref_dx.abs().max().item()


def test_768_stem_backward_with_dropout_masks_matches_autograd():
    """fc1_backward(masks=...) = autograd of oracle.fc1(masks=...): the training mode of the 768-d stem (TransMIL.py:123-126)."""
    m = build_model(2, 768)
    sd = _sd64(m)
    x = make_input("randn", 2, 30, 768, 12).double()
    g = torch.Generator().manual_seed(6)
    m0 = (torch.rand(2, 30, 768, generator=g) >= 0.6).double() / 0.4
    m4 = (torch.rand(2, 30, 512, generator=g) >= 0.6).double() / 0.4
    dh = torch.randn(2, 30, 512, generator=g).double()
    prm = {k: v.clone().requires_grad_(True) for k, v in sd.items() if k.startswith("_fc1.")}
    xr = x.clone().requires_grad_(True)
    (O.fc1(prm, xr, masks=(m0, m4)) * dh).sum().backward()
    dx, got = Bk.fc1_backward(sd, x, dh, masks=(m0, m4))
    for k, v in prm.items():
        assert (got[k].reshape(v.shape) - v.grad).abs().max().item() <= 1e-9 * max(v.grad.abs().max().item(), 1e-30), k
    assert (dx - xr.grad).abs().max().item() <= 1e-9 * xr.grad.abs().max().item()
