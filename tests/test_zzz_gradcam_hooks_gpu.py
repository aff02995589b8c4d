"""GPU suite: GradCAM-style consumer of `model.layer2.norm` (code/train_loop.py:87, code/test_visualize.py:121-127) -- a forward
hook that keeps the activation and registers a gradient hook on it -- served by the library's autograd path in eval mode.
Sorted after the established suites: written when no device was at hand; the same choreography and kernels pass on the host
(tests/test_backward_emu.py::test_gradcam_style_hook_on_layer2_norm)."""
import pytest
import torch
import torch.nn.functional as F

from common import LOGIT_TOL, build_model, make_input
from oracle import transmil_backward as OB
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu


def _close(a, b, rtol, what):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    err, bound = (a - b).abs().max().item(), rtol * max(b.abs().max().item(), 1e-6)
    assert err <= bound, f"{what}: {err:.3e} > {bound:.3e}"


def test_hook_on_layer2_norm_sees_activation_and_gradient(built_lib):
    D, nc, n, B = 512, 3, 300, 1
    m = build_model(nc, D).cuda()          # eval mode, autograd on: what a CAM computation runs under
    x = make_input("randn", B, n, D, 79)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(15))
    seen = {}

    def hook(mod, inp, out):
        seen["act"] = out.detach().clone()
        out.register_hook(lambda gr: seen.__setitem__("grad", gr.detach().clone()))

    handle = m.layer2.norm.register_forward_hook(hook)
    logits = m(x.cuda())
    assert logits.requires_grad
    (logits * dlogits.cuda()).sum().backward()
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    sd64 = {k: v.double() for k, v in sd.items()}
    ref_logits, st = O.forward(sd64, x.double(), return_stages=True, dtype=torch.float64)
    assert (logits.detach().cpu().double() - ref_logits).abs().max().item() <= LOGIT_TOL
    h2 = st["h2"]
    xn = F.layer_norm(h2, (512,), sd64["layer2.norm.weight"], sd64["layer2.norm.bias"], 1e-5).requires_grad_(True)
    h3 = h2 + O.nystrom(sd64, "layer2.attn.", xn)
    out = F.linear(F.layer_norm(h3, (512,), sd64["norm.weight"], sd64["norm.bias"], 1e-5)[:, 0], sd64["_fc.weight"], sd64["_fc.bias"])
    (out * dlogits.double()).sum().backward()
    assert seen["act"].shape == (B, O.plan(n)["S"], 512) and seen["act"].is_cuda
    _close(seen["act"], xn, 1e-3, "hooked activation")   # the tensor-core forward keeps LayerNorm's output as fp16 (11-bit mantissa)
    _close(seen["grad"], xn.grad, 1e-2, "gradient of the hooked activation")
    ref, _ = OB.backward(sd, x, dlogits.double())
    for k, p in m.named_parameters():
        _close(p.grad, ref[k].reshape(p.shape), 1e-2, k)   # tensor-core engine: measured worst 4.3e-3 (profiles/r2/grad_budget.txt)
    handle.remove()
    with torch.no_grad():
        assert not m(x.cuda()).requires_grad            # back on the inference path
