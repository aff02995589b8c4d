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


def test_hook_on_final_norm_and_hooks_without_a_graph(built_lib):
    """code/visualize_mil.py:227 / gradcam_sus.py target the FINAL LayerNorm (`model.norm`): the hook sees LayerNorm(h3) on every
    row and a gradient that lives in the cls row only (TransMIL.py:202-204).  Without a graph (no_grad, or the eval-mode
    `return_attn=True` call of visualize_mil.py:580 made while the GradCAM object's hooks are alive) the hooks still fire, with the
    activation only -- as in the reference."""
    D, nc, n, B = 512, 3, 300, 1
    m = build_model(nc, D).cuda()
    x = make_input("randn", B, n, D, 84)
    dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(20))
    seen = {}

    def hook(mod, inp, out):
        seen["module"], seen["input"], seen["act"] = mod, inp[0].detach().clone(), out.detach().clone()
        if out.requires_grad:                      # pytorch_grad_cam's save_gradient guards the same way
            out.register_hook(lambda gr: seen.__setitem__("grad", gr.detach().clone()))

    handle = m.norm.register_forward_hook(hook)
    logits = m(x.cuda())
    assert logits.requires_grad and seen["module"] is m.norm
    (logits * dlogits.cuda()).sum().backward()
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    sd64 = {k: v.double() for k, v in sd.items()}
    ref_logits, st = O.forward(sd64, x.double(), return_stages=True, dtype=torch.float64)
    hn = F.layer_norm(st["h3"], (512,), sd64["norm.weight"], sd64["norm.bias"], 1e-5)
    assert (logits.detach().cpu().double() - ref_logits).abs().max().item() <= LOGIT_TOL
    _close(seen["input"], st["h3"], 2.5e-3, "final-norm hook input (h3)")
    _close(seen["act"], hn, 2.5e-3, "final-norm hook output")
    gref = torch.zeros_like(hn)
    gref[:, 0] = dlogits.double() @ sd64["_fc.weight"]
    _close(seen["grad"], gref, 1e-2, "gradient of the final-norm activation")
    assert seen["grad"][:, 1:].abs().max().item() == 0.0
    ref, _ = OB.backward(sd, x, dlogits.double())
    for k, p in m.named_parameters():
        _close(p.grad, ref[k].reshape(p.shape), 1e-2, k)
    # no graph: activation only
    seen.clear()
    with torch.no_grad():
        out = m(x.cuda())
    assert not out.requires_grad and "grad" not in seen
    _close(seen["act"], hn, 2.5e-3, "final-norm hook output under no_grad")
    assert (out.cpu().double() - ref_logits).abs().max().item() <= LOGIT_TOL
    # the visualize_mil.py flow: hooks alive, autograd on, eval(), return_attn=True
    seen.clear()
    logits2, (attn, padding) = m(x.cuda(), return_attn=True)
    assert "act" in seen and "grad" not in seen and padding == O.plan(n)["padding"]
    assert torch.equal(logits2, out)
    row = attn[0, :, padding + 1, padding + 1:padding + 1 + 64]
    assert tuple(row.shape) == (8, 64) and torch.isfinite(row).all()
    handle.remove()
    # a layer hook without a graph: LayerNorm of the layer input
    handle = m.layer1.norm.register_forward_hook(hook)
    seen.clear()
    with torch.no_grad():
        m(x.cuda())
    xn1 = F.layer_norm(st["h0"], (512,), sd64["layer1.norm.weight"], sd64["layer1.norm.bias"], 1e-5)
    assert seen["module"] is m.layer1.norm
    _close(seen["act"], xn1, 2.5e-3, "layer1.norm hook output under no_grad")
    handle.remove()
    seen.clear()
    with torch.no_grad():
        m(x.cuda())
    assert not seen
