"""CPU suite: the host-side mirror keeps the reference's plugin interface
(constructor, names, state_dict layout, loader) and refuses to run without CUDA."""
import pytest
import torch

from common import build_model, checksum, load_golden

EXPECTED_2048 = {
    "cls_token": (1, 1, 512),
    "pos_layer.proj.weight": (512, 1, 7, 7), "pos_layer.proj.bias": (512,),
    "pos_layer.proj1.weight": (512, 1, 5, 5), "pos_layer.proj1.bias": (512,),
    "pos_layer.proj2.weight": (512, 1, 3, 3), "pos_layer.proj2.bias": (512,),
    "_fc1.0.weight": (1024, 2048), "_fc1.0.bias": (1024,), "_fc1.2.weight": (1024,), "_fc1.2.bias": (1024,),
    "_fc1.3.weight": (512, 1024), "_fc1.3.bias": (512,),
    "layer1.norm.weight": (512,), "layer1.norm.bias": (512,), "layer1.attn.to_qkv.weight": (1536, 512),
    "layer1.attn.to_out.0.weight": (512, 512), "layer1.attn.to_out.0.bias": (512,), "layer1.attn.res_conv.weight": (8, 1, 33, 1),
    "layer2.norm.weight": (512,), "layer2.norm.bias": (512,), "layer2.attn.to_qkv.weight": (1536, 512),
    "layer2.attn.to_out.0.weight": (512, 512), "layer2.attn.to_out.0.bias": (512,), "layer2.attn.res_conv.weight": (8, 1, 33, 1),
    "norm.weight": (512,), "norm.bias": (512,), "_fc.weight": (2, 512), "_fc.bias": (2,),
}


def test_state_dict_layout_matches_reference():
    m = build_model(2, 2048)
    sd = m.state_dict()
    assert {k: tuple(v.shape) for k, v in sd.items()} == EXPECTED_2048
    assert sum(v.numel() for v in sd.values()) == 4772370          # SURVEY 2.1 parameter census
    assert m.n_classes == 2
    assert isinstance(m.layer2.norm, torch.nn.LayerNorm)           # GradCAM hook point (train_loop.py:87)


def test_seeded_init_identical_to_reference():
    gold = load_golden()
    seen = set()
    for case in gold["cases"]:
        key = (case["in_features"], case["n_classes"])
        if key in seen:
            continue
        seen.add(key)
        sd = build_model(key[1], key[0], gold["weight_seed"]).state_dict()
        assert set(sd) == set(case["param_checksums"])
        for k, v in case["param_checksums"].items():
            assert checksum(sd[k]) == v, (key, k)


def test_plugin_loader_selects_by_yaml_name():
    from transmil_b200 import plugin
    import transmil_b200 as T
    cfg = {"Model": {"name": "TransMIL", "n_classes": 2, "backbone": "resnet18"}, "Data": {"feature_extractor": "retccl"}}
    m = plugin.load_model(plugin.model_cfg_from_yaml(cfg))           # train.py:392-393 -> in_features 2048
    assert isinstance(m, T.TransMIL) and m.in_features == 2048 and m.n_classes == 2
    with pytest.raises(ValueError, match="Invalid Module File Name or Invalid Class Name!"):
        plugin.load_model({"name": "NoSuchModel"})
    import models                                                    # code/models/__init__.py:3
    assert models.TransMIL is T.TransMIL


def test_forward_argument_contract_and_no_cpu_fallback():
    import transmil_b200 as T
    m = build_model(2, 512)
    with pytest.raises(TypeError):
        m()
    with pytest.raises(ValueError):
        m(torch.zeros(1, 10, 77))
    if not torch.cuda.is_available():
        with pytest.raises(T.capi.TransmilError, match="no CPU path"):
            m(torch.zeros(1, 10, 512))
        with pytest.raises(T.capi.TransmilError):
            m(data=torch.zeros(10, 512))
    m.train()   # training path (library forward + backward behind autograd): CUDA only as well
    with pytest.raises(T.capi.TransmilError, match="no CPU path"):
        m(torch.zeros(1, 10, 512))
    with pytest.raises(NotImplementedError):
        m(torch.zeros(1, 10, 512), return_attn=True)


def test_hooks_on_the_gradcam_layers_are_never_skipped_silently():
    """train_loop.py:87 hooks model.layer2.norm, visualize_mil.py:227 the final model.norm.  Forward hooks are served by the autograd
    path (tests/test_backward_emu.py, tests/test_zzz_gradcam_hooks_gpu.py) or, without a graph, fired with the kept activations;
    other hook kinds make the forward say so instead of skipping them."""
    import transmil_b200 as T
    m = build_model(2, 512)
    for norm, name in ((m.layer2.norm, "layer2"), (m.layer1.norm, "layer1"), (m.norm, "norm")):
        handle = norm.register_forward_hook(lambda mod, inp, out: None)
        assert m._hooked_norms() == [name]
        if not torch.cuda.is_available():
            with pytest.raises(T.capi.TransmilError, match="no CPU path"):      # autograd on: routed to the (CUDA-only) training path
                m(torch.zeros(1, 10, 512))
            with torch.no_grad(), pytest.raises(T.capi.TransmilError, match="no CPU path"):   # no graph: the inference path, hooks fired
                m(torch.zeros(1, 10, 512))
        handle.remove()
    assert m._hooked_norms() == []
    handle = m.layer1.norm.register_forward_pre_hook(lambda mod, inp: None)
    with pytest.raises(NotImplementedError, match="only forward hooks"):
        m(torch.zeros(1, 10, 512))
    handle.remove()
    handle = m.norm.register_full_backward_hook(lambda mod, gi, go: None)
    with pytest.raises(NotImplementedError, match="only forward hooks are supported on norm"):
        m(torch.zeros(1, 10, 512))
    handle.remove()


def test_load_reference_checkpoint_keys_roundtrip():
    """Consumers strip the Lightning 'model.' prefix and load_state_dict into a bare TransMIL
    (code/visualize_mil.py:165-172)."""
    src = build_model(3, 1024, seed=1)
    ckpt = {"state_dict": {"model." + k: v.clone() for k, v in src.state_dict().items()}}
    dst = build_model(3, 1024, seed=2)
    stripped = {k[len("model."):]: v for k, v in ckpt["state_dict"].items()}
    missing, unexpected = dst.load_state_dict(stripped, strict=True)
    assert not missing and not unexpected
    assert all(torch.equal(a, b) for a, b in zip(src.state_dict().values(), dst.state_dict().values()))


def test_reference_checkpoint_loads_with_the_model_prefix(tmp_path):
    """visualize_mil.py:165-172: Lightning checkpoints hold the TransMIL parameters under `model.`; keys are renamed and
    loaded strictly, foreign keys only pass with strict=False."""
    import torch
    from common import build_model
    from transmil_b200 import plugin
    src = build_model(3, 1024, seed=7)
    ckpt = {"state_dict": {"model." + k: v.clone() for k, v in src.state_dict().items()}, "epoch": 3}
    dst = build_model(3, 1024, seed=8)
    assert not torch.equal(dst.state_dict()["_fc.weight"], src.state_dict()["_fc.weight"])
    plugin.load_reference_checkpoint(dst, ckpt)
    for k, v in src.state_dict().items():
        assert torch.equal(dst.state_dict()[k], v), k
    path = tmp_path / "epoch3.ckpt"
    torch.save(ckpt, path)
    dst2 = build_model(3, 1024, seed=9)
    plugin.load_reference_checkpoint(dst2, str(path))
    assert torch.equal(dst2.state_dict()["layer1.attn.to_qkv.weight"], src.state_dict()["layer1.attn.to_qkv.weight"])
    ckpt["state_dict"]["loss.weight"] = torch.ones(3)
    with pytest.raises(RuntimeError):
        plugin.load_reference_checkpoint(build_model(3, 1024, seed=1), ckpt)
    plugin.load_reference_checkpoint(build_model(3, 1024, seed=1), ckpt, strict=False)


def test_validation_step_mirror_follows_the_reference():
    """model_interface.py:432-470 + utils/utils.py:242-247, with a host stand-in for the model call."""
    from transmil_b200 import plugin

    class Stub(torch.nn.Module):
        n_classes = 3

        def forward(self, x):
            assert not torch.is_grad_enabled() and x.dtype == torch.float32 and x.is_contiguous()
            assert not self.training          # Lightning's validation loop: eval() for the duration of the step
            return x.sum(1)[:, :3]

    x = torch.randn(2, 5, 4, dtype=torch.float64)
    label = torch.tensor([2, 0])
    stub = Stub().train()
    out = plugin.validation_step(stub, (x, label, (("slide_a", "slide_b"), None, ("p1", "p2"))))
    assert stub.training              # ... and the mode it came in is restored
    logits = x.float().sum(1)[:, :3]
    assert torch.equal(out["logits"], logits) and torch.equal(out["Y_hat"], logits.argmax(1))
    assert torch.allclose(out["Y_prob"], logits.softmax(1)) and out["label"].dtype == torch.int32
    # the reference's loop form of the loss
    sm = [torch.softmax(logits[i], dim=0) for i in range(2)]
    ref = -torch.sum(torch.tensor([torch.log(sm[i][label[i]]) for i in range(2)])) / 2
    assert torch.allclose(out["loss"], ref) and not out["loss"].requires_grad
    assert out["name"] == ("slide_a", "slide_b") and out["patient"] == ("p1", "p2")
