"""Optimizer step and gradient exchange of the training path, CPU suite.

The fused RAdam / Lookahead kernels (csrc/backward.cu) run through the host harness of tests/cuda_emu and are compared with
torch.optim.RAdam -- the optimizer the reference's factory builds (code/MyOptimizer/optim_factory.py:78-80) -- and with the
Lookahead update rule of code/MyOptimizer/lookahead.py:28-52; the one-bucket gradient allreduce runs over gloo, world_size 2.
"""
import os
import socket
from types import SimpleNamespace

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from common import build_model
from test_backward_emu import K, T  # noqa: F401  (fixtures: the host build of the kernel source)


def _twin_models():
    a, b = build_model(2, 1024, seed=3), build_model(2, 1024, seed=3)
    a.train(); b.train()
    return a, b


def _fake_grads(models, step):
    g = torch.Generator().manual_seed(100 + step)
    grads = [torch.randn(p.shape, generator=g) * 0.1 for p in models[0].parameters()]
    for m in models:
        for p, gr in zip(m.parameters(), grads):
            p.grad = gr.clone()


def test_factory_groups_follow_the_reference(K):
    from transmil_b200 import optim
    m = build_model(2, 2048)
    opt = optim.create_optimizer(SimpleNamespace(opt="radam", lr=1e-4, weight_decay=0.01), m, lib=K.L)
    no_decay, decay = opt.param_groups
    assert no_decay["weight_decay"] == 0.0 and decay["weight_decay"] == 0.01
    names = {id(p): n for n, p in m.named_parameters()}
    assert all(names[id(p)].endswith(".bias") or p.dim() == 1 for p in no_decay["params"])
    assert all(p.dim() > 1 and not names[id(p)].endswith(".bias") for p in decay["params"])
    assert len(no_decay["params"]) + len(decay["params"]) == len(list(m.parameters())) == 29
    assert isinstance(optim.create_optimizer(SimpleNamespace(opt="lookahead_radam", lr=1e-4, weight_decay=0.01), m, lib=K.L), optim.FusedLookahead)
    with pytest.raises(ValueError):
        optim.create_optimizer(SimpleNamespace(opt="adamw", lr=1e-4, weight_decay=0.01), m, lib=K.L)


def test_fused_radam_matches_torch_radam(K):
    """12 steps: the unrectified phase (rho_t <= 5, steps 1-5) and the rectified one, both weight-decay groups."""
    from transmil_b200 import optim
    a, b = _twin_models()
    ours = optim.create_optimizer(SimpleNamespace(opt="radam", lr=1e-2, weight_decay=0.01), a, lib=K.L)
    ref = torch.optim.RAdam(optim.add_weight_decay(b, 0.01), lr=1e-2, weight_decay=0.0)
    for step in range(12):
        _fake_grads((a, b), step)
        ours.step()
        ref.step()
        for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
            err = (pa - pb).abs().max().item()
            assert err <= 2e-6 * max(1.0, pb.abs().max().item()), f"step {step + 1} {n}: {err:.3e}"


def test_lookahead_rule(K):
    from transmil_b200 import optim
    a, b = _twin_models()
    ours = optim.create_optimizer(SimpleNamespace(opt="lookahead_radam", lr=1e-2, weight_decay=0.01), a, lib=K.L)
    ref = torch.optim.RAdam(optim.add_weight_decay(b, 0.01), lr=1e-2, weight_decay=0.0)
    slow, alpha, k = {}, 0.5, 6
    for step in range(13):
        _fake_grads((a, b), step)
        ours.step()
        ref.step()
        if (step + 1) % k == 0:           # lookahead.py:28-41, 47-52
            with torch.no_grad():
                for p in b.parameters():
                    if id(p) not in slow:
                        slow[id(p)] = p.data.clone()
                    slow[id(p)].add_(p.data - slow[id(p)], alpha=alpha)
                    p.data.copy_(slow[id(p)])
        for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
            assert (pa - pb).abs().max().item() <= 2e-6 * max(1.0, pb.abs().max().item()), f"step {step + 1} {n}"


def test_state_dict_round_trip_resumes_mid_training(K):
    """save after 8 steps (past one Lookahead sync), load into a fresh optimizer, continue: identical to an
    uninterrupted run -- the per-parameter `step` (bias correction / rectification) and the slow buffers survive."""
    from transmil_b200 import optim
    import copy
    args = SimpleNamespace(opt="lookahead_radam", lr=1e-2, weight_decay=0.01)
    a, b = _twin_models()
    oa, ob = optim.create_optimizer(args, a, lib=K.L), optim.create_optimizer(args, b, lib=K.L)
    for step in range(8):
        _fake_grads((a, b), step)
        oa.step(); ob.step()
    sd = copy.deepcopy(ob.state_dict())
    assert set(sd) == {"state", "slow_state", "param_groups"} and len(sd["slow_state"]) == len(list(b.parameters()))
    assert all(float(st["step"]) == 8.0 for st in sd["state"].values())
    assert sd["param_groups"][0]["lookahead_step"] == 8
    c = build_model(2, 1024, seed=4); c.train()
    c.load_state_dict(b.state_dict())
    oc = optim.create_optimizer(args, c, lib=K.L)
    oc.load_state_dict(sd)
    for step in range(8, 14):
        _fake_grads((a, c), step)
        oa.step(); oc.step()
        for (n, pa), pc in zip(a.named_parameters(), c.parameters()):
            assert torch.equal(pa, pc), f"step {step + 1} {n}"


def test_torch_radam_checkpoint_loads_and_schedulers_accept_the_wrapper(K):
    from transmil_b200 import optim
    a, b = _twin_models()
    ours = optim.create_optimizer(SimpleNamespace(opt="radam", lr=1e-2, weight_decay=0.01), a, lib=K.L)
    ref = torch.optim.RAdam(optim.add_weight_decay(b, 0.01), lr=1e-2, weight_decay=0.0)
    for step in range(7):
        _fake_grads((b,), step)
        ref.step()
    a.load_state_dict(b.state_dict())
    import copy
    ours.load_state_dict(copy.deepcopy(ref.state_dict()))   # a checkpoint of the optimizer the reference builds (a copy, as torch.load
                                                            # gives: state_dict() itself hands out the live `step` tensors)
    for step in range(7, 10):
        _fake_grads((a, b), step)
        ours.step(); ref.step()
        for (n, pa), pb in zip(a.named_parameters(), b.parameters()):
            assert (pa - pb).abs().max().item() <= 2e-6 * max(1.0, pb.abs().max().item()), f"step {step + 1} {n}"
    la = optim.create_optimizer(SimpleNamespace(opt="lookahead_radam", lr=1e-2, weight_decay=0.01), a, lib=K.L)
    assert isinstance(la, torch.optim.Optimizer)
    sched = torch.optim.lr_scheduler.ReduceLROnPlateau(la, mode="min", factor=0.5, patience=0)     # model_interface.py:873
    sched.step(1.0); sched.step(2.0)
    assert all(abs(g["lr"] - 5e-3) < 1e-12 for g in la.base.param_groups)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from transmil_b200 import optim
        m = build_model(2, 1024, seed=3)
        g = torch.Generator().manual_seed(50 + rank)
        for p in m.parameters():
            p.grad = torch.randn(p.shape, generator=g)
        n = optim.allreduce_gradients(m.parameters())
        out[rank] = (n, [p.grad.clone() for p in m.parameters()])
    finally:
        dist.destroy_process_group()


def test_one_bucket_gradient_allreduce_gloo_world2():
    world = 2
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
        res = dict(out)
    m = build_model(2, 1024, seed=3)
    # every rank drew all its tensors from ONE generator: rebuild rank by rank
    per_rank = []
    for r in range(world):
        g = torch.Generator().manual_seed(50 + r)
        per_rank.append([torch.randn(p.shape, generator=g) for p in m.parameters()])
    expect = [sum(t) / world for t in zip(*per_rank)]
    for r in range(world):
        n, grads = res[r]
        assert n == sum(p.numel() for p in m.parameters())
        for a, b in zip(grads, expect):
            assert torch.allclose(a, b, atol=1e-6)
    assert all(torch.equal(a, b) for a, b in zip(res[0][1], res[1][1]))
