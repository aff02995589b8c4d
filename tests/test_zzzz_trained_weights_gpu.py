"""GPU suite: the north_star gates (logits 1e-3, Y_prob 1e-4) on weights that have been TRAINED, not only on the default init
(trained checkpoints are sharper than the seeded init).  No trained checkpoint can travel to the box, so the test
makes one: a few dozen RAdam steps of the library's own training path on separable synthetic bags (loss drops, the attention
rows sharpen), then the eval forward of the resulting state_dict -- loaded through load_reference_checkpoint from a Lightning
style file ('model.' prefix) -- against the CPU oracle on held-out bags, for both engines."""
import os
from types import SimpleNamespace

import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, make_input
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("D,n", [(1024, 1500), (2048, 1000)])
def test_trained_weights_stay_inside_the_gates(built_lib, tmp_path, D, n):
    import transmil_b200 as T
    from transmil_b200 import optim, plugin
    torch.manual_seed(2021)
    m = T.TransMIL(n_classes=2, in_features=D).cuda().train()
    opt = optim.create_optimizer(SimpleNamespace(opt="lookahead_radam", lr=1e-3, weight_decay=0.01), m)
    g = torch.Generator().manual_seed(3)
    B = 8
    y = torch.arange(B) % 2
    # class 1 bags carry a shifted block of patches: learnable, so the weights really move
    x = torch.rand(B, n, D, generator=g)
    x[y == 1, : n // 4, : D // 8] += 0.75
    x, yc = x.cuda(), y.cuda()
    losses = []
    for _ in range(40):
        opt.zero_grad(set_to_none=True)
        loss, _, _ = plugin.training_step(m, (x, yc, None))
        loss.sum().backward()
        opt.step()
        losses.append(float(loss.detach().sum()))
    assert losses[-1] < 0.8 * losses[0], losses[::8]
    # checkpoint round trip the way the reference's consumers do it (Lightning state_dict, 'model.' prefix)
    ckpt = os.path.join(tmp_path, "trained.ckpt")
    torch.save({"state_dict": {"model." + k: v.detach().cpu() for k, v in m.state_dict().items()}}, ckpt)
    torch.manual_seed(1)
    fresh = T.TransMIL(n_classes=2, in_features=D).cuda().eval()
    plugin.load_reference_checkpoint(fresh, ckpt)
    sd = {k: v.detach().cpu() for k, v in fresh.state_dict().items()}
    moved = max((sd[k] - v.cpu()).abs().max().item() for k, v in T.TransMIL(2, D).state_dict().items() if k in sd)
    assert moved > 1e-2
    for seed in (11, 12):
        xe = make_input("rand", 1, n + 37 * seed, D, seed)
        if seed == 12:
            xe[:, : n // 4, : D // 8] += 0.75
        ref = O.forward(sd, xe)
        for engine in ("fp32", "tensor-core"):
            fresh.engine = engine
            with torch.no_grad():
                got = fresh(xe.cuda()).cpu()
            dl = (got - ref).abs().max().item()
            dp = (got.softmax(1) - ref.softmax(1)).abs().max().item()
            print(f"trained weights D={D} seed={seed} {engine}: dlogit {dl:.2e} dprob {dp:.2e} logits {ref.tolist()}")
            assert dl <= LOGIT_TOL and dp <= PROB_TOL, (engine, dl, dp)
