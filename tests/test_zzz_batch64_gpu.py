"""GPU suite: the reference's default training batch shape, 64 bags x 1000 patches x 2048 features
(code/datasets/feature_dataloader.py:27,388-401), eval forward of the whole batch in one call against the golden vectors of the
unmodified reference module (dense attention and all; tests/golden/make_golden.py --append).  The pinv init scalar is a maximum
over ALL 512 (bag, head) matrices of the batch (nystrom-attention 0.0.11), so this is also the widest batch-coupling case.
Sorted after the established suites: the case was added when no device was at hand."""
import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, build_model, load_golden, make_input

pytestmark = pytest.mark.gpu
CASES = [c for c in load_golden()["cases"] if c["B"] == 64]


@pytest.mark.parametrize("engine", ["default", "simt"])
def test_batch_of_64_bags_matches_reference_golden(built_lib, engine):
    import transmil_b200 as T
    assert len(CASES) == 1
    case = CASES[0]
    m = build_model(case["n_classes"], case["in_features"]).cuda()
    m.flags = T.capi.TM_FLAG_SIMT_ONLY if engine == "simt" else 0
    x = make_input(case["gen"], case["B"], case["n"], case["in_features"], case["input_seed"]).cuda()
    with torch.no_grad():
        logits = m(x).cpu().double()
    ref = torch.tensor(case["logits"], dtype=torch.float64)
    assert logits.shape == ref.shape == (64, 2)
    assert (logits - ref).abs().max().item() <= LOGIT_TOL
    assert (logits.softmax(1) - torch.tensor(case["y_prob"], dtype=torch.float64)).abs().max().item() <= PROB_TOL
