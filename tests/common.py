import json
import os

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden", "transmil_golden.json")

LOGIT_TOL = 1e-3   # BASELINE.json north_star: logits within 1e-3 max-abs of the fp32 reference
PROB_TOL = 1e-4    # ... and Y_prob within 1e-4


def load_golden():
    with open(GOLDEN) as f:
        return json.load(f)


def make_input(gen, B, n, D, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand if gen == "rand" else torch.randn)(B, n, D, generator=g)


def checksum(t):
    t = t.double()
    return [float(t.sum()), float(t.abs().sum())]


def build_model(n_classes, in_features, seed=2021):
    import transmil_b200 as T
    torch.manual_seed(seed)
    m = T.TransMIL(n_classes=n_classes, in_features=in_features)
    m.eval()
    return m
