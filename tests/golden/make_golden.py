"""Generates tests/golden/transmil_golden.json.

Run in the BUILD container (needs /root/reference):  python tests/golden/make_golden.py

Every case executes the UNMODIFIED reference module /root/reference/code/models/TransMIL.py
(loaded by oracle/ref_loader.py; its un-vendored dependency nystrom_attention 0.0.11 is
supplied by the restatement in oracle/nystrom_attention) in eval mode, strict fp32, on a
seeded synthetic bag, and records logits, Y_prob (model_interface.py:323-324), the
`padding` it reports, and float64 checksums of per-stage tensors taken from the oracle
restatement (which tests/test_oracle.py proves bit-identical to the reference module).
Cases marked source="oracle" are too large for the reference's dense [8,Sp,Sp] attention
to be comfortable on the build host and come from the (pinned) restatement instead.

Weights are NOT stored: they are re-created from `weight_seed` by constructing the module
under torch.manual_seed (same nn sub-modules in the same order => identical init); the
fixture stores a float64 checksum of every parameter so a drifting init is detected.
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader, transmil_oracle as O  # noqa: E402

WEIGHT_SEED = 2021  # General.seed, Camelyon/TransMIL.yaml:3

#        D     classes n      B  gen      source
CASES = [
    (2048, 2, 1, 1, "randn", "reference"),
    (2048, 2, 2, 1, "randn", "reference"),
    (2048, 2, 3, 1, "randn", "reference"),
    (2048, 2, 4, 1, "randn", "reference"),
    (512, 3, 255, 1, "randn", "reference"),
    (512, 3, 256, 1, "randn", "reference"),
    (512, 3, 257, 1, "rand", "reference"),
    (2048, 2, 1000, 1, "rand", "reference"),
    (2048, 2, 1000, 4, "randn", "reference"),
    (1024, 2, 1023, 1, "randn", "reference"),
    (1024, 2, 1024, 2, "rand", "reference"),
    (768, 6, 1025, 1, "randn", "reference"),
    (384, 6, 3000, 1, "rand", "reference"),
    (1024, 2, 6000, 1, "rand", "reference"),     # BASELINE config 1 (Camelyon, CPU-runnable)
    (2048, 3, 6000, 1, "randn", "reference"),    # TransMIL.py:216 __main__ smoke shape
    (2048, 2, 20000, 1, "rand", "oracle"),       # BASELINE config 2 bag size
    (2048, 2, 1000, 64, "rand", "reference"),    # the reference's default training batch: 64 bags x 1000 patches
]                                                # (feature_dataloader.py:27,388-401); appended with --append


def make_input(gen, B, n, D, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.rand if gen == "rand" else torch.randn)(B, n, D, generator=g)


def checksum(t):
    t = t.double()
    return [float(t.sum()), float(t.abs().sum())]


def main():
    torch.set_num_threads(os.cpu_count())
    torch.set_float32_matmul_precision("highest")
    out = {"weight_seed": WEIGHT_SEED, "cases": []}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "transmil_golden.json")
    if "--append" in sys.argv:      # keep the cases already in the fixture byte for byte, compute only the new tail of CASES
        with open(path) as f:
            out = json.load(f)
        assert out["weight_seed"] == WEIGHT_SEED and len(out["cases"]) <= len(CASES)
    models = {}
    for i, (D, nc, n, B, gen, source) in enumerate(CASES):
        if i < len(out["cases"]):
            c = out["cases"][i]
            assert (c["in_features"], c["n_classes"], c["n"], c["B"], c["gen"], c["source"]) == (D, nc, n, B, gen, source)
            continue
        key = (D, nc)
        if key not in models:
            models[key] = ref_loader.build_reference(nc, D, seed=WEIGHT_SEED)
        model = models[key]
        sd = model.state_dict()
        x = make_input(gen, B, n, D, 1000 + i)
        if source == "reference":
            logits, (attn, padding) = ref_loader.reference_forward(model, x, return_attn=True)
            p = O.plan(n)
            row = padding + 1
            attn_probe = attn[0, :, row, padding + 1:padding + 1 + min(n, 64)].double()  # visualize_mil.py:581 indexing
            attn_probe = [[float(v) for v in r] for r in attn_probe]
            del attn
        else:
            logits = O.forward(sd, x)
            padding = O.plan(n)["padding"]
            attn_probe = None
        lo2, stages = O.forward(sd, x, return_stages=True)
        assert torch.equal(lo2, logits), "oracle restatement drifted from the reference module"
        case = {
            "in_features": D, "n_classes": nc, "n": n, "B": B, "gen": gen, "input_seed": 1000 + i, "source": source,
            "plan": O.plan(n),
            "logits": [[float(v) for v in r] for r in logits.double()],
            "y_prob": [[float(v) for v in r] for r in O.y_prob(logits).double()],
            "padding": int(padding),
            "attn_probe": attn_probe,
            "param_checksums": {k: checksum(v) for k, v in sd.items()},
            "stage_checksums": {
                "h0": checksum(stages["h0"]), "h1": checksum(stages["h1"]), "h2": checksum(stages["h2"]),
                "h3_cls": checksum(stages["h3"][:, 0]),
                "q_lm1": checksum(stages["layer1"]["q_lm"]), "k_lm1": checksum(stages["layer1"]["k_lm"]),
                "attn21": checksum(stages["layer1"]["attn2"]), "Z1": checksum(stages["layer1"]["Z"]),
                "W31": checksum(stages["layer1"]["W3"]),
            },
        }
        out["cases"].append(case)
        print(i, D, nc, n, B, gen, source, case["logits"][0])
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
