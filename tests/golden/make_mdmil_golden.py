"""Generates tests/golden/mdmil_golden.json.

Run in the BUILD container (needs /root/reference):  python tests/golden/make_mdmil_golden.py

Every case executes the UNMODIFIED reference module /root/reference/code/models/MDMIL.py (class MDMIL, :60-113; loaded by
path with oracle/ on sys.path so its `nystrom_attention` import resolves to the restated package) in eval mode, strict fp32,
on a seeded synthetic bag, and records logits, Y_prob and a probe row of the dense attn2 it returns.  The oracle restatement
(oracle/transmil_oracle.py, with MDMIL's `_fc2` classifier mapped to `_fc`) is asserted within 2e-6 on every case: it is
bit-identical at B=1; at B>1 MDMIL's PPEG feeds its convolutions a non-contiguous view (MDMIL.py:40 has no .contiguous(),
TransMIL.py:68 has one) and torch takes another conv path there (3e-7 on the B=2 case).
Weights are re-created from `weight_seed` (same sub-modules in the same order); parameter checksums are stored.
"""
import importlib.util
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader, transmil_oracle as O  # noqa: E402

REF_FILE = "/root/reference/code/models/MDMIL.py"
WEIGHT_SEED = 2021
ORACLE_TOL = 2e-6
#        classes n     B  gen
CASES = [(2, 1, 1, "randn"), (2, 257, 2, "randn"), (6, 1000, 1, "rand"), (3, 3000, 1, "rand"), (2, 6000, 1, "randn")]


def load_reference_mdmil():
    if os.path.dirname(ref_loader.__file__) not in sys.path:
        sys.path.insert(0, os.path.dirname(ref_loader.__file__))
    spec = importlib.util.spec_from_file_location("_reference_MDMIL", REF_FILE)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def oracle_state(sd):
    """MDMIL state_dict -> the TransMIL key names the oracle reads."""
    return {k.replace("_fc2.", "_fc."): v for k, v in sd.items()}


def checksum(t):
    t = t.double()
    return [float(t.sum()), float(t.abs().sum())]


def main():
    torch.set_float32_matmul_precision("highest")
    mod = load_reference_mdmil()
    out = {"weight_seed": WEIGHT_SEED, "cases": []}
    for i, (nc, n, B, gen) in enumerate(CASES):
        torch.manual_seed(WEIGHT_SEED)
        model = mod.MDMIL(n_classes=nc).eval()
        g = torch.Generator().manual_seed(3000 + i)
        x = (torch.rand if gen == "rand" else torch.randn)(B, n, 1024, generator=g)
        with ref_loader._NoCuda(), torch.no_grad():
            logits, attn2 = model(x)
        sd = model.state_dict()
        err = float((O.forward(oracle_state(sd), x) - logits).abs().max())
        print("oracle vs reference MDMIL max abs", err)
        assert err <= ORACLE_TOL, "oracle restatement drifted from the reference MDMIL"
        p = O.plan(n)
        pad = p["padding"]
        assert tuple(attn2.shape) == (B, 8, p["Sp"], p["Sp"])
        probe = attn2[0, :, pad + 1, pad + 1:pad + 1 + min(n, 64)].double()
        out["cases"].append({
            "n_classes": nc, "n": n, "B": B, "gen": gen, "input_seed": 3000 + i, "plan": p,
            "logits": [[float(v) for v in r] for r in logits.double()],
            "y_prob": [[float(v) for v in r] for r in torch.softmax(logits, 1).double()],
            "attn_probe": [[float(v) for v in r] for r in probe],
            "param_checksums": {k: checksum(v) for k, v in sd.items()},
        })
        print(i, nc, n, B, gen, out["cases"][-1]["logits"][0])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mdmil_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
