"""Generates tests/golden/ctmil_golden.json.

Run in the BUILD container (needs /root/reference):  python tests/golden/make_ctmil_golden.py

Every case executes the UNMODIFIED reference module /root/reference/code/models/CTMIL.py (class CTMIL, :76-165; loaded by path
as a member of a synthetic package so that its `from ._transformer import ...` resolves to the reference's own file, with
oracle/ on sys.path so its `nystrom_attention` import resolves to the restated package) in eval mode, strict fp32, on seeded
synthetic image tiles [1,N,C,H,H], and records the logits [N, n_classes] (one row per tile), Y_prob, a checksum and a probe
of the tokens the conv stem hands to the Nystrom path, and the parameter checksums of the seeded init.  The oracle restatement
(oracle/transmil_oracle.py ctmil_forward) is asserted within 2e-6 on every case.  The commented-out Nystrom forward of
TransformerMIL.py:155-183 is the same code on TransformerMIL's parameters: the generator checks that a TransformerMIL
carrying the CTMIL case's weights gives the same logits through the oracle.
`.cuda()` at CTMIL.py:150 is neutralised on this CPU-only host for the duration of the call (ref_loader._NoCuda).
"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_loader, transmil_oracle as O  # noqa: E402

WEIGHT_SEED = 2021
ORACLE_TOL = 2e-6
#        classes in_features N  H   gen
CASES = [(2, 768, 3, 20, "randn"),     # 20 -> 10 -> 5: 25 tokens per tile
         (3, 64, 5, 50, "rand"),       # the reference's own tile size (CTMIL.py:170): 50 -> 25 -> 13: 169 tokens, no _fc1 built
         (6, 1024, 2, 30, "randn"),    # 30 -> 15 -> 8: 64 tokens
         (2, 128, 20, 34, "randn")]    # 34 -> 17 -> 9: 81 tokens, 20 tiles = 20 bags coupled through the pinv scalar


def checksum(t):
    t = t.double()
    return [float(t.sum()), float(t.abs().sum())]


def main():
    torch.set_float32_matmul_precision("highest")
    mod = ref_loader.load_reference_sibling("CTMIL")
    out = {"weight_seed": WEIGHT_SEED, "cases": []}
    for i, (nc, D, N, H, gen) in enumerate(CASES):
        torch.manual_seed(WEIGHT_SEED)
        model = mod.CTMIL(n_classes=nc, in_features=D).eval()
        g = torch.Generator().manual_seed(4000 + i)
        x = (torch.rand if gen == "rand" else torch.randn)(1, N, D, H, H, generator=g)
        with ref_loader._NoCuda(), torch.no_grad():
            logits = model(x)
        sd = model.state_dict()
        mine = O.ctmil_forward(sd, x)
        err = float((mine - logits).abs().max())
        print("oracle vs reference CTMIL max abs", err)
        assert err <= ORACLE_TOL, "oracle restatement drifted from the reference CTMIL"
        tok = O.ctmil_tokens(sd, x)
        out["cases"].append({
            "n_classes": nc, "in_features": D, "N": N, "H": H, "gen": gen, "input_seed": 4000 + i,
            "n_tokens": int(tok.shape[1]), "plan": O.plan(int(tok.shape[1])),
            "logits": [[float(v) for v in r] for r in logits.double()],
            "y_prob": [[float(v) for v in r] for r in torch.softmax(logits, 1).double()],
            "tokens_checksum": checksum(tok), "tokens_probe": [float(v) for v in tok[0, 0, :16].double()],
            "param_checksums": {k: checksum(v) for k, v in sd.items()},
        })
        print(i, nc, D, N, H, gen, out["cases"][-1]["logits"][0])
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ctmil_golden.json")
    with open(path, "w") as f:
        json.dump(out, f, indent=1)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
