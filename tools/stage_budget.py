"""MEASUREMENT TOOL: per-stage error of the tensor-core engine (and the fp32 CUDA-core engine) against the CPU oracle on the bags of
tests/test_parity_gpu.py::test_every_stage_against_oracle -- the numbers the per-stage budgets of that test are taken from.
    python tools/stage_budget.py > profiles/r2/stage_budget.txt"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import transmil_b200 as T  # noqa: E402
from common import build_model, make_input  # noqa: E402
from oracle import transmil_oracle as O  # noqa: E402

CASES = [(2048, 2, 1000, 2, "rand"), (1024, 3, 777, 1, "randn"), (768, 2, 300, 1, "randn"), (512, 2, 5, 3, "randn"), (384, 6, 2500, 1, "rand"),
         (1024, 2, 6000, 1, "rand"), (2048, 2, 20000, 1, "rand")]


def err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return (a - b).abs().max().item(), b.abs().max().item()


def main():
    for engine in ("tc", "simt"):
        worst = {}
        for D, nc, n, B, gen in CASES:
            m = build_model(nc, D).cuda()
            m.flags = T.capi.TM_FLAG_KEEP_STAGES | (T.capi.TM_FLAG_SIMT_ONLY if engine == "simt" else 0)
            x = make_input(gen, B, n, D, 42)
            with torch.no_grad():
                logits = m(x.cuda())
            sd = {k: v.cpu() for k, v in m.state_dict().items()}
            ref, st = O.forward(sd, x, return_stages=True)
            l2 = st["layer2"]
            pairs = [("h0", m.stage("h0"), st["h0"]), ("q_lm1", m.stage("q_lm1"), st["layer1"]["q_lm"]), ("k_lm1", m.stage("k_lm1"), st["layer1"]["k_lm"]),
                     ("attn21", m.stage("attn21"), st["layer1"]["attn2"]), ("Z1", m.stage("Z1"), st["layer1"]["Z"]), ("W31", m.stage("W31"), st["layer1"]["W3"]),
                     ("h1", m.stage("h1"), st["h1"]), ("h2", m.stage("h2"), st["h2"]), ("xn", m.stage("xn")[:, l2["P"]:], l2["xn"]),
                     ("q", m.stage("q"), l2["q"]), ("k", m.stage("k"), l2["k"]), ("v", m.stage("v"), l2["v"]), ("merged", m.stage("merged"), l2["merged"]),
                     ("Z", m.stage("Z"), l2["Z"]), ("h3cls", m.stage("h3")[:, 0], st["h3"][:, 0]), ("logits", logits, ref)]
            line = []
            for name, got, want in pairs:
                e, mx = err(got, want)
                line.append(f"{name} {e:.2e}/{mx:.2e}")
                w = worst.setdefault(name, [0.0, 0.0])
                w[0] = max(w[0], e); w[1] = max(w[1], e / max(mx, 1e-30))
            print(f"{engine} D={D} nc={nc} n={n} B={B} {gen}: " + "  ".join(line))
        print(f"{engine} WORST abs / rel-to-max: " + "  ".join(f"{k} {v[0]:.2e}/{v[1]:.2e}" for k, v in worst.items()))


if __name__ == "__main__":
    main()
