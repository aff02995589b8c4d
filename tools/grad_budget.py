"""MEASUREMENT TOOL: max-abs error of every parameter gradient relative to the tensor's largest entry, library backward vs the
float64 backward oracle, per engine (fp32 CUDA-core kernels / tcgen05 tf32 contractions on the tensor-core forward) on the bags
of tests/test_zz_backward_gpu.py.  The numbers behind GRAD_RTOL of that test.
    python tools/grad_budget.py > profiles/r2/grad_budget.txt"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import transmil_b200 as T  # noqa: E402
from common import build_model, make_input  # noqa: E402
from oracle import transmil_backward as OB  # noqa: E402

CASES = [(512, 3, 300, 2), (2048, 2, 1000, 2), (1024, 2, 2500, 1)]


def main():
    L = T.capi.lib()
    for engine, code in (("simt", 1), ("tc", 0)):
        prev, prev_fwd = L.tm_bgemm_engine(code), T.training.FORWARD_ENGINE
        T.training.FORWARD_ENGINE = engine
        worst = {}
        try:
            for D, nc, n, B in CASES:
                m = build_model(nc, D).cuda()
                m.keep_graph = True
                x = make_input("randn", B, n, D, 77)
                dlogits = torch.randn(B, nc, generator=torch.Generator().manual_seed(12))
                logits = m(x.cuda())
                sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
                logits.backward(dlogits.cuda())
                ref, _ = OB.backward(sd, x, dlogits.double())
                line = []
                for k, p in m.named_parameters():
                    r = ref[k].reshape(p.shape).double()
                    e = (p.grad.double().cpu() - r).abs().max().item() / max(r.abs().max().item(), 1e-30)
                    worst[k] = max(worst.get(k, 0.0), e)
                    line.append(e)
                print(f"{engine} D={D} nc={nc} n={n} B={B}: worst parameter {max(line):.2e}, median {sorted(line)[len(line) // 2]:.2e}")
        finally:
            L.tm_bgemm_engine(prev)
            T.training.FORWARD_ENGINE = prev_fwd
        top = sorted(worst.items(), key=lambda kv: -kv[1])
        print(f"{engine} WORST over the cases: " + "  ".join(f"{k} {v:.2e}" for k, v in top[:8]))
        print(f"{engine} all: " + "  ".join(f"{k} {v:.1e}" for k, v in worst.items()))


if __name__ == "__main__":
    main()
