"""Throughput sweep over the BASELINE.json bag sizes on one GPU (device-resident bags, eval forward):
n in {6k, 20k, 50k, 100k, 200k} x D in {1024, 2048}, plus the reference's default training-shaped batch 64 x 1000 x 2048.
Prints one JSON line per case: ms per step, slides/s, tokens/s, algorithmic TFLOP/s (minimal association)
and its fraction of the measured sustained bf16 peak."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import transmil_b200 as T  # noqa: E402


def flops(B, n, D):
    p = T.capi.make_plan(n)
    S, Sp, N = int(p.S), int(p.Sp), int(p.N)
    f_fc1 = 2.0 * n * (2048 * 1024 + 1024 * 512) if D == 2048 else 2.0 * n * D * 512
    f_layer = 2.0 * Sp * 512 * 1536 + 4 * (2.0 * Sp * 256 * 512) + 2.0 * S * 512 * 512 + 2.0 * 33 * Sp * 512 + 8 * (2 * 2.0 * 256 * 256 * 64 + 6 * 4 * 2.0 * 256 ** 3)
    return B * (f_fc1 + 2 * f_layer + 2.0 * 83 * N * 512)


def main():
    peak = 1354.0
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops_sustained"]
    except Exception:
        pass
    cases = [(1, n, D) for D in (1024, 2048) for n in (6000, 20000, 50000, 100000, 200000)] + [(64, 1000, 2048), (1, 1000, 2048)]
    for B, n, D in cases:
        torch.manual_seed(2021)
        m = T.TransMIL(n_classes=2, in_features=D).eval().cuda()
        x = torch.rand(B, n, D, generator=torch.Generator().manual_seed(1)).cuda()
        steps = 20 if n <= 50000 else 10
        with torch.no_grad():
            for _ in range(3):
                m(x)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                m(x)
            e1.record()
            torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        tf = flops(B, n, D) / (ms * 1e-3) / 1e12
        print(json.dumps({"B": B, "n": n, "D": D, "ms_per_step": round(ms, 4), "slides_per_s": round(B / ms * 1e3, 1),
                          "tokens_per_s": round(B * n / ms * 1e3), "alg_tflops": round(tf, 1), "frac_of_measured_sustained_bf16": round(tf / peak, 3)}), flush=True)
        del m, x
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
