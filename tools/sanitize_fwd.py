"""Small forwards for compute-sanitizer (memcheck / racecheck): fp32 and fp16 bags, ragged sizes, two back-to-back forwards
each so that programmatic launches cross the forward boundary.  Prints the logits; exits non-zero on a CUDA error."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import build_model, make_input  # noqa: E402

for D, n, B, half in ((1024, 3000, 1, False), (1024, 777, 1, True), (2048, 300, 2, False)):
    m = build_model(2, D).cuda()
    x = make_input("rand", B, n, D, 3).cuda()
    if half:
        x = x.half()
    with torch.no_grad():
        for _ in range(2):
            y = m(x)
    torch.cuda.synchronize()
    print(D, n, B, half, y.cpu().tolist(), flush=True)
print("sanitize_fwd done")
