"""MEASUREMENT TOOL: time of tm_pinv_fwd (attn2 + prep + cluster kernel) at B = 1 / 8 / 64 and its error against float64, for the
cluster kernel of csrc/pinv_pipe.cu.  Run under `timeout`."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import transmil_b200 as T  # noqa: E402
from test_pinv_gpu import _pinv64  # noqa: E402

L = T.capi.lib()
st = torch.cuda.current_stream().cuda_stream
for B in (1, 8, 64):
    g = torch.Generator().manual_seed(B)
    q = (torch.randn(B, 8, 256, 64, generator=g) * 0.3).cuda()
    k = (torch.randn(B, 8, 256, 64, generator=g) * 0.3).cuda()
    attn2 = torch.empty(B, 8, 256, 256, device="cuda")
    Z = torch.empty_like(attn2)
    sc = torch.empty(int(L.tm_pinv_scratch_floats(B)), device="cuda")
    for _ in range(5):
        T.capi.check(L.tm_pinv_fwd(q.data_ptr(), k.data_ptr(), B, attn2.data_ptr(), Z.data_ptr(), sc.data_ptr(), 0, st))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        T.capi.check(L.tm_pinv_fwd(q.data_ptr(), k.data_ptr(), B, attn2.data_ptr(), Z.data_ptr(), sc.data_ptr(), 0, st))
    e1.record()
    torch.cuda.synchronize()
    ref = _pinv64(attn2)
    err = (Z.double() - ref).abs().max().item() / ref.abs().max().item()
    print(f"B={B}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us per tm_pinv_fwd, max|Z - Z64| / max|Z64| = {err:.2e}", flush=True)
