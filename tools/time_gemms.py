"""Times the three big projections of one 50k bag on their own (CUDA events, 20 launches): fc1 (tf32), qkv16, to_out.
With TM_GEMM_DBG=1|2 in the environment the TMA-store epilogues are (partly) switched off: measurement of what bounds them."""
import ctypes as C, os, sys, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from common import build_model, make_input
import transmil_b200 as T
L = T.capi.lib()
st = torch.cuda.current_stream().cuda_stream
B, S, P = 1, 50177, 255
Sp = S + P
merged = torch.randn(B, Sp, 512).half().cuda(); w = (torch.randn(512, 512) * 0.04).half().cuda(); b = torch.randn(512).cuda()
h = torch.randn(B, S, 512).cuda(); ho = torch.empty_like(h)
xn = torch.randn(B, Sp, 512).half().cuda(); wq = (torch.randn(1536, 512) * 0.04).half().cuda()
q = torch.empty(B, 8, Sp, 64, dtype=torch.float16, device="cuda"); k = torch.empty_like(q); v = torch.empty_like(q)
m = build_model(2, 1024).cuda(); wts = m._weights()
x = make_input("rand", 1, 50000, 1024, 1).cuda(); h0 = torch.empty(1, S, 512, device="cuda")
def run(fn, name):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    print(os.environ.get("TM_GEMM_DBG", "0"), name, "%.1f us" % (e0.elapsed_time(e1) * 50), flush=True)
run(lambda: T.capi.check(L.tm_fc1_fwd(C.byref(wts), x.data_ptr(), 1, 50000, h0.data_ptr(), 0, 0, st)), "fc1 tf32")
run(lambda: T.capi.check(L.tm_qkv16_fwd(xn.data_ptr(), wq.data_ptr(), B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), T.capi.TM_FLAG_F16_ACT, st)), "qkv16")
run(lambda: T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), wq.data_ptr(), B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), 0, 1024, 0, None, T.capi.TM_FLAG_F16_ACT, st)), "qk16")
run(lambda: T.capi.check(L.tm_to_out_residual_fwd(merged.data_ptr(), w.data_ptr(), b.data_ptr(), h.data_ptr(), B, S, P, ho.data_ptr(), T.capi.TM_FLAG_F16_ACT, st)), "to_out")
