import sys, torch
sys.path.insert(0, '/root/repo')
import transmil_b200 as T
L = T.capi.lib()
st = torch.cuda.current_stream().cuda_stream
B, S, P = 1, 50177, 255
Sp = S + P
merged = torch.randn(B, Sp, 512).half().cuda(); w = (torch.randn(512, 512) * 0.04).half().cuda(); b = torch.randn(512).cuda()
h = torch.randn(B, S, 512).cuda(); ho = torch.empty_like(h)
xn = torch.randn(B, Sp, 512).half().cuda(); wq = (torch.randn(1536, 512) * 0.04).half().cuda()
q = torch.empty(B, 8, Sp, 64, dtype=torch.float16, device="cuda"); k = torch.empty_like(q); v = torch.empty_like(q)
def run(fn, name):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    print(name, "%.1f us" % (e0.elapsed_time(e1) * 50))
run(lambda: T.capi.check(L.tm_to_out_residual_fwd(merged.data_ptr(), w.data_ptr(), b.data_ptr(), h.data_ptr(), B, S, P, ho.data_ptr(), T.capi.TM_FLAG_F16_ACT, st)), "to_out")
run(lambda: T.capi.check(L.tm_qkv16_fwd(xn.data_ptr(), wq.data_ptr(), B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), T.capi.TM_FLAG_F16_ACT, st)), "qkv16")
