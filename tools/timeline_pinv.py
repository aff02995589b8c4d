"""Per-step clock64 timeline of pinv_cluster_kernel (build with TM_EXTRA_NVCC_FLAGS=-DTM_PINV_TIMELINE)."""
import sys, torch
sys.path.insert(0, '/root/repo')
import transmil_b200 as T
L = T.capi.lib()
g = torch.Generator().manual_seed(0)
qlm = (torch.randn(1, 8, 256, 64, generator=g) * 0.3).cuda()
klm = (torch.randn(1, 8, 256, 64, generator=g) * 0.3).cuda()
attn2 = torch.empty(1, 8, 256, 256, device="cuda"); Z = torch.empty_like(attn2)
scratch = torch.empty(int(L.tm_pinv_scratch_floats(1)), device="cuda")
st = torch.cuda.current_stream().cuda_stream
for i in range(3):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    T.capi.check(L.tm_pinv_fwd(qlm.data_ptr(), klm.data_ptr(), 1, attn2.data_ptr(), Z.data_ptr(), scratch.data_ptr(), 0, st))
    e1.record(); torch.cuda.synchronize()
    print("pinv ms", e0.elapsed_time(e1), flush=True)
A = attn2[0, 0].double(); Zd = Z[0, 0].double()
print("||AZA-A||/||A||", float((A @ Zd @ A - A).norm() / A.norm()))
# reference: the same 6 iterations in fp32 torch on the kernel's own attn2
A = attn2[0]
absA = A.abs()
Zr = A.transpose(-1, -2) / (absA.sum(-1).max() * absA.sum(-2).max())
I = torch.eye(256, device="cuda")[None]
for _ in range(6):
    XZ = A @ Zr
    Zr = 0.25 * Zr @ (13 * I - XZ @ (15 * I - XZ @ (7 * I - XZ)))
print("max|Z - Z_ref| / max|Z_ref|", float((Z[0] - Zr).abs().max() / Zr.abs().max()))
