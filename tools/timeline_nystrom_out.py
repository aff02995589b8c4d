import torch, sys
sys.path.insert(0, '/root/repo')
import transmil_b200 as T
L = T.capi.lib()
B, Sp = 1, 50432
g = torch.Generator().manual_seed(0)
q = (torch.randn(B, 8, Sp, 64, generator=g) * 0.3).half().cuda()
klm = (torch.randn(B, 8, 256, 64, generator=g) * 0.3).half().cuda()
mt16 = torch.randn(B, 8, 64, 256, generator=g).half().cuda()
v = torch.randn(B, 8, Sp, 64, generator=g).half().cuda()
w = (torch.randn(8, 1, 33, 1, generator=g) * 0.2).cuda()
merged = torch.empty(B, Sp, 512, dtype=torch.float16, device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    T.capi.check(L.tm_nystrom_out_tc_fwd(q.data_ptr(), klm.data_ptr(), mt16.data_ptr(), v.data_ptr(), w.data_ptr(), B, Sp, merged.data_ptr(), T.capi.TM_FLAG_F16_ACT, st))
    torch.cuda.synchronize()
