import torch, sys
sys.path.insert(0, '/root/repo')
import transmil_b200 as T
L = T.capi.lib()
B, Sp = 1, 50432
g = torch.Generator().manual_seed(0)
qlm = (torch.randn(B, 8, 256, 64, generator=g) * 0.3).half().cuda()
k = (torch.randn(B, 8, Sp, 64, generator=g) * 0.3).half().cuda()
v = torch.randn(B, 8, Sp, 64, generator=g).half().cuda()
W3 = torch.empty(B, 8, 256, 64, device="cuda")
scratch = torch.empty(int(L.tm_attn3v_tc_scratch_floats(B, Sp)), device="cuda")
st = torch.cuda.current_stream().cuda_stream
for _ in range(2):
    T.capi.check(L.tm_attn3v_tc_fwd(qlm.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, W3.data_ptr(), scratch.data_ptr(), 0, st))
    torch.cuda.synchronize()
