import sys, ctypes as C, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/tests')
from common import build_model, make_input
import transmil_b200 as T
L = T.capi.lib()
st = torch.cuda.current_stream().cuda_stream
def sync(tag):
    try:
        torch.cuda.synchronize(); print(tag, "ok", flush=True)
    except Exception as e:
        print(tag, "FAILED", str(e)[:200], flush=True); sys.exit(1)
which = sys.argv[1] if len(sys.argv) > 1 else "all"
B, Sp = 1, 3072
if which in ("qkv", "all"):
    xn = torch.randn(B, Sp, 512).half().cuda(); w = (torch.randn(1536, 512) * 0.04).half().cuda()
    q = torch.empty(B, 8, Sp, 64, dtype=torch.float16, device="cuda"); k = torch.empty_like(q); v = torch.empty_like(q)
    T.capi.check(L.tm_qkv16_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), T.capi.TM_FLAG_F16_ACT, st))
    sync("qkv16")
    ref = (xn.float() @ w.float().T).view(B, Sp, 3, 8, 64).permute(2, 0, 3, 1, 4)
    print("qkv err", float((q.float() - ref[0] * 0.125).abs().max()), float((k.float() - ref[1]).abs().max()), float((v.float() - ref[2]).abs().max()))
if which in ("out", "all"):
    S, P = 3000, 72
    merged = torch.randn(B, Sp, 512).half().cuda(); w = (torch.randn(512, 512) * 0.04).half().cuda(); b = torch.randn(512).cuda()
    h = torch.randn(B, S, 512).cuda(); ho = torch.full_like(h, float("nan"))
    T.capi.check(L.tm_to_out_residual_fwd(merged.data_ptr(), w.data_ptr(), b.data_ptr(), h.data_ptr(), B, S, P, ho.data_ptr(), T.capi.TM_FLAG_F16_ACT, st))
    sync("to_out")
    ref = h + (merged.float() @ w.float().T + b)[:, P:]
    print("to_out err", float((ho - ref).abs().max()))
if which in ("fc1", "all"):
    m = build_model(2, 1024).cuda()
    x = make_input("rand", 1, 3000, 1024, 1).cuda()
    wts = m._weights()
    h0 = torch.full((1, 3026, 512), float("nan"), device="cuda")
    T.capi.check(L.tm_fc1_fwd(C.byref(wts), x.data_ptr(), 1, 3000, h0.data_ptr(), 0, 0, st))
    sync("fc1")
    ref = torch.nn.functional.gelu(x @ m._fc1[0].weight.T + m._fc1[0].bias)
    print("fc1 err", float((h0[:, 1:3001] - ref).abs().max()), "dup equal", bool(torch.equal(h0[:, 3001:3026], h0[:, 1:26])), "cls", float((h0[:,0]-m.cls_token[0]).abs().max()))
