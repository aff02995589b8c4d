"""Which pinv engine for many (bag, head) pairs?  The cluster kernel is latency-oriented (4 SMs per head, 24 dependent
steps); with hundreds of heads it runs in waves.  Compares it with the batched tcgen05 GEMM chain (TM_FLAG_PINV_GEMM_CHAIN)."""
import sys, torch
sys.path.insert(0, '/root/repo')
import transmil_b200 as T
for B, n, D in [(64, 1000, 2048), (16, 1000, 2048), (8, 6000, 1024), (4, 20000, 1024), (2, 50000, 1024)]:
    torch.manual_seed(2021)
    m = T.TransMIL(n_classes=2, in_features=D).eval().cuda()
    x = torch.rand(B, n, D, generator=torch.Generator().manual_seed(1)).cuda()
    out = {}
    for name, fl in [("cluster", 0), ("gemm_chain", T.capi.TM_FLAG_PINV_GEMM_CHAIN)]:
        m.flags = fl
        with torch.no_grad():
            for _ in range(3):
                y = m(x)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(10):
                y = m(x)
            e1.record(); torch.cuda.synchronize()
        out[name] = (e0.elapsed_time(e1) / 10, y.cpu())
    print(f"B={B} n={n} D={D}: cluster {out['cluster'][0]:.3f} ms, gemm_chain {out['gemm_chain'][0]:.3f} ms, max|dlogit| {float((out['cluster'][1]-out['gemm_chain'][1]).abs().max()):.2e}", flush=True)
