"""GPU suite: the tcgen05 strided batched GEMM of the backward (csrc/bgemm_tc.cu) against a float64 product and against the
fp32 CUDA-core engine, in all four transposition forms, with ragged shapes, batches, split-K, beta * L (aliasing C) and a
broadcast bias row.  Sorted early (file name) and run with its own timeout by tools/gpu_*.sh: a hang here must not cost the
whole suite."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(x):
    xi = x.contiguous().view(torch.int32)
    r = ((xi >> 13) & 1) + 0xFFF
    return ((xi + r) & ~0x1FFF).view(torch.float32)


def _operand(rows, cols, transposed, batch, gen, scale=1.0):
    """Logical [batch, rows, cols] operand (tf32-exact values); stored transposed when asked.  Returns (tensor, (s_row, s_col), stride)."""
    t = _tf32(torch.randn(batch, rows, cols, generator=gen) * scale)
    if transposed:
        store = t.transpose(1, 2).contiguous().cuda()      # [batch, cols, rows]
        return t, store, (1, rows), rows * cols
    return t, t.contiguous().cuda(), (cols, 1), rows * cols


def _run(L, T, A, a_str, B, b_str, C, M, N, K, batch, strides, alpha=1.0, beta=0.0, Lt=None, ldl=None, strideL=None, engine=0):
    st = torch.cuda.current_stream().cuda_stream
    prev = L.tm_bgemm_engine(engine)
    try:
        nsc = int(L.tm_bgemm_scratch_floats(M, N, K, batch))
        sc = torch.empty(max(nsc, 1), device="cuda")
        T.capi.check(L.tm_bgemm_ex(A.data_ptr(), B.data_ptr(), C.data_ptr(), Lt.data_ptr() if Lt is not None else None, M, N, K,
                                   a_str[0], a_str[1], b_str[0], b_str[1], N, N if ldl is None else ldl, batch, strides[0], strides[1],
                                   M * N, M * N if strideL is None else strideL, alpha, beta, sc.data_ptr() if nsc else None, nsc, st))
        torch.cuda.synchronize()
    finally:
        L.tm_bgemm_engine(prev)
    return nsc


FORMS = [(False, False), (False, True), (True, False), (True, True)]   # (A stored [K,M], B stored [N,K])
FEEDS = {"tma": 0, "cpasync": 2}                                        # tm_bgemm_engine codes of the two operand feeds


@pytest.mark.parametrize("feed", ["tma", "cpasync"])
@pytest.mark.parametrize("tA,tB", FORMS)
@pytest.mark.parametrize("M,N,K,batch", [(128, 64, 32, 1), (256, 256, 256, 8), (1280, 64, 256, 3), (256, 64, 1280, 4),
                                         (130, 100, 68, 2), (77, 300, 1000, 1), (512, 512, 8192, 1), (1536, 512, 4096, 1),
                                         (1000, 512, 1024, 1), (256, 1280, 64, 5)])
def test_bgemm_tc_all_forms(built_lib, tA, tB, M, N, K, batch, feed):
    import transmil_b200 as T
    L = T.capi.lib()
    g = torch.Generator().manual_seed(M + 3 * N + 5 * K + batch + (2 if tA else 0) + (1 if tB else 0))
    # A logical [M,K]: stored [M,K] (k contiguous) or [K,M]; B logical [K,N]: stored [K,N] (n contiguous) or [N,K]
    A_log, A_st, a_rc, sA = _operand(M, K, tA, batch, g)
    B_log, B_st, b_rc, sB = _operand(K, N, tB, batch, g, 0.1)
    a_str = (a_rc[0], a_rc[1])                  # (sam, sak)
    b_str = (b_rc[0], b_rc[1])                  # (sbk, sbn)
    ref = A_log.double() @ B_log.double()
    scale = ref.abs().max().item()
    C_tc = torch.full((batch, M, N), float("nan"), device="cuda")
    C_simt = torch.empty((batch, M, N), device="cuda")
    _run(L, T, A_st, a_str, B_st, b_str, C_tc, M, N, K, batch, (sA, sB), engine=FEEDS[feed])
    _run(L, T, A_st, a_str, B_st, b_str, C_simt, M, N, K, batch, (sA, sB), engine=1)
    assert torch.isfinite(C_tc).all()
    e_simt = (C_simt.cpu().double() - ref).abs().max().item() / scale
    e_tc = (C_tc.cpu().double() - ref).abs().max().item() / scale
    assert e_simt < 2e-5, e_simt
    assert e_tc < 2e-5, f"tcgen05 bgemm error {e_tc:.3e} (operands are tf32-exact, accumulation is fp32)"


def test_bgemm_tc_split_k_is_used_and_deterministic(built_lib):
    import transmil_b200 as T
    L = T.capi.lib()
    M, N, K = 1536, 512, 81920 // 4
    g = torch.Generator().manual_seed(7)
    dY = _tf32(torch.randn(K, M, generator=g) * 0.1).cuda()      # A stored [K,M]: dW = dY^T X
    X = _tf32(torch.randn(K, N, generator=g)).cuda()
    out = [torch.empty(M, N, device="cuda") for _ in range(2)]
    for o in out:
        nsc = _run(L, T, dY, (1, M), X, (N, 1), o, M, N, K, 1, (0, 0))
        assert nsc > 0, "a weight-gradient shape must be split along K"
    assert torch.equal(out[0], out[1])
    ref = dY.double().T @ X.double()
    assert (out[0].double() - ref).abs().max().item() / ref.abs().max().item() < 2e-5


def test_bgemm_tc_epilogue_terms(built_lib):
    """alpha, beta * L with L aliasing C (the pinv backward accumulates in place), a broadcast bias row (ldl = 0) and an
    unaligned / odd-N output."""
    import transmil_b200 as T
    L = T.capi.lib()
    g = torch.Generator().manual_seed(11)
    M, N, K, batch = 256, 256, 256, 4
    A = _tf32(torch.randn(batch, M, K, generator=g)).cuda()
    B = _tf32(torch.randn(batch, K, N, generator=g) * 0.1).cuda()
    C0 = torch.randn(batch, M, N, generator=g).cuda()
    C = C0.clone()
    _run(L, T, A, (K, 1), B, (N, 1), C, M, N, K, batch, (M * K, K * N), alpha=-0.25, beta=1.0, Lt=C)
    ref = -0.25 * (A.double() @ B.double()) + C0.double()
    assert (C.double() - ref).abs().max().item() / ref.abs().max().item() < 2e-5
    bias = torch.randn(N, generator=g).cuda()
    C2 = torch.empty(batch, M, N, device="cuda")
    _run(L, T, A, (K, 1), B, (N, 1), C2, M, N, K, batch, (M * K, K * N), beta=1.0, Lt=bias, ldl=0, strideL=0)
    ref2 = A.double() @ B.double() + bias.double()
    assert (C2.double() - ref2).abs().max().item() / ref2.abs().max().item() < 2e-5
    # odd N, unaligned K: element path of the loaders and scalar stores
    M3, N3, K3 = 200, 37, 101
    A3 = _tf32(torch.randn(M3, K3, generator=g)).cuda()
    B3 = _tf32(torch.randn(K3, N3, generator=g)).cuda()
    C3 = torch.full((M3, N3), float("nan"), device="cuda")
    _run(L, T, A3, (K3, 1), B3, (N3, 1), C3, M3, N3, K3, 1, (0, 0))
    ref3 = A3.double() @ B3.double()
    assert (C3.double() - ref3).abs().max().item() / ref3.abs().max().item() < 2e-5


def test_pinv_bwd_engines_agree(built_lib):
    """tm_pinv_bwd (97 batched 256^3 products, every transposition form, + diag * I) on the tensor cores vs the fp32 engine."""
    import transmil_b200 as T
    from transmil_b200 import training
    L = training.bind(T.capi.lib())
    nb = 8
    g = torch.Generator().manual_seed(3)
    A2 = torch.softmax(torch.randn(nb, 256, 256, generator=g) * 0.3, -1).cuda()
    dZ0 = (torch.randn(nb, 256, 256, generator=g) * 1e-3).cuda()
    st = torch.cuda.current_stream().cuda_stream
    outs = []
    for engine in (1, 0, 2):
        prev = L.tm_bgemm_engine(engine)
        dZ = dZ0.clone()
        dA = torch.empty_like(A2)
        sc = torch.empty(int(L.tm_pinv_bwd_scratch_floats(nb)), device="cuda")
        T.capi.check(L.tm_pinv_bwd(A2.data_ptr(), dZ.data_ptr(), nb, dA.data_ptr(), sc.data_ptr(), st))
        torch.cuda.synchronize()
        L.tm_bgemm_engine(prev)
        outs.append(dA)
    for o in outs[1:]:
        rel = (outs[0] - o).abs().max().item() / outs[0].abs().max().item()
        assert torch.isfinite(o).all()
        assert rel < 5e-3, f"pinv backward, tf32 tensor cores vs fp32: {rel:.3e}"
    assert torch.equal(outs[1], outs[2]), "TMA feed and loader-warp feed must give the same bits (same operands, same order)"
