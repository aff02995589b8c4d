"""GPU suite: the tcgen05/TMA GEMM against torch fp32 matmul and the CUDA-core engine."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _tf32(x):
    xi = x.contiguous().view(torch.int32)
    r = ((xi >> 13) & 1) + 0xFFF
    return ((xi + r) & ~0x1FFF).view(torch.float32)


@pytest.mark.parametrize("M,N,K", [(128, 256, 32), (128, 256, 64), (256, 512, 512), (1000, 512, 1024), (5000, 1536, 512),
                                   (130, 768, 768), (4099, 1024, 2048), (777, 512, 384), (300, 512, 784)])
def test_gemm_tc_matches_fp32(built_lib, M, N, K):
    import transmil_b200 as T
    L = T.capi.lib()
    g = torch.Generator().manual_seed(M * 7 + N + K)
    A = _tf32(torch.randn(M, K, generator=g)).cuda()
    W = _tf32(torch.randn(N, K, generator=g) * 0.05).cuda()
    C_tc = torch.full((M, N), float("nan"), device="cuda")
    C_simt = torch.empty((M, N), device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    T.capi.check(L.tm_gemm_nt(A.data_ptr(), W.data_ptr(), C_tc.data_ptr(), M, N, K, 0, st))
    T.capi.check(L.tm_gemm_nt(A.data_ptr(), W.data_ptr(), C_simt.data_ptr(), M, N, K, T.capi.TM_FLAG_SIMT_ONLY, st))
    torch.cuda.synchronize()
    ref = (A.double() @ W.double().T)
    scale = ref.abs().max().item()
    assert torch.isfinite(C_tc).all()
    e_simt = (C_simt.double() - ref).abs().max().item() / scale
    e_tc = (C_tc.double() - ref).abs().max().item() / scale
    assert e_simt < 1e-5, e_simt
    assert e_tc < 1e-5, f"tcgen05 GEMM error {e_tc:.3e} (operands are tf32-exact, accumulation is fp32)"


def test_gemm_tc_truncates_unrounded_operands(built_lib):
    """Unrounded fp32 operands are truncated to tf32 by the tensor core: error ~2^-10 relative, not garbage."""
    import transmil_b200 as T
    L = T.capi.lib()
    g = torch.Generator().manual_seed(1)
    A = torch.rand(512, 1024, generator=g).cuda()
    W = (torch.randn(512, 1024, generator=g) * 0.03).cuda()
    C = torch.empty(512, 512, device="cuda")
    T.capi.check(L.tm_gemm_nt(A.data_ptr(), W.data_ptr(), C.data_ptr(), 512, 512, 1024, 0, torch.cuda.current_stream().cuda_stream))
    ref = A.double() @ W.double().T
    rel = (C.double() - ref).abs().max().item() / ref.abs().max().item()
    assert rel < 5e-3


@pytest.mark.parametrize("B,Sp", [(1, 3072), (2, 1280), (1, 50432)])
def test_qkv_column_slices_equal_the_full_projection(built_lib, B, Sp):
    """tm_transmil_fwd projects q,k first and v later beside the pinv chain (tm_qkv16_cols_fwd, capped grid): the slices
    must write exactly the bits of the one-launch projection, and nothing outside their own outputs."""
    import transmil_b200 as T
    L = T.capi.lib()
    st = torch.cuda.current_stream().cuda_stream
    g = torch.Generator().manual_seed(5)
    xn = torch.randn(B, Sp, 512, generator=g).half().cuda()
    w = (torch.randn(1536, 512, generator=g) * 0.04).half().cuda()
    shape = (B, 8, Sp, 64)
    full = [torch.empty(shape, dtype=torch.float16, device="cuda") for _ in range(3)]
    T.capi.check(L.tm_qkv16_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *[t.data_ptr() for t in full], T.capi.TM_FLAG_F16_ACT, st))
    ref = (xn.float() @ w.float().T).view(B, Sp, 3, 8, 64).permute(2, 0, 3, 1, 4)
    assert (full[0].float() - ref[0] * 0.125).abs().max().item() < 2e-2 and (full[2].float() - ref[2]).abs().max().item() < 5e-2
    part = [torch.full(shape, 7.0, dtype=torch.float16, device="cuda") for _ in range(3)]
    ptrs = [t.data_ptr() for t in part]
    T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 0, 1024, 0, None, T.capi.TM_FLAG_F16_ACT, st))
    assert torch.equal(part[0], full[0]) and torch.equal(part[1], full[1])
    assert bool((part[2] == 7.0).all())                       # v untouched by the q,k slice
    T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 1024, 512, 100, None, T.capi.TM_FLAG_F16_ACT, st))
    assert torch.equal(part[2], full[2]) and torch.equal(part[0], full[0])
    rc = L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 256, 512, 0, None, T.capi.TM_FLAG_F16_ACT, st)
    assert rc != 0


@pytest.mark.parametrize("B,Sp", [(1, 50432), (2, 32768), (1, 33280), (1, 200704)])
def test_landmarks_from_the_projection_epilogue(built_lib, B, Sp):
    """l >= 128: the q,k projection leaves per-tile column sums and tm_landmarks_parts_fwd finishes the landmark means;
    they must agree with the separate pass over the stored fp16 q / k (same values, different fp32 summation order) and
    with a float64 mean of those values."""
    import transmil_b200 as T
    L = T.capi.lib()
    st = torch.cuda.current_stream().cuda_stream
    l = Sp // 256
    g = torch.Generator().manual_seed(9)
    xn = torch.randn(B, Sp, 512, generator=g).half().cuda()
    xn[:, :37] = 0                                              # front pad rows
    w = (torch.randn(1536, 512, generator=g) * 0.04).half().cuda()
    shape = (B, 8, Sp, 64)
    q, k, v = (torch.empty(shape, dtype=torch.float16, device="cuda") for _ in range(3))
    part = torch.full((B * (Sp // 128) * 8 * 1024,), float("nan"), device="cuda")
    T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, q.data_ptr(), k.data_ptr(), v.data_ptr(), 0, 1024, 0,
                                     part.data_ptr(), T.capi.TM_FLAG_F16_ACT, st))
    assert not bool(torch.isnan(part).any())
    lm = [torch.empty(B, 8, 256, 64, device="cuda") for _ in range(4)]
    lm16 = [torch.empty(B, 8, 256, 64, dtype=torch.float16, device="cuda") for _ in range(4)]
    T.capi.check(L.tm_landmarks_parts_fwd(part.data_ptr(), B, Sp, l, lm[0].data_ptr(), lm[1].data_ptr(), lm16[0].data_ptr(),
                                          lm16[1].data_ptr(), st))
    T.capi.check(L.tm_landmarks16_fwd(q.data_ptr(), k.data_ptr(), B, Sp, l, lm[2].data_ptr(), lm[3].data_ptr(), lm16[2].data_ptr(),
                                      lm16[3].data_ptr(), st))
    for t, src in ((lm[0], q), (lm[1], k)):
        ref = src.double().view(B, 8, 256, l, 64).sum(3) / l
        assert (t.double() - ref).abs().max().item() < 2e-6
    assert (lm[0] - lm[2]).abs().max().item() < 2e-6 and (lm[1] - lm[3]).abs().max().item() < 2e-6
    assert (lm16[0].float() - lm16[2].float()).abs().max().item() < 2e-3
    # the same call twice gives the same bits (fixed summation order)
    lm2 = [torch.empty(B, 8, 256, 64, device="cuda") for _ in range(2)]
    T.capi.check(L.tm_landmarks_parts_fwd(part.data_ptr(), B, Sp, l, lm2[0].data_ptr(), lm2[1].data_ptr(), lm16[2].data_ptr(),
                                          lm16[3].data_ptr(), st))
    assert torch.equal(lm2[0], lm[0]) and torch.equal(lm2[1], lm[1])
    rc = L.tm_landmarks_parts_fwd(part.data_ptr(), 1, 2560, 10, lm[0].data_ptr(), lm[1].data_ptr(), lm16[0].data_ptr(), lm16[1].data_ptr(), st)
    assert rc != 0
