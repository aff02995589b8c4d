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
    T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 0, 1024, 0, T.capi.TM_FLAG_F16_ACT, st))
    assert torch.equal(part[0], full[0]) and torch.equal(part[1], full[1])
    assert bool((part[2] == 7.0).all())                       # v untouched by the q,k slice
    T.capi.check(L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 1024, 512, 100, T.capi.TM_FLAG_F16_ACT, st))
    assert torch.equal(part[2], full[2]) and torch.equal(part[0], full[0])
    rc = L.tm_qkv16_cols_fwd(xn.data_ptr(), w.data_ptr(), B, Sp, *ptrs, 256, 512, 0, T.capi.TM_FLAG_F16_ACT, st)
    assert rc != 0
