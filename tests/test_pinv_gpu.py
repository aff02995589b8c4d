"""GPU suite: the cluster kernel of the pseudo-inverse chain (csrc/pinv_pipe.cu, fp16 operands / fp32 accumulate) against the fp32
CUDA-core chain of the same library and against a float64 restatement of moore_penrose_iter_pinv (nystrom-attention 0.0.11:
global init scalar, exactly 6 iterations).  Sorted early and cheap: a hang here must not cost the whole suite."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _pinv64(A):
    A = A.double()
    absA = A.abs()
    Z = A.transpose(-1, -2) / (absA.sum(-1).max() * absA.sum(-2).max())
    eye = torch.eye(256, dtype=torch.float64, device=A.device)
    for _ in range(6):
        XZ = A @ Z
        Z = 0.25 * Z @ (13 * eye - XZ @ (15 * eye - XZ @ (7 * eye - XZ)))
    return Z


def _run(T, q, k, flags):
    L = T.capi.lib()
    B = q.shape[0]
    attn2 = torch.full((B, 8, 256, 256), float("nan"), device="cuda")
    Z = torch.full_like(attn2, float("nan"))
    sc = torch.empty(int(L.tm_pinv_scratch_floats(B)), device="cuda")
    T.capi.check(L.tm_pinv_fwd(q.data_ptr(), k.data_ptr(), B, attn2.data_ptr(), Z.data_ptr(), sc.data_ptr(), flags,
                               torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return attn2, Z


@pytest.mark.parametrize("B,scale", [(1, 0.3), (2, 0.1), (3, 0.5), (8, 0.3), (24, 0.2)])
def test_pinv_cluster_kernel_against_fp32_chain_and_float64(built_lib, B, scale):
    import transmil_b200 as T
    g = torch.Generator().manual_seed(B)
    q = (torch.randn(B, 8, 256, 64, generator=g) * scale).cuda()
    k = (torch.randn(B, 8, 256, 64, generator=g) * scale).cuda()
    a_tc, z_tc = _run(T, q, k, 0)
    a_f32, z_f32 = _run(T, q, k, T.capi.TM_FLAG_SIMT_ONLY)
    assert torch.isfinite(z_tc).all()
    assert torch.equal(a_tc, a_f32)                      # attn2 is computed in fp32 by both engines
    ref = _pinv64(a_f32)
    mx = ref.abs().max().item()
    e_f32 = (z_f32.double() - ref).abs().max().item() / mx
    e_tc = (z_tc.double() - ref).abs().max().item() / mx
    assert e_f32 < 1e-4, e_f32
    assert e_tc < 5e-3, f"fp16-operand cluster kernel vs float64 chain: {e_tc:.3e} of max|Z|"
    again = _run(T, q, k, 0)[1]
    assert torch.equal(again, z_tc), "the cluster kernel must be deterministic run to run"
