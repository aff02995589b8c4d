"""tm_attn3v_tc_fwd_static: the attn3 @ v kernels with their chunk partials in the library's static device array instead
of caller-owned scratch (no allocation inside the call; include/transmil_b200.h).  No counterpart in the reference.
Collected last on purpose.

The GPU test calls it through the C ABI several times in a row -- with a different launch geometry in between, so that the
shared array holds another call's partials -- and compares every result with tm_attn3v_tc_fwd on the same inputs (same
kernels, same chunking: the tolerance of tests/test_attn_tc_gpu.py, 3e-3, is the gate; the measured difference is 0) and
with a torch fp32 reference of the op."""
import warnings

import pytest
import torch


def _lib():
    import transmil_b200 as T
    return T, T.capi.lib()


TOL = 3e-3   # tests/test_attn_tc_gpu.py::test_attn3v_tc


def _inputs(B, Sp, seed):
    g = torch.Generator().manual_seed(seed)
    qlm = (torch.randn(B, 8, 256, 64, generator=g) * 0.3).half().cuda()
    k = (torch.randn(B, 8, Sp, 64, generator=g) * 0.3).half().cuda()
    k[:, :, :37] = 0          # zero-score front-pad rows take part in the softmax
    v = torch.randn(B, 8, Sp, 64, generator=g).half().cuda()
    return qlm, k, v


@pytest.mark.gpu
def test_static_scratch_entry_matches_caller_scratch_entry(built_lib):
    T, L = _lib()
    st = torch.cuda.current_stream().cuda_stream
    shapes = [(1, 1280, 11), (2, 2560, 12), (1, 1280, 13), (1, 50432, 14), (18, 256, 15)]
    # everything the calls need exists before the first of them: the static-scratch calls themselves allocate nothing
    cases = []
    for B, Sp, seed in shapes:
        qlm, k, v = _inputs(B, Sp, seed)
        want = torch.full((B, 8, 256, 64), float("nan"), device="cuda")
        scratch = torch.empty(int(L.tm_attn3v_tc_scratch_floats(B, Sp)), device="cuda")
        T.capi.check(L.tm_attn3v_tc_fwd(qlm.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, want.data_ptr(), scratch.data_ptr(), 0, st))
        got = torch.full((B, 8, 256, 64), float("nan"), device="cuda")
        cases.append((B, Sp, qlm, k, v, want, got))
    torch.cuda.synchronize()
    free_before = None
    for i, (B, Sp, qlm, k, v, want, got) in enumerate(cases):      # back to back on one stream, no sync in between
        T.capi.check(L.tm_attn3v_tc_fwd_static(qlm.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, got.data_ptr(), 0, st))
        if i == 0:                                                  # the first call looked the array's address up
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info()[0]
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    if free_after < free_before:
        warnings.warn(f"free device memory went from {free_before} to {free_after} bytes across the static-scratch calls")
    for B, Sp, qlm, k, v, want, got in cases:
        assert torch.isfinite(got).all(), (B, Sp)
        d = (got - want).abs().max().item()
        assert d <= TOL, (B, Sp, d)
        ref = (qlm.float() @ k.float().transpose(-1, -2)).softmax(-1) @ v.float()
        err = (got - ref).abs().max().item()
        assert err < TOL, (B, Sp, err)
        print(f"static vs caller scratch B={B} Sp={Sp}: max|d|={d:.3g} bit-identical={torch.equal(got, want)} vs torch {err:.3g}")
    # second round over the same buffers: the array now holds the partials of the (18, 256) call
    for B, Sp, qlm, k, v, want, got in cases[:2]:
        got.fill_(float("nan"))
        T.capi.check(L.tm_attn3v_tc_fwd_static(qlm.data_ptr(), k.data_ptr(), v.data_ptr(), B, Sp, got.data_ptr(), 0, st))
    torch.cuda.synchronize()
    for B, Sp, qlm, k, v, want, got in cases[:2]:
        d = (got - want).abs().max().item()
        assert d <= TOL, (B, Sp, d)


def test_static_scratch_entry_rejects_sizes_beyond_the_array(built_lib):
    """B = 19 needs 152 (bag, head) partial slots, the array holds 148: refused before anything touches a device (so this
    also runs in the CPU suite)."""
    T, L = _lib()
    with pytest.raises(ValueError, match="static array holds"):
        T.capi.check(L.tm_attn3v_tc_fwd_static(1, 1, 1, 19, 256, 1, 0, None))
    with pytest.raises(ValueError, match="multiple of 256"):
        T.capi.check(L.tm_attn3v_tc_fwd_static(1, 1, 1, 1, 100, 1, 0, None))
    with pytest.raises(ValueError, match="NULL"):
        T.capi.check(L.tm_attn3v_tc_fwd_static(1, 1, 1, 1, 256, None, 0, None))
