"""GPU suite, SURVEY 8(f) row N2: bags stored as half-precision feature files.  The reference upcasts them with x.float()
(TransMIL.py:174), which is exact, so the oracle is run on the upcast halves; the CUDA path reads the halves directly
(TM_FLAG_X_F16: fp16 x fp16 -> f32 first contraction).  Same tolerances as the fp32 bags (logits 1e-3, Y_prob 1e-4)."""
import ctypes as C

import pytest
import torch

from common import LOGIT_TOL, PROB_TOL, build_model, make_input
from oracle import transmil_oracle as O

pytestmark = pytest.mark.gpu


def _oracle_logits(m, x32):
    sd = {k: v.detach().cpu() for k, v in m.state_dict().items()}
    return O.forward(sd, x32.cpu())


@pytest.mark.parametrize("D,nc,n,B,gen", [(1024, 2, 6000, 1, "rand"), (2048, 3, 1000, 2, "rand"), (768, 2, 300, 1, "randn"),
                                          (512, 2, 257, 1, "randn"), (1024, 2, 1, 1, "rand")])
def test_half_bags_match_oracle_on_the_upcast_halves(built_lib, D, nc, n, B, gen):
    m = build_model(nc, D).cuda()
    x16 = make_input(gen, B, n, D, 4242).half()
    ref = _oracle_logits(m, x16.float()).double()
    with torch.no_grad():
        got = m(x16.cuda())
        got_host = m(x16.pin_memory())          # host halves: copied as halves
    assert got.dtype == torch.float32 and got.shape == (B, nc)
    assert (got.double().cpu() - ref).abs().max().item() <= LOGIT_TOL
    assert (torch.softmax(got.double().cpu(), 1) - torch.softmax(ref, 1)).abs().max().item() <= PROB_TOL
    assert torch.equal(got, got_host)


def test_half_bags_through_the_slide_loop_and_the_simt_engine_fallback(built_lib):
    import transmil_b200 as T
    m = build_model(2, 1024).cuda()
    bags = [make_input("rand", 1, n, 1024, 7 + i).half().pin_memory() for i, n in enumerate([3000, 500, 6000])]
    with torch.no_grad():
        direct = [m(b.cuda()).cpu() for b in bags]
    got = list(T.StreamedInference(m)(bags))
    for g, d in zip(got, direct):
        assert torch.equal(g, d)
    # mixed dtypes reuse the same byte slots
    mixed = [bags[0], bags[1].float().pin_memory(), bags[2]]
    got = list(T.StreamedInference(m)(mixed))
    assert torch.equal(got[0], direct[0]) and torch.equal(got[2], direct[2])
    # the fp32 CUDA-core cross-check engine has no fp16 first contraction: halves are upcast on the device instead
    m.flags = T.capi.TM_FLAG_SIMT_ONLY
    with torch.no_grad():
        simt = m(bags[1].cuda()).cpu()
    assert (simt - direct[1]).abs().max().item() <= LOGIT_TOL


def test_x_f16_flag_is_refused_by_the_simt_engine_at_the_c_abi(built_lib):
    import transmil_b200 as T
    L = T.capi.lib()
    m = build_model(2, 1024).cuda()
    w = m._weights()
    B, n = 1, 300
    flags = T.capi.TM_FLAG_SIMT_ONLY | T.capi.TM_FLAG_X_F16
    nbytes = C.c_size_t()
    T.capi.check(L.tm_workspace_bytes(B, n, C.byref(w), flags, C.byref(nbytes)))
    ws = torch.empty(nbytes.value, dtype=torch.uint8, device="cuda")
    x16 = torch.rand(B, n, 1024, device="cuda").half()
    logits = torch.empty(B, 2, device="cuda")
    rc = L.tm_transmil_fwd(C.byref(w), x16.data_ptr(), B, n, logits.data_ptr(), ws.data_ptr(), ws.numel(), flags,
                           torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"TM_FLAG_X_F16" in L.tm_last_error()
