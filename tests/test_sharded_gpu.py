"""GPU suite of the token-sharded path: every kernel that produces or consumes data crossing a rank boundary,
checked on ONE GPU by splitting a sequence into shards by hand and comparing with the unsharded kernels; the whole sharded
forward over a 1-rank NCCL group against the unsharded forward and the oracle; and (when the box has >= 2 GPUs) the real
multi-rank run launched with torchrun."""
import ctypes as C
import os
import subprocess
import sys

import pytest
import torch
import torch.nn.functional as F

from common import LOGIT_TOL, PROB_TOL, ROOT, build_model, make_input

pytestmark = pytest.mark.gpu


def _lib():
    import transmil_b200 as T
    return T, T.capi.lib()


def _st():
    return torch.cuda.current_stream().cuda_stream


@pytest.mark.parametrize("n,cuts", [(3000, [0, 1000, 3025]), (50000, [0, 17000, 33333, 50176]), (300, [0, 324])])
def test_ppeg_rows_equal_full_ppeg(built_lib, n, cuts):
    """tm_ppeg_rows_fwd on token ranges (own rows + 3-row halo only) reassembles to tm_ppeg_fwd bit for bit, and the full
    kernel matches torch's three depthwise convs (TransMIL.py:60-75)."""
    T, L = _lib()
    m = build_model(2, 1024).cuda()
    w = m._weights()
    p = T.capi.make_plan(n)
    G, N, S = int(p.G), int(p.N), int(p.S)
    cuts = [min(c, N) for c in cuts]
    h = torch.randn(1, S, 512, generator=torch.Generator().manual_seed(n)).cuda()
    full = torch.full_like(h, float("nan"))
    packed = torch.empty(50 * 512, device="cuda")
    T.capi.check(L.tm_ppeg_fwd(C.byref(w), h.data_ptr(), 1, G, full.data_ptr(), packed.data_ptr(), _st()))
    img = h[:, 1:].transpose(1, 2).reshape(1, 512, G, G)
    pl = m.pos_layer
    ref = (F.conv2d(img, pl.proj.weight, pl.proj.bias, 1, 3, groups=512) + img + F.conv2d(img, pl.proj1.weight, pl.proj1.bias, 1, 2, groups=512)
           + F.conv2d(img, pl.proj2.weight, pl.proj2.bias, 1, 1, groups=512)).flatten(2).transpose(1, 2)
    assert torch.equal(full[:, 0], h[:, 0])
    assert (full[:, 1:] - ref).abs().max().item() < 2e-5
    for a, b in zip(cuts[:-1], cuts[1:]):
        y0, y1 = max(0, a // G - 3), min(G, (b - 1) // G + 4)
        local = h[0, 1 + y0 * G: 1 + y1 * G].clone()          # only the rows a rank would hold
        out = torch.full((b - a, 512), float("nan"), device="cuda")
        T.capi.check(L.tm_ppeg_rows_fwd(C.byref(w), local.data_ptr(), y0, y1, G, a, b, out.data_ptr(), packed.data_ptr(), _st()))
        torch.cuda.synchronize()
        assert torch.equal(out, full[0, 1 + a: 1 + b])
    with pytest.raises(ValueError):
        T.capi.check(L.tm_ppeg_rows_fwd(C.byref(w), h.data_ptr(), 5, G, G, 0, N, full.data_ptr(), packed.data_ptr(), _st()))


@pytest.mark.parametrize("Sp,l,cut", [(2560, 10, 1024), (50432, 197, 25088), (1280, 5, 256)])
def test_landmark_sums_and_attn3v_partials_merge(built_lib, Sp, l, cut):
    T, L = _lib()
    g = torch.Generator().manual_seed(Sp)
    q = (torch.randn(8, Sp, 64, generator=g) * 0.3).half().cuda()
    k = (torch.randn(8, Sp, 64, generator=g) * 0.3).half().cuda()
    v = torch.randn(8, Sp, 64, generator=g).half().cuda()
    # reference: the unsharded kernels
    q_lm, k_lm = torch.empty(8, 256, 64, device="cuda"), torch.empty(8, 256, 64, device="cuda")
    qlm16, klm16 = torch.empty(8, 256, 64, dtype=torch.float16, device="cuda"), torch.empty(8, 256, 64, dtype=torch.float16, device="cuda")
    T.capi.check(L.tm_landmarks16_fwd(q.data_ptr(), k.data_ptr(), 1, Sp, l, q_lm.data_ptr(), k_lm.data_ptr(), qlm16.data_ptr(), klm16.data_ptr(), _st()))
    W3 = torch.empty(8, 256, 64, device="cuda")
    scratch = torch.empty(int(L.tm_attn3v_tc_scratch_floats(1, Sp)), device="cuda")
    T.capi.check(L.tm_attn3v_tc_fwd(qlm16.data_ptr(), k.data_ptr(), v.data_ptr(), 1, Sp, W3.data_ptr(), scratch.data_ptr(), 0, _st()))
    # two shards [0,cut) and [cut,Sp)
    pf = int(L.tm_attn3v_partial_floats(1))
    total = torch.zeros(2, 8, 256, 64, device="cuda")
    gathered = torch.empty(2, pf, device="cuda")
    shards = [(0, cut), (cut, Sp)]
    parts = [(q[:, a:b].contiguous(), k[:, a:b].contiguous(), v[:, a:b].contiguous()) for a, b in shards]
    for (a, b), (qs, ks, vs) in zip(shards, parts):
        sums = torch.full((2, 8, 256, 64), float("nan"), device="cuda")
        T.capi.check(L.tm_landmark_sums16_fwd(qs.data_ptr(), ks.data_ptr(), 1, b - a, a, l, sums.data_ptr(), _st()))
        total += sums                                        # the allreduce
    q_lm2, k_lm2 = torch.empty_like(q_lm), torch.empty_like(k_lm)
    qlm16b, klm16b = torch.empty_like(qlm16), torch.empty_like(klm16)
    T.capi.check(L.tm_landmark_finish_fwd(total.data_ptr(), 1, l, q_lm2.data_ptr(), k_lm2.data_ptr(), qlm16b.data_ptr(), klm16b.data_ptr(), _st()))
    torch.cuda.synchronize()
    ref_q = q.float().reshape(8, 256, l, 64).sum(2) / l
    assert (q_lm2 - ref_q).abs().max().item() < 1e-6 and (q_lm2 - q_lm).abs().max().item() < 1e-6
    assert (k_lm2 - k_lm).abs().max().item() < 1e-6
    assert (qlm16b.float() - qlm16.float()).abs().max().item() < 1e-3
    for i, ((a, b), (qs, ks, vs)) in enumerate(zip(shards, parts)):
        sc = torch.empty(int(L.tm_attn3v_tc_scratch_floats(1, b - a)), device="cuda")
        T.capi.check(L.tm_attn3v_tc_partial_fwd(qlm16.data_ptr(), ks.data_ptr(), vs.data_ptr(), 1, b - a, gathered[i].data_ptr(), sc.data_ptr(), _st()))
    W3m = torch.full_like(W3, float("nan"))
    T.capi.check(L.tm_attn3v_merge_fwd(gathered.data_ptr(), 2, 1, W3m.data_ptr(), _st()))
    torch.cuda.synchronize()
    ref = (qlm16.float() @ k.float().transpose(-1, -2)).softmax(-1) @ v.float()
    assert (W3m - ref).abs().max().item() < 3e-3
    assert (W3m - W3).abs().max().item() < 1e-3


@pytest.mark.parametrize("Sp,cut", [(2560, 1024), (6144, 3072)])
def test_conv_halo_fix_closes_the_shard_boundary(built_lib, Sp, cut):
    T, L = _lib()
    g = torch.Generator().manual_seed(Sp + 1)
    q = (torch.randn(8, Sp, 64, generator=g) * 0.3).half().cuda()
    klm = (torch.randn(8, 256, 64, generator=g) * 0.3).half().cuda()
    mt16 = torch.randn(8, 64, 256, generator=g).half().cuda()
    v = torch.randn(8, Sp, 64, generator=g).half().cuda()
    w = (torch.randn(8, 1, 33, 1, generator=g) * 0.2).cuda()
    F16 = T.capi.TM_FLAG_F16_ACT
    full = torch.empty(Sp, 512, dtype=torch.float16, device="cuda")
    T.capi.check(L.tm_nystrom_out_tc_fwd(q.data_ptr(), klm.data_ptr(), mt16.data_ptr(), v.data_ptr(), w.data_ptr(), 1, Sp, full.data_ptr(), F16, _st()))
    outs = []
    for a, b in [(0, cut), (cut, Sp)]:
        qs, vs = q[:, a:b].contiguous(), v[:, a:b].contiguous()
        mg = torch.empty(b - a, 512, dtype=torch.float16, device="cuda")
        T.capi.check(L.tm_nystrom_out_tc_fwd(qs.data_ptr(), klm.data_ptr(), mt16.data_ptr(), vs.data_ptr(), w.data_ptr(), 1, b - a, mg.data_ptr(), F16, _st()))
        vp = v[:, a - 16:a].contiguous() if a > 0 else None
        vn = v[:, b:b + 16].contiguous() if b < Sp else None
        raw = mg.clone()
        T.capi.check(L.tm_conv_halo_fix_fwd(vp.data_ptr() if vp is not None else 0, vn.data_ptr() if vn is not None else 0, w.data_ptr(), 1, b - a, mg.data_ptr(), _st()))
        outs.append((raw, mg))
    torch.cuda.synchronize()
    raw = torch.cat([o[0] for o in outs]).float()
    fixed = torch.cat([o[1] for o in outs]).float()
    assert (raw - full.float()).abs().max().item() > 0.05          # without the halo the boundary rows are wrong
    assert (fixed - full.float()).abs().max().item() < 6e-3        # fp16 double rounding on 32 rows
    inner = torch.ones(Sp, dtype=torch.bool)
    inner[cut - 16: cut + 16] = False
    assert torch.equal(fixed[inner], full.float()[inner])


def test_fc1_rows_equal_rows_of_fc1(built_lib):
    T, L = _lib()
    for D in (1024, 2048):
        m = build_model(2, D).cuda()
        w = m._weights()
        n = 3000
        p = T.capi.make_plan(n)
        x = make_input("rand", 1, n, D, 3).cuda()
        h0 = torch.empty(1, int(p.S), 512, device="cuda")
        f0 = 1024 if D == 2048 else 0
        sc = torch.empty(max(1, n * f0), device="cuda")
        T.capi.check(L.tm_fc1_fwd(C.byref(w), x.data_ptr(), 1, n, h0.data_ptr(), sc.data_ptr(), 0, _st()))
        rows = x[0, 700:2011].contiguous()
        out = torch.empty(rows.shape[0], 512, device="cuda")
        T.capi.check(L.tm_fc1_rows_fwd(C.byref(w), rows.data_ptr(), rows.shape[0], out.data_ptr(), sc.data_ptr(), 0, _st()))
        torch.cuda.synchronize()
        assert (out - h0[0, 701:2012]).abs().max().item() < 1e-6


@pytest.fixture(scope="module")
def one_rank_group():
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29531")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    yield
    dist.destroy_process_group()


@pytest.mark.parametrize("transport", ["peer", "nccl"])
@pytest.mark.parametrize("n,D", [(3000, 1024), (6000, 2048), (20000, 512)])
def test_sharded_forward_one_rank_matches_unsharded_and_oracle(built_lib, one_rank_group, n, D, transport):
    from oracle import transmil_oracle as O
    from transmil_b200.sharded import TokenShardedTransMIL
    m = build_model(2, D).cuda()
    x = make_input("rand", 1, n, D, 11)
    with torch.no_grad():
        plain = m(x.cuda()).cpu()
    sharded = TokenShardedTransMIL(m, cuda_graph=True, transport=transport)
    got = sharded(x).cpu()                 # first call: eager
    again = sharded(x).cpu()               # second call: captured into a CUDA graph and replayed
    third = sharded(x * 1.0).cpu()         # replay with a fresh input tensor
    assert torch.equal(again, got) and torch.equal(third, got)
    other = sharded(make_input("rand", 1, n, D, 12)).cpu()
    assert not torch.equal(other, got)     # the replay really reads its new input
    ref = O.forward({k: v.cpu() for k, v in m.state_dict().items()}, x)
    assert (got - ref).abs().max().item() < LOGIT_TOL
    assert (torch.softmax(got, 1) - torch.softmax(ref, 1)).abs().max().item() < PROB_TOL
    assert (got - plain).abs().max().item() < 2e-4
    if transport == "peer":
        assert sharded.comm.status() == 0


@pytest.mark.parametrize("transport", ["peer", "nccl"])
def test_sharded_path_follows_parameter_updates(built_lib, one_rank_group, transport):
    """the operand copies of the GEMM weights are re-derived when a parameter changes after TokenShardedTransMIL was
    built (load_state_dict / optimizer step), with and without a captured graph."""
    from transmil_b200.sharded import TokenShardedTransMIL
    m = build_model(2, 1024).cuda()
    x = make_input("rand", 1, 3000, 1024, 21)
    for graph in (False, True):
        sharded = TokenShardedTransMIL(m, cuda_graph=graph, transport=transport)
        first = sharded(x).cpu()
        sharded(x)
        other = build_model(2, 1024, seed=77).state_dict()
        m.load_state_dict(other)
        with torch.no_grad():
            plain = m(x.cuda()).cpu()
        got = sharded(x).cpu()
        assert not torch.equal(got, first)
        assert (got - plain).abs().max().item() < 2e-4, "the sharded path kept operand copies of the old weights"
        m.load_state_dict(build_model(2, 1024).state_dict())


@pytest.mark.parametrize("world", [2, 4])
def test_sharded_forward_multi_gpu(built_lib, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(ROOT, "tests", "run_sharded_gpu.py")], capture_output=True, text=True, timeout=240)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "SHARDED-OK" in r.stdout
