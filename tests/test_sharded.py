"""Token-sharded path, CPU suite: the shard plan's integers, and the host logic (halo exchanges, landmark
allreduce, attn3.v allgather + merge, PPEG halos) driven over gloo with world_size 2 and 3 by the torch restatement of the
per-kernel ops in tests/sharded_cpu_ops.py, compared with the oracle's single-process forward."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from common import build_model, make_input


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_shard_plan_integers():
    import transmil_b200 as T
    from transmil_b200.sharded import ShardPlan, bag_plan
    from oracle.transmil_oracle import plan as oracle_plan
    for n in (1, 2, 3, 4, 255, 256, 257, 1000, 6000, 20000, 50000, 100000, 200000, 299999):
        assert bag_plan(n) == oracle_plan(n)
        assert bag_plan(n) == T.capi.make_plan(n).as_dict()
    for n, world in [(6000, 2), (20000, 8), (50000, 4), (100000, 8), (200000, 2), (200000, 4), (200000, 8), (3000, 3)]:
        sp = ShardPlan(n, world)
        p = sp.p
        assert sp.shards[0].row0 == 0 and sp.shards[-1].row1 == p["Sp"]
        tok, rows_of_x = 0, []
        for r, s in enumerate(sp.shards):
            assert s.rows % 256 == 0 and s.rows > 0
            assert s.row0 == (sp.shards[r - 1].row1 if r else 0)
            assert s.tok0 == tok and s.S_local == s.ntok + (1 if r == 0 else 0) and s.rows == s.S_local + s.P_local
            tok = s.tok1
            # every token's source row of x: t < n ? t : t - n   (TransMIL.py:180)
            src = [t if t < n else t - n for t in range(s.tok0, s.tok1)]
            got = [i for a, b in s.x_ranges for i in range(a, b)]
            assert src == got
            # halo covers 3 image rows on both sides (or the image border)
            assert s.y_in0 == max(0, s.tok0 // p["G"] - 3) and s.y_in1 == min(p["G"], (s.tok1 - 1) // p["G"] + 4)
            assert s.halo_lo == s.tok0 - s.y_in0 * p["G"] and s.halo_hi == s.y_in1 * p["G"] - s.tok1
        assert tok == p["N"]
    with pytest.raises(ValueError):
        ShardPlan(300, 4)     # 2 blocks of 256 rows cannot feed 4 ranks
    with pytest.raises(ValueError):
        ShardPlan(5000, 20)   # 256-row shards: the 3-row PPEG halo (up to 4*71 tokens) would span more than one neighbour


def test_c_shard_plan_equals_python_shard_plan(built_lib):
    """tm_make_shard_plan (the plan tm_transmil_fwd_sharded runs on) holds the integers of ShardPlan / RankShard, and refuses
    the same bags."""
    import transmil_b200 as T
    from transmil_b200.sharded import ShardPlan, shard_plan_c
    cases = [(n, w) for n in (600, 3000, 6000, 20000, 49999, 50000, 100000, 200000, 1000000) for w in (1, 2, 3, 4, 8, 16)]
    checked = 0
    for n, world in cases:
        try:
            sp = ShardPlan(n, world)
        except ValueError:
            with pytest.raises(ValueError):
                shard_plan_c(n, world, 0)
            continue
        for r, s in enumerate(sp.shards):
            c = shard_plan_c(n, world, r)
            assert c.bag.as_dict() == sp.p
            got = {k: int(getattr(c, k)) for k in ("world", "rank", "row0", "rows", "P_local", "S_local", "has_cls", "tok0", "ntok",
                                                   "y_in0", "y_in1", "halo_lo", "halo_hi")}
            want = dict(world=world, rank=r, row0=s.row0, rows=s.rows, P_local=s.P_local, S_local=s.S_local, has_cls=int(s.has_cls),
                        tok0=s.tok0, ntok=s.ntok, y_in0=s.y_in0, y_in1=s.y_in1, halo_lo=s.halo_lo, halo_hi=s.halo_hi)
            assert got == want, (n, world, r)
            assert c.x_ranges() == [tuple(t) for t in s.x_ranges]
            assert int(c.n_up) == (sp.shards[r + 1].halo_lo if r + 1 < world else 0)
            assert int(c.n_dn) == (sp.shards[r - 1].halo_hi if r > 0 else 0)
            checked += 1
    assert checked > 100
    with pytest.raises(ValueError):
        shard_plan_c(50000, 17, 0)       # TM_SHARD_MAX_WORLD
    with pytest.raises(ValueError):
        shard_plan_c(50000, 4, 4)


def _worker(rank, world, port, n, D, n_classes, sd, x, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sharded_cpu_ops import OracleOps
        from transmil_b200.sharded import ShardPlan, sharded_forward
        torch.set_num_threads(2)
        plan = ShardPlan(n, world)
        logits = sharded_forward(OracleOps(sd, n_classes), plan, plan.take(x, rank))
        out[rank] = logits.clone()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n,D", [(2, 3000, 1024), (3, 3000, 512), (2, 1500, 2048)])
def test_sharded_host_logic_over_gloo(world, n, D):
    from oracle import transmil_oracle as O
    m = build_model(3, D)
    sd = {k: v.detach().clone() for k, v in m.state_dict().items()}
    x = make_input("randn", 1, n, D, 7)
    ref = O.forward(sd, x)
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), n, D, 3, sd, x, out), nprocs=world, join=True)
    for r in range(world):
        assert torch.allclose(out[r], ref, atol=2e-5, rtol=0), (r, out[r], ref)
        assert torch.equal(out[r], out[0])
