"""Per-stage time of the token-sharded forward on every rank (CUDA events at the stage boundaries of tm_transmil_fwd_sharded,
pinv / attn3.v overlap off while profiling), plus the un-profiled step time of the three ways to run it.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/shard_profile.py --patches 200000 --features 2048
"""
import argparse
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import transmil_b200 as T  # noqa: E402
from transmil_b200.sharded import TokenShardedTransMIL  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--patches", type=int, default=200000)
    ap.add_argument("--features", type=int, default=2048)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--nccl", action="store_true", help="also time the Python-issued NCCL schedule")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    L = T.capi.lib()
    torch.manual_seed(2021)
    model = T.TransMIL(n_classes=2, in_features=a.features).eval().to(dev)
    n = a.patches
    x_full = torch.rand(1, n, a.features, generator=torch.Generator().manual_seed(4321))
    out = {"n": n, "features": a.features, "world": world}

    def timed(sh, x):
        for _ in range(3):
            sh(x, n=n)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.steps):
            sh(x, n=n)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    with torch.no_grad():
        def note(msg):
            print(f"[rank {rank}] {msg}", file=sys.stderr, flush=True)
        eager = TokenShardedTransMIL(model)
        x = eager.plan(n).take(x_full, rank).to(dev)
        note("peer comm up, timing eager")
        out["peer_eager_ms"] = timed(eager, x)
        note(f"eager {out['peer_eager_ms']:.3f} ms, status {eager.comm.status()}")
        graphed = TokenShardedTransMIL(model, cuda_graph=True, static_input=True)
        out["peer_graph_ms"] = timed(graphed, x)
        note(f"graph {out['peer_graph_ms']:.3f} ms, status {graphed.comm.status()}")
        if a.nccl:
            out["nccl_ms"] = timed(TokenShardedTransMIL(model, transport="nccl"), x)
        note("variants done, stage profile")
        # stage profile (events serialise the side stream: the sum is the no-overlap time)
        L.tm_profile_enable(1)
        T.capi.profile_read()
        for _ in range(a.steps):
            eager(x, n=n)
        torch.cuda.synchronize()
        prof = T.capi.profile_read()
        L.tm_profile_enable(0)
        mine = {k: round(1e3 * v[0] / a.steps, 1) for k, v in prof.items()}       # us per forward
        mine["total_us"] = round(sum(mine.values()), 1)
        allp = [None] * world
        dist.all_gather_object(allp, mine)
        out["status"] = eager.comm.status() | graphed.comm.status()
    if rank == 0:
        print(json.dumps(out))
        keys = list(allp[0].keys())
        print("stage (us per forward)".ljust(24) + "".join(f"rank{r}".rjust(9) for r in range(world)))
        for k in keys:
            print(k.ljust(24) + "".join(f"{p.get(k, 0):9.1f}" for p in allp))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
