"""Times one TRAINING step of the drop-in TransMIL (BASELINE config 2: DeepGraft/TransMIL_feat_norm_rest.yaml -- 2048-d bags,
RAdam lr 1e-4 wd 0.01, CE against the one-hot label) on one B200, or bag-data-parallel under torchrun (gradients summed with
one allreduce per step over NCCL).  Prints one JSON line (same keys as bench.py where they apply).  bench.py imports
`measure()` for its `train_step` sub-object; the headline metric of bench.py stays the 50 k-patch inference bag.

  python tools/bench_train.py [--bags 64 --patches 1000 | --bags 1 --patches 20000] [--steps 5 --warmup 3]
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/bench_train.py ...
"""
import argparse
import json
import os
import sys
from types import SimpleNamespace

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import transmil_b200 as T  # noqa: E402
from transmil_b200 import optim, plugin  # noqa: E402


def measure(bags=64, patches=1000, features=2048, classes=2, steps=5, warmup=3, opt_name="lookahead_radam", engine="tc", rank=0, world=1,
            dev=None):
    """One rank's part of the measurement; every rank of the group must call it (the gradient allreduce is collective).
    Returns the result dict on every rank (timings are the max over ranks)."""
    dev = dev if dev is not None else torch.device("cuda", torch.cuda.current_device())
    L = T.capi.lib()
    prev_engine, prev_fwd = L.tm_bgemm_engine(-1), T.training.FORWARD_ENGINE
    if engine == "simt":
        L.tm_bgemm_engine(1)
        T.training.FORWARD_ENGINE = "simt"
    try:
        torch.manual_seed(2021)
        model = T.TransMIL(n_classes=classes, in_features=features).to(dev).train()
        opt = optim.create_optimizer(SimpleNamespace(opt=opt_name, lr=1e-4, weight_decay=0.01), model)
        g = torch.Generator().manual_seed(1000 + rank)
        x = torch.rand(bags, patches, features, generator=g).to(dev)
        y = torch.randint(0, classes, (bags,), generator=g).to(dev)
        ar_events = []      # (start, end) around the gradient exchange of every timed step; read after the timed region

        def one_step(timed=False):
            opt.zero_grad(set_to_none=True)
            loss, _, _ = plugin.training_step(model, (x, y, None))
            loss.sum().backward()
            if timed and world > 1:
                ar_events.append((torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)))
                ar_events[-1][0].record()
            optim.allreduce_gradients(model.parameters())
            if timed and world > 1:
                ar_events[-1][1].record()
            opt.step()
            return loss

        for _ in range(warmup):
            one_step()
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        L.tm_reset_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            loss = one_step(timed=True)
        e1.record()
        torch.cuda.synchronize(dev)
        ar_ms = [a.elapsed_time(b) for a, b in ar_events]
        t = torch.tensor([e0.elapsed_time(e1) / steps, sum(ar_ms) / max(len(ar_ms), 1)], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ar = float(t[0]), float(t[1])
        launches = int(L.tm_launch_count())      # process-wide counter: forward, backward (autograd thread) and optimizer launches
        n_grad = sum(p.numel() for p in model.parameters())
        return {"metric": "training bags/s", "engine": engine, "value": bags * world / (ms / 1e3), "unit": "bags/s",
                "tokens_per_s": bags * patches * world / (ms / 1e3), "n_gpus": world, "steps": steps, "warmup": warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
                "dtype": "f32 storage and master weights; tf32 (fc1, every backward contraction) / fp16 (attention forward) tensor-core operands, "
                         "f32 accumulate" if engine == "tc" else "f32",
                "data": "synthetic", "gpu_launches": launches, "loss": float(loss.sum().item()),
                "allreduce": {"ms_per_step": ar if world > 1 else 0.0, "bytes": n_grad * 4, "backend": "nccl" if world > 1 else None,
                              "calls_per_step": 1 if world > 1 else 0},
                "config": {"workload": f"train step (forward + backward + {opt_name}), {bags} bags x {patches} x {features} per GPU, "
                                       f"n_classes={classes}, CE one-hot, to_out Dropout(0.7) active",
                           "parallelism": f"bag-dp{world}"}}
    finally:
        L.tm_bgemm_engine(prev_engine)
        T.training.FORWARD_ENGINE = prev_fwd


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bags", type=int, default=64)       # reference default: 64 bags of 1000 patches (feature_dataloader.py:27,388-401)
    ap.add_argument("--patches", type=int, default=1000)
    ap.add_argument("--features", type=int, default=2048)
    ap.add_argument("--classes", type=int, default=2)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--opt", default="lookahead_radam")
    ap.add_argument("--engine", default="tc", choices=["tc", "simt"])   # tcgen05 forward + backward contractions, or the fp32 cross-check
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    if world > 1:
        dist.init_process_group("nccl")
    line = measure(a.bags, a.patches, a.features, a.classes, a.steps, a.warmup, a.opt, a.engine, rank, world)
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
