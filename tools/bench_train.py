"""Times one TRAINING step of the drop-in TransMIL (BASELINE config 2: DeepGraft/TransMIL_feat_norm_rest.yaml -- 2048-d bags,
RAdam lr 1e-4 wd 0.01, CE against the one-hot label) on one B200, or bag-data-parallel under torchrun (gradients summed with
one allreduce per step).  Prints one JSON line (same keys as bench.py where they apply).  Not part of the default bench: the
headline metric of bench.py stays the 50 k-patch inference bag.

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
    torch.manual_seed(2021)
    model = T.TransMIL(n_classes=a.classes, in_features=a.features).cuda().train()
    opt = optim.create_optimizer(SimpleNamespace(opt=a.opt, lr=1e-4, weight_decay=0.01), model)
    g = torch.Generator().manual_seed(1000 + rank)
    x = torch.rand(a.bags, a.patches, a.features, generator=g).cuda()
    y = torch.randint(0, a.classes, (a.bags,), generator=g).cuda()
    L = T.capi.lib()
    if a.engine == "simt":
        L.tm_bgemm_engine(1)
        T.training.FORWARD_ENGINE = "simt"

    def one_step():
        opt.zero_grad(set_to_none=True)
        loss, _, _ = plugin.training_step(model, (x, y, None))
        loss.sum().backward()
        optim.allreduce_gradients(model.parameters())
        opt.step()
        return loss

    for _ in range(a.warmup):
        one_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    L.tm_reset_launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        loss = one_step()
    e1.record()
    torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / a.steps], device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = ms.item()
        print(json.dumps({"metric": "training bags/s", "engine": a.engine, "value": a.bags * world / (ms / 1e3), "unit": "bags/s",
                          "tokens_per_s": a.bags * a.patches * world / (ms / 1e3), "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                          "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "dtype": "f32 storage; tf32 / fp16 tensor-core contractions, f32 accumulate" if a.engine == "tc" else "f32", "data": "synthetic",
                          "gpu_launches": int(L.tm_launch_count()), "loss": float(loss.sum().item()),
                          "config": {"workload": f"train step, {a.bags} bags x {a.patches} x {a.features} per GPU, {a.opt}, CE one-hot",
                                     "parallelism": f"bag-dp{world}"}}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
