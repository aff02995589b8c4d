#!/usr/bin/env python
"""bench.py -- TransMIL bag-level inference throughput on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W [--impl reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one pass of the hot path (TransMIL.forward, TransMIL.py:167-211) over one batch of
synthetic bags per GPU: by default ONE 50,000 x 1024 bag (BASELINE config 3, the size the
"ms per 50k-patch bag" metric is quoted on); whole bags are data-parallel across ranks with no
data-path collective (weak scaling).  Rank 0 prints ONE JSON line.

  value     slides/s with the bag already resident in HBM (CUDA events, max over ranks)
  e2e       slides/s through the public API (`model(x_pinned_host)` + logits.cpu()): H2D of the
            bag and D2H of the logits inside the timed region
  roofline  the dominant kernel stage of the step, timed live with CUDA events on the launch
            stream inside the timed region (tm_profile_*), algorithmic FLOPs/bytes from
            SURVEY.md 8(d) / DESIGN.md, peak from MEASURED_PEAKS.json
  cpu_baseline  the CPU oracle (a port of the reference forward: the reference module itself
            cannot travel to the GPU box) on the host cores, bounded sample of the same workload

--impl reference times that CPU oracle for K steps of the same workload (rank 0 only).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

if os.environ.get("NCCL_DEBUG", "").upper() not in ("INFO", "TRACE"):
    os.environ["NCCL_DEBUG"] = "WARN"   # keep NCCL's version banner off stdout: rank 0 prints exactly one JSON line

import torch  # noqa: E402

METRIC = "transmil_inference_slides_per_s"
UNIT = "slides/s"
WEIGHT_SEED = 2021


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--patches", dest="n", type=int, default=50000,
                    help="patches per bag (use --patches under torchrun: its parser treats --n as an abbreviation of its own options)")
    ap.add_argument("--features", type=int, default=1024, help="in_features D")
    ap.add_argument("--bags", type=int, default=1, help="bags per step per GPU")
    ap.add_argument("--classes", type=int, default=2)
    ap.add_argument("--simt", action="store_true", help="force the fp32 CUDA-core cross-check engine")
    ap.add_argument("--shard", default="bags", choices=["bags", "tokens"],
                    help="bags: whole bags data-parallel across ranks (default, weak scaling); tokens: ONE bag per step, "
                         "token-sharded across all ranks (BASELINE config 4, strong scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--clock-hold", type=float, default=1.5, help="seconds of extra identical steps run under the clock sampler")
    return ap.parse_args()


def workload_config(a, world):
    return {"workload": f"BASELINE config 3: {a.bags} synthetic bag(s) of {a.n} x {a.features} fp32 patch features per step per GPU, "
                        f"n_classes={a.classes}, eval forward (TransMIL.py:167-211)",
            "n": a.n, "in_features": a.features, "bags_per_step_per_gpu": a.bags, "n_classes": a.classes,
            "parallelism": f"bag-dp{world}", "weights": f"default init, torch.manual_seed({WEIGHT_SEED})",
            "inputs": "torch.rand (code/sustainability_test.py:45), seed 1234+rank",
            "l2": "inputs larger than L2: every step streams the bag plus >1 GB of activations (126 MB L2)"}


def make_bag(a, rank):
    g = torch.Generator().manual_seed(1234 + rank)
    return torch.rand(a.bags, a.n, a.features, generator=g)


# ---- algorithmic cost model (DESIGN.md "Kernels") -------------------
def stage_cost(a):
    from oracle.transmil_oracle import plan
    p = plan(a.n)
    B, n, D, S, Sp = a.bags, a.n, a.features, p["S"], p["Sp"]
    F0 = {2048: 1024, 768: 768}.get(D, 0)
    c = {}
    if F0:
        c["fc1_gemm0"] = ("tensor", 2.0 * B * n * D * F0)
        c["fc1_ln"] = ("hbm", 2.0 * B * n * F0 * 4)
        c["fc1_gemm1"] = ("tensor", 2.0 * B * n * F0 * 512)
    else:
        c["fc1_gemm1"] = ("tensor", 2.0 * B * n * D * 512)
    c["ln_pad"] = ("hbm", 1.0 * B * (S + Sp) * 512 * 4)
    c["qkv"] = ("tensor", 2.0 * B * Sp * 512 * 1536)
    c["landmarks"] = ("hbm", 2.0 * B * Sp * 512 * 4)
    c["pinv"] = ("tensor", B * 8 * (2.0 * 256 * 256 * 64 + 6 * 4 * 2.0 * 256 ** 3))
    c["attn3v"] = ("tensor", 2 * (2.0 * B * Sp * 256 * 512))
    c["zw"] = ("tensor", B * 8 * 2.0 * 256 * 256 * 64)
    c["nystrom_out"] = ("tensor", 2 * (2.0 * B * Sp * 256 * 512) + 2.0 * 33 * B * Sp * 512)
    c["to_out"] = ("tensor", 2.0 * B * S * 512 * 512)
    c["ppeg"] = ("hbm", 2.0 * B * S * 512 * 4)
    c["head"] = ("hbm", 512.0 * 4 * B)
    c["fc1_tail"] = ("hbm", 512.0 * 4 * B)
    return c


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(path):
        with open(path) as f:
            m = json.load(f)
        return {"hbm": m["hbm_gbs"], "tensor": m["bf16_tflops_sustained"], "src": "measured (MEASURED_PEAKS.json: hbm_gbs, bf16_tflops_sustained)"}
    return {"hbm": 6650.0, "tensor": 1590.0, "src": "fallback (B200_PROFILING.md)"}


def ncu_traffic(a, stage):
    """dram bytes per launch of `stage` from the committed `ncu --set full` capture of this workload (None for other workloads)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")) as f:
            t = json.load(f)
        w = t["workload"]
        if (w["n"], w["in_features"], w["bags"]) != (a.n, a.features, a.bags):
            return None
        return t["stages"][stage]["dram_bytes_per_launch"]
    except Exception:
        return None


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "20", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([s.strip() for s in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---- the CPU arm: oracle port of the reference forward -------------------------------------
def cpu_state_dict(a):
    """Reference-layout weights from the seed.  The constructor only builds nn parameter containers
    (no CUDA, no kernels), in the reference's creation order, so the init equals the reference's."""
    import transmil_b200 as T
    torch.manual_seed(WEIGHT_SEED)
    return T.TransMIL(n_classes=a.classes, in_features=a.features).state_dict()


def run_reference_arm(a, rank, world):
    if rank != 0:
        return
    from oracle import transmil_oracle as O
    cores = os.cpu_count()
    torch.set_num_threads(cores)
    torch.set_float32_matmul_precision("highest")
    sd = cpu_state_dict(a)
    x = make_bag(a, 0)
    for _ in range(max(1, min(a.warmup, 1))):
        O.forward(sd, x)
    t0 = time.perf_counter()
    for _ in range(a.steps):
        logits = O.forward(sd, x)
    dt = time.perf_counter() - t0
    val = a.steps * a.bags / dt
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
            "ms_per_step": 1e3 * dt / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic", "config": workload_config(a, world),
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                             "sample": f"{a.steps} steps x {a.bags} bag(s) of {a.n}x{a.features} through oracle/transmil_oracle.py "
                                       f"(CPU port of the reference forward, dense attn skipped: logits identical), 1 warm-up"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "logits": logits.tolist()[0]}
    print(json.dumps(line), flush=True)


def cpu_baseline(a):
    from oracle import transmil_oracle as O
    cores = os.cpu_count()
    torch.set_num_threads(cores)
    torch.set_float32_matmul_precision("highest")
    sd = cpu_state_dict(a)
    x = make_bag(a, 0)
    t0 = time.perf_counter()
    ref = O.forward(sd, x)                       # warm-up, also the parity reference of this run
    t_warm = time.perf_counter() - t0
    reps = max(1, min(5, int(20.0 / max(t_warm, 1e-3))))
    t0 = time.perf_counter()
    for _ in range(reps):
        O.forward(sd, x)
    dt = (time.perf_counter() - t0) / reps
    return ref, {"value": a.bags / dt, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_bag": 1e3 * dt / a.bags,
                 "sample": f"{reps} pass(es) over the bench bag ({a.bags}x{a.n}x{a.features}) through oracle/transmil_oracle.py after 1 warm-up"}


def run_token_sharded(a, rank, local_rank, world, dev, model, barrier):
    """BASELINE config 4: one bag per step, token-sharded over all ranks (transmil-deepgraft_b200/sharded.py).  Each rank
    holds only its rows of the bag (device-resident for `value`; pinned host rows copied in for `e2e`)."""
    import torch.distributed as dist
    import transmil_b200 as T
    from transmil_b200.sharded import TokenShardedTransMIL
    if a.bags != 1:
        raise SystemExit("--shard tokens runs one bag per step")
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    L = T.capi.lib()
    sharded = TokenShardedTransMIL(model)
    plan = sharded.plan(a.n)
    x_host = plan.take(make_bag(a, 0), rank).contiguous().pin_memory()     # the same bag on every rank, this rank's rows
    x_dev = x_host.to(dev)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    for _ in range(max(a.warmup, 3)):
        logits = sharded(x_dev, n=a.n)
    barrier()
    L.tm_reset_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(a.steps):
        logits = sharded(x_dev, n=a.n)
    ev1.record()
    barrier()
    ms = ev0.elapsed_time(ev1)
    launches = int(L.tm_launch_count())
    for _ in range(2):
        sharded(x_host, n=a.n).cpu()
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        sharded(x_host, n=a.n).cpu()
    e1.record()
    barrier()
    e2e_ms = e0.elapsed_time(e1)
    clocks = None
    if sampler:
        clocks = sampler.stop()
    t = torch.tensor([ms, e2e_ms], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms = float(t[0]), float(t[1])
    with torch.no_grad():
        plain = model(make_bag(a, 0).to(dev)) if rank == 0 else None
    if rank == 0:
        value = a.steps / (ms / 1e3)
        cfg = workload_config(a, world)
        cfg["workload"] = (f"BASELINE config 4: ONE synthetic bag of {a.n} x {a.features} fp32 patch features per step, token-sharded over "
                           f"{world} GPU(s) (landmark allreduce + attn3.v allgather/merge + conv and PPEG halos), n_classes={a.classes}, eval forward")
        cfg["parallelism"] = f"token-shard{world}"
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32 storage; tf32 (fc1) / fp16 tensor-core contractions with f32 accumulate", "data": "synthetic", "config": cfg,
                "tokens_per_s": value * a.n, "ms_per_bag": ms / a.steps, "gpu_launches": launches, "clocks": clocks,
                "e2e": {"value": a.steps / (e2e_ms / 1e3), "unit": UNIT, "h2d_bytes_per_step": int(x_host.numel() * 4 * world),
                        "d2h_bytes_per_step": a.classes * 4 * world, "ms_per_step": e2e_ms / a.steps,
                        "api": "TokenShardedTransMIL(model)(x_rows_pinned_host, n) -> logits.cpu() on every rank"},
                "parity": {"max_abs_dlogit_vs_unsharded": float((logits - plain).abs().max())}}
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


def main():
    a = parse()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if a.impl == "reference":
        run_reference_arm(a, rank, world)
        return
    import torch.distributed as dist
    import transmil_b200 as T
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU path)"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    L = T.capi.lib()

    torch.manual_seed(WEIGHT_SEED)
    model = T.TransMIL(n_classes=a.classes, in_features=a.features).eval().to(dev)
    if a.simt:
        model.flags = T.capi.TM_FLAG_SIMT_ONLY

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    if a.shard == "tokens":
        return run_token_sharded(a, rank, local_rank, world, dev, model, barrier)
    x_host = make_bag(a, rank).pin_memory()
    x_dev = x_host.to(dev)

    with torch.no_grad():
        # nvidia-smi samples every 20 ms from the warm-up through both timed regions; a K-step region of ~1 ms steps is
        # shorter than one sample period, so the same step also runs for a further `--clock-hold` seconds (untimed) after
        # the timed regions with the sampler still attached ("hold_s" in the clocks object).
        sampler = ClockSampler(local_rank) if rank == 0 else None
        for _ in range(max(a.warmup, 3)):
            logits = model(x_dev)
        barrier()
        # ---- timed region 1: device-resident bag -------------------------------------------
        L.tm_reset_launch_count()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record()
        for _ in range(a.steps):
            logits = model(x_dev)
        ev1.record()
        barrier()
        ms = ev0.elapsed_time(ev1)
        launches = int(L.tm_launch_count())
        # ---- stage pass: the same K steps again with a CUDA event on the launch stream at every stage boundary.  The
        # events need every stage on one stream, so the library runs the pinv chain and attn3.v back to back here
        # (in the timed region above they overlap on two streams); `stage_pass_ms_per_step` shows the difference.
        L.tm_profile_enable(1)
        T.capi.profile_read()
        pv0, pv1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        pv0.record()
        for _ in range(a.steps):
            model(x_dev)
        pv1.record()
        barrier()
        stage_ms = pv0.elapsed_time(pv1)
        stages = T.capi.profile_read()
        L.tm_profile_enable(0)
        # ---- timed region 2: end to end from pinned host memory ----------------------------
        e2e_ms = e2e_serial_ms = None
        if not a.no_e2e:
            for _ in range(2):
                model(x_host).cpu()
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                out_host = model(x_host).cpu()
            e1.record()
            barrier()
            e2e_serial_ms = e0.elapsed_time(e1)
            # the same K slides through the slide-loop API: the copy of slide i+1 overlaps the forward of slide i
            loop = T.StreamedInference(model)
            for out_host in loop([x_host] * 3):
                pass
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for out_host in loop(x_host for _ in range(a.steps)):
                pass
            e1.record()
            barrier()
            e2e_ms = e0.elapsed_time(e1)
            # the same slides as half-precision feature files (fp16 on the host, read as fp16 by fc1)
            x16_host = x_host.half().pin_memory()
            x16_dev = x16_host.to(dev)
            ref32 = model(x16_dev.float())
            for _ in range(3):
                got16 = model(x16_dev)
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.steps):
                model(x16_dev)
            e1.record()
            barrier()
            f16_dev_ms = e0.elapsed_time(e1)
            for out_host in loop([x16_host] * 3):
                pass
            barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for out_host in loop(x16_host for _ in range(a.steps)):
                pass
            e1.record()
            barrier()
            f16_e2e_ms = e0.elapsed_time(e1)
            f16 = {"note": "same bag stored as fp16 feature files (SURVEY 8f/N2): fp16 on the host and in HBM, fc1 contracts fp16 x fp16 -> f32; "
                           "not the headline (the reference feeds fp32 tensors)",
                   "device_ms_per_step": f16_dev_ms / a.steps, "e2e_ms_per_step": f16_e2e_ms / a.steps,
                   "e2e_value": a.steps * a.bags / (f16_e2e_ms / 1e3), "h2d_bytes_per_step": a.bags * a.n * a.features * 2,
                   "max_abs_dlogit_vs_f32_read_of_the_same_halves": float((got16 - ref32).abs().max())}
            del x16_dev
        clocks = None
        if sampler:
            t_hold = time.perf_counter()
            while time.perf_counter() - t_hold < a.clock_hold:
                for _ in range(20):
                    model(x_dev)
                torch.cuda.synchronize(dev)
            clocks = sampler.stop()
            clocks["hold_s"] = a.clock_hold
    t = torch.tensor([ms, e2e_ms if e2e_ms is not None else 0.0, e2e_serial_ms if e2e_ms is not None else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_ms_max, e2e_serial_max = float(t[0]), float(t[1]), float(t[2])

    if rank == 0:
        total_bags = a.steps * a.bags * world
        value = total_bags / (ms / 1e3)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": ms / a.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32 storage; tf32 (fc1) / fp16 tensor-core contractions with f32 accumulate" if not a.simt else "f32",
                "data": "synthetic", "config": workload_config(a, world),
                "tokens_per_s": value * a.n, "ms_per_bag": ms / (a.steps * a.bags),
                "gpu_launches": launches, "clocks": clocks}
        if e2e_ms is not None:
            line["e2e"] = {"value": total_bags / (e2e_ms_max / 1e3), "unit": UNIT,
                           "h2d_bytes_per_step": a.bags * a.n * a.features * 4, "d2h_bytes_per_step": a.bags * a.classes * 4,
                           "ms_per_step": e2e_ms_max / a.steps,
                           "api": "StreamedInference(model)(pinned host bags) -> host logits, one slide at a time: every step copies "
                                  "its 50k x D bag host->device and its logits back; the copy of slide i+1 overlaps the forward of slide i",
                           "serial": {"value": total_bags / (e2e_serial_max / 1e3), "ms_per_step": e2e_serial_max / a.steps,
                                      "api": "model(x_pinned_host).cpu() per step, copy then compute"},
                           "f16_features": f16}
        # roofline of the dominant stage
        cost = stage_cost(a)
        pk = peaks()
        per_stage = {k: {"ms_per_launch": v[0] / v[1], "launches_per_step": v[1] / a.steps, "share": v[0] / stage_ms} for k, v in stages.items()}
        line["stage_pass_ms_per_step"] = stage_ms / a.steps
        if per_stage:
            top = max(per_stage, key=lambda k: per_stage[k]["share"])
            bound, alg = cost[top]
            dur_s = per_stage[top]["ms_per_launch"] / 1e3
            achieved = alg / dur_s / (1e12 if bound == "tensor" else 1e9)
            line["roofline"] = {"kernel": top, "bound": bound, "achieved": achieved, "peak": pk[bound],
                                "unit": "TFLOP/s" if bound == "tensor" else "GB/s", "frac": achieved / pk[bound],
                                "traffic": ncu_traffic(a, top), "peak_source": pk["src"], "algorithmic_per_launch": alg,
                                "ms_per_launch": per_stage[top]["ms_per_launch"],
                                "timing": "CUDA events on the launch stream at stage boundaries, K steps of the same workload run "
                                          "right after the timed region with the two-stream pinv/attn3v overlap off"}
            detail = {}
            for k, v in per_stage.items():
                b, alg_k = cost[k]
                ach = alg_k / (v["ms_per_launch"] / 1e3) / (1e12 if b == "tensor" else 1e9)
                detail[k] = {"ms_per_launch": round(v["ms_per_launch"], 5), "share": round(v["share"], 4), "bound": b,
                             "achieved": round(ach, 2), "frac": round(ach / pk[b], 4)}
            line["stages"] = detail
        if not a.no_cpu_baseline and world == 1:
            ref, cb = cpu_baseline(a)
            line["cpu_baseline"] = cb
            got = logits.detach().cpu()
            line["parity"] = {"max_abs_dlogit": float((got - ref).abs().max()),
                              "max_abs_dprob": float((torch.softmax(got, 1) - torch.softmax(ref, 1)).abs().max())}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
