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
    ap.add_argument("--no-extras", action="store_true", help="skip the sub-objects torch_eager / train_step / cfg5 / token_shard")
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
    """stage -> (algorithmic FLOPs, algorithmic HBM bytes) of ONE launch.  Bytes use the element sizes the tensor-core engine
    really stores: fp32 for the bag, the residual stream h and the fc1 scratch; fp16 for xn, q, k, v, merged (SIMT engine: fp32
    everywhere).  Minimal association, halo re-reads and weights not credited."""
    from oracle.transmil_oracle import plan
    p = plan(a.n)
    B, n, D, S, Sp = a.bags, a.n, a.features, p["S"], p["Sp"]
    F0 = {2048: 1024, 768: 768}.get(D, 0)
    e16 = 4 if a.simt else 2
    c = {}
    if F0:
        c["fc1_gemm0"] = (2.0 * B * n * D * F0, 4.0 * B * n * (D + F0))
        c["fc1_ln"] = (0.0, 2.0 * B * n * F0 * 4)
        c["fc1_gemm1"] = (2.0 * B * n * F0 * 512, 4.0 * B * (n * F0 + S * 512))
    else:
        c["fc1_gemm1"] = (2.0 * B * n * D * 512, 4.0 * B * (n * D + S * 512))
    c["ln_pad"] = (0.0, 1.0 * B * 512 * (S * 4 + Sp * e16))
    c["qkv"] = (2.0 * B * Sp * 512 * 1536, 1.0 * B * Sp * 512 * e16 * 4)
    c["landmarks"] = (0.0, 2.0 * B * Sp * 512 * e16)
    c["pinv"] = (B * 8 * (2.0 * 256 * 256 * 64 + 6 * 4 * 2.0 * 256 ** 3), B * 8 * (2 * 256 * 64 + 2 * 256 * 256) * 4.0)
    c["attn3v"] = (2 * (2.0 * B * Sp * 256 * 512), 2.0 * B * Sp * 512 * e16)
    c["zw"] = (B * 8 * 2.0 * 256 * 256 * 64, B * 8 * (256 * 256 + 2 * 256 * 64) * 4.0)
    c["nystrom_out"] = (2 * (2.0 * B * Sp * 256 * 512) + 2.0 * 33 * B * Sp * 512, 3.0 * B * Sp * 512 * e16)
    c["to_out"] = (2.0 * B * S * 512 * 512, 1.0 * B * 512 * (Sp * e16 + 2 * S * 4))
    c["ppeg"] = (2.0 * 49 * B * S * 512, 2.0 * B * S * 512 * 4)
    c["head"] = (2.0 * 512 * a.classes * B, 512.0 * 4 * B)
    c["fc1_tail"] = (0.0, 512.0 * 4 * B)
    return c


def stage_roofline(cost, ms_per_launch, pk):
    """Both roofline fractions of one stage; `bound` = the roof that sets the larger ideal time."""
    flops, nbytes = cost
    t = ms_per_launch / 1e3
    tf, gb = flops / t / 1e12, nbytes / t / 1e9
    bound = "tensor" if flops / (pk["tensor"] * 1e12) >= nbytes / (pk["hbm"] * 1e9) else "hbm"
    return {"bound": bound, "achieved": tf if bound == "tensor" else gb, "unit": "TFLOP/s" if bound == "tensor" else "GB/s",
            "frac": (tf / pk["tensor"]) if bound == "tensor" else (gb / pk["hbm"]),
            "tflops": tf, "frac_tensor": tf / pk["tensor"], "gbs": gb, "frac_hbm": gb / pk["hbm"]}


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
        path = os.path.join(ROOT, "profiles", "r2_ncu_traffic.json")
        if not os.path.isfile(path):
            path = os.path.join(ROOT, "profiles", "r1_ncu_traffic.json")
        with open(path) as f:
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
    out = {"value": a.bags / dt, "unit": UNIT, "cores": cores, "kind": "port", "ms_per_bag": 1e3 * dt / a.bags,
           "sample": f"{reps} pass(es) over the bench bag ({a.bags}x{a.n}x{a.features}) through oracle/transmil_oracle.py after 1 warm-up"}
    if cores != 8:     # the reference's own setting: torch.set_num_threads(8), code/train.py:93
        torch.set_num_threads(8)
        O.forward(sd, x)
        reps8 = max(1, min(3, int(10.0 / max(dt, 1e-3))))
        t0 = time.perf_counter()
        for _ in range(reps8):
            O.forward(sd, x)
        dt8 = (time.perf_counter() - t0) / reps8
        out["threads_8"] = {"value": a.bags / dt8, "ms_per_bag": 1e3 * dt8 / a.bags, "cores": 8,
                            "note": "torch.set_num_threads(8), the reference's own setting (code/train.py:93)"}
        torch.set_num_threads(cores)
    return ref, out


def token_shard_measure(n, features, classes, steps, warmup, rank, world, dev, barrier, with_oracle=True, with_e2e=True):
    """BASELINE config 4: ONE bag per step, token-sharded over all ranks (transmil-deepgraft_b200/sharded.py).  Every rank
    holds only its rows of the bag.  Every rank of the group calls this; returns the result dict on rank 0 (None elsewhere).
    world == 1 gives the N = 1 point of the strong-scaling curve: the ordinary whole-graph forward of the same bag."""
    import torch.distributed as dist
    import transmil_b200 as T
    from transmil_b200.sharded import TokenShardedTransMIL
    L = T.capi.lib()
    torch.manual_seed(WEIGHT_SEED)
    model = T.TransMIL(n_classes=classes, in_features=features).eval().to(dev)
    g = torch.Generator().manual_seed(4321)
    x_full = torch.rand(1, n, features, generator=g)          # the same bag on every rank
    res = {"workload": f"BASELINE config 4: ONE synthetic bag of {n} x {features} fp32 patch features per step, token-sharded over {world} "
                       "GPU(s) (landmark sums + attn3.v partials all-to-all, conv and PPEG halos to the neighbours), eval forward",
           "n": n, "in_features": features, "n_gpus": world, "scaling": "strong", "steps": steps}
    with torch.no_grad():
        # single-GPU time of the same bag (rank 0): the denominator of the speed-up and the unsharded parity reference
        plain = None
        single_ms = 0.0
        if rank == 0:
            xd = x_full.to(dev)
            for _ in range(max(warmup, 3)):
                plain = model(xd)
            torch.cuda.synchronize(dev)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                plain = model(xd)
            e1.record()
            torch.cuda.synchronize(dev)
            single_ms = e0.elapsed_time(e1) / steps
            del xd
        barrier()      # the other ranks stay idle (no host pinning, no collectives spinning) while rank 0 times the single-GPU bag
        if world > 1:
            def timed(sharded, xin, sync_host=False):
                for _ in range(max(warmup, 3)):
                    out = sharded(xin, n=n)
                    if sync_host:
                        out.cpu()
                barrier()
                L.tm_reset_launch_count()
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                ev0.record()
                for _ in range(steps):
                    out = sharded(xin, n=n)
                    if sync_host:
                        out.cpu()
                ev1.record()
                barrier()
                t = torch.tensor([ev0.elapsed_time(ev1) / steps], device=dev, dtype=torch.float64)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                return float(t[0]), int(L.tm_launch_count()) / steps, out

            # three ways to run the same sharded bag: the library call replayed from a CUDA graph (headline), the same call
            # issued eagerly, and the Python-issued schedule over NCCL (round-1 form, kept as the cross-check)
            via_nccl = TokenShardedTransMIL(model, transport="nccl")
            x_host = via_nccl.plan(n).take(x_full, rank).contiguous().pin_memory()
            x_dev = x_host.to(dev)
            ms_nccl, launches_nccl, logits_nccl = timed(via_nccl, x_dev)
            nr = torch.ones(1, device=dev)
            dist.all_reduce(nr)
            res["comm_nranks_seen"] = int(nr.item())
            try:
                graphed = TokenShardedTransMIL(model, cuda_graph=True, static_input=True)   # replays read x_dev in place
                eager = TokenShardedTransMIL(model)
            except T.capi.TransmilError as e:          # raised on every rank together (PeerComm): report the NCCL form alone
                graphed = None
                res.update({"peer_transport_error": str(e)[:300], "ms_per_step": ms_nccl, "launches_per_step_rank0": launches_nccl,
                            "transport": "NCCL through torch.distributed, schedule issued from Python (peer-memory transport unavailable)"})
                logits = logits_nccl
            if graphed is not None:
                ms, launches, logits = timed(graphed, x_dev)
                ms_eager, launches_eager, logits_eager = timed(eager, x_dev)
                status = graphed.comm.status() | eager.comm.status()
                res.update({"ms_per_step": ms, "launches_per_step_rank0": launches_eager, "graph_replays_per_step": 1,
                            "transport": "tm_transmil_fwd_sharded: one C call per rank and bag, captured into a CUDA graph; exchanges are "
                                         "flag-guarded pull kernels over NVLink peer memory (CUDA IPC), no collective library on the data path",
                            "exchange_timeouts": status,
                            "variants_ms_per_step": {"peer_graph": ms, "peer_eager": ms_eager, "nccl_python_schedule": ms_nccl},
                            "variants_launches_per_step_rank0": {"peer_eager": launches_eager, "nccl_python_schedule": launches_nccl},
                            "max_abs_dlogit_peer_vs_nccl": float((logits - logits_nccl).abs().max()),
                            "graph_equals_eager_bits": bool(torch.equal(logits, logits_eager))})
            if with_e2e:
                runner = graphed if graphed is not None else via_nccl
                e2e_ms, _, _ = timed(runner, x_host, sync_host=True)
                res["e2e"] = {"ms_per_step": e2e_ms, "h2d_bytes_per_step": int(x_host.numel() * 4 * world), "d2h_bytes_per_step": classes * 4 * world,
                              "api": "TokenShardedTransMIL(model, cuda_graph=True)(x_rows_pinned_host, n) -> logits.cpu() on every rank"}
            barrier()
        else:
            logits = plain
            res.update({"ms_per_step": single_ms, "comm_nranks_seen": 1})
    if rank != 0:
        return None
    res["single_gpu_ms_per_step"] = single_ms
    res["speedup_vs_single_gpu"] = single_ms / res["ms_per_step"] if res["ms_per_step"] > 0 else None
    res["tokens_per_s"] = n / (res["ms_per_step"] / 1e3)
    par = {"max_abs_dlogit_vs_unsharded": float((logits - plain).abs().max())}
    if with_oracle:
        from oracle import transmil_oracle as O
        torch.set_num_threads(os.cpu_count())
        torch.set_float32_matmul_precision("highest")
        ref = O.forward({k: v.detach().cpu() for k, v in model.state_dict().items()}, x_full)
        got = logits.detach().cpu()
        par["max_abs_dlogit_vs_oracle"] = float((got - ref).abs().max())
        par["max_abs_dprob_vs_oracle"] = float((torch.softmax(got, 1) - torch.softmax(ref, 1)).abs().max())
    res["parity"] = par
    return res


def run_token_sharded(a, rank, local_rank, world, dev, model, barrier):
    """`--shard tokens`: the config-4 measurement as the headline of the line (`"scaling": "strong"`)."""
    import torch.distributed as dist
    if a.bags != 1:
        raise SystemExit("--shard tokens runs one bag per step")
    if not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29577")
        dist.init_process_group("nccl", rank=0, world_size=1, device_id=dev)
    sampler = ClockSampler(local_rank) if rank == 0 else None
    r = token_shard_measure(a.n, a.features, a.classes, a.steps, a.warmup, rank, world, dev, barrier, with_oracle=not a.no_cpu_baseline)
    clocks = sampler.stop() if sampler else None
    if rank == 0:
        cfg = workload_config(a, world)
        cfg["workload"] = r["workload"]
        cfg["parallelism"] = f"token-shard{world}"
        line = {"metric": METRIC, "value": 1e3 / r["ms_per_step"], "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": max(a.warmup, 3),
                "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32 storage; tf32 (fc1) / fp16 tensor-core contractions with f32 accumulate", "data": "synthetic", "config": cfg,
                "tokens_per_s": r["tokens_per_s"], "ms_per_bag": r["ms_per_step"], "gpu_launches": int(r.get("launches_per_step_rank0", 0) * a.steps),
                "clocks": clocks, "parity": r["parity"], "token_shard": r}
        if "e2e" in r:
            line["e2e"] = dict(r["e2e"], value=1e3 / r["e2e"]["ms_per_step"], unit=UNIT)
        print(json.dumps(line), flush=True)
    dist.destroy_process_group()


def run_extras(a, rank, world, dev, barrier, model, x_dev, oracle_logits):
    """Sub-objects of the JSON line for the BASELINE configs the headline does not cover.  Run AFTER every
    timed region of the headline; each part is independent and a failure is recorded, not raised.  Every rank calls this (the
    training allreduce, config 5's gather and the token-sharded bag are collective)."""
    out = {}
    tools = os.path.join(ROOT, "tools")
    if tools not in sys.path:
        sys.path.insert(0, tools)

    def part(name, fn, collective=True):
        try:
            if collective or rank == 0:
                r = fn()
                if rank == 0 and r is not None:
                    out[name] = r
        except Exception as e:                      # noqa: BLE001 -- a measurement tool must not take the headline down
            if rank == 0:
                out[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
        barrier()

    # D6: stock torch eager on the same bag (rank 0 only; the other ranks wait at the barrier)
    def eager():
        import bench_torch_eager
        with torch.no_grad():
            return bench_torch_eager.measure(model, x_dev, steps=5, warmup=2, oracle_logits=oracle_logits)
    part("torch_eager", eager, collective=False)

    # E4: the same bag with the second layer's out-projection restricted to the cls row's tiles (model.skip_unread_rows).  A
    # separate number: it does FEWER flops than the headline (the reference computes those rows), same logits bit for bit.
    def e4():
        with torch.no_grad():
            full = model(x_dev)
            model.skip_unread_rows = True
            try:
                for _ in range(3):
                    got = model(x_dev)
                torch.cuda.synchronize(dev)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(20):
                    got = model(x_dev)
                e1.record()
                torch.cuda.synchronize(dev)
            finally:
                model.skip_unread_rows = False
        ms = e0.elapsed_time(e1) / 20
        return {"ms_per_step": ms, "value": x_dev.shape[0] / (ms / 1e3), "unit": UNIT, "logits_bit_identical_to_full_path": bool(torch.equal(full, got)),
                "note": "TM_FLAG_CLS_ONLY_L2: layer-2 nystrom_out / to_out run for the cls row's 128-row tiles only (TransMIL.py:197-204 reads "
                        "h[:, 0] only); q, k, v, landmarks, pinv, attn3.v complete.  Fewer flops than the headline step: not comparable with it"}
    part("e4_cls_only", e4, collective=False)

    # config 2: training step (reference-default batch and one 20 000-patch bag), gradient allreduce over NCCL at N > 1
    def train():
        import bench_train
        r = {}
        for key, (bags, patches) in (("batch64x1000x2048", (64, 1000)), ("bag1x20000x2048", (1, 20000))):
            r[key] = bench_train.measure(bags, patches, 2048, 2, steps=5, warmup=3, rank=rank, world=world, dev=dev)
            torch.cuda.empty_cache()
        return r
    part("train_step", train)

    # config 5: 256 variable-length bags, n_classes = 6, whole bags placed longest-first
    def cfg5():
        import bench_cfg5
        return bench_cfg5.measure(256, 2048, 6, 5, rank=rank, world=world, with_parity=True)
    part("cfg5", cfg5)

    # config 4: one 200 000 x 2048 bag token-sharded over all ranks (N = 1: the single-GPU point of the curve)
    def shard():
        torch.cuda.empty_cache()
        return token_shard_measure(200000, 2048, 2, steps=5, warmup=3, rank=rank, world=world, dev=dev, barrier=barrier,
                                   with_oracle=True, with_e2e=False)
    part("token_shard", shard)
    return out


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

    # parity of THIS run's logits against the CPU oracle (rank 0's bag; at every world size), and the CPU baseline (N = 1 only)
    ref = cb = None
    if rank == 0 and not a.no_cpu_baseline:
        if world == 1:
            ref, cb = cpu_baseline(a)
        else:
            from oracle import transmil_oracle as O
            torch.set_num_threads(os.cpu_count())
            torch.set_float32_matmul_precision("highest")
            ref = O.forward(cpu_state_dict(a), make_bag(a, 0))
    extras = {}
    if not a.no_extras:
        extras = run_extras(a, rank, world, dev, barrier, model, x_dev, ref)

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
            h2d = a.bags * a.n * a.features * 4
            line["e2e"] = {"value": total_bags / (e2e_ms_max / 1e3), "unit": UNIT,
                           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": a.bags * a.classes * 4,
                           "ms_per_step": e2e_ms_max / a.steps,
                           "h2d_gbs_per_gpu": h2d / (e2e_ms_max / a.steps / 1e3) / 1e9,
                           "bound": "host->device copy (PCIe): the copy of slide i+1 runs under the forward of slide i, a step costs "
                                    "max(copy, forward) = the copy; with N GPUs the host side of the copies is shared",
                           "api": "StreamedInference(model)(pinned host bags) -> host logits, one slide at a time: every step copies "
                                  "its 50k x D bag host->device and its logits back; the copy of slide i+1 overlaps the forward of slide i",
                           "serial": {"value": total_bags / (e2e_serial_max / 1e3), "ms_per_step": e2e_serial_max / a.steps,
                                      "api": "model(x_pinned_host).cpu() per step, copy then compute"},
                           "f16_features": f16}
        # roofline of the dominant stage; every stage with both fractions
        cost = stage_cost(a)
        pk = peaks()
        per_stage = {k: {"ms_per_launch": v[0] / v[1], "launches_per_step": v[1] / a.steps, "share": v[0] / stage_ms} for k, v in stages.items()}
        line["stage_pass_ms_per_step"] = stage_ms / a.steps
        if per_stage:
            top = max(per_stage, key=lambda k: per_stage[k]["share"])
            rl = stage_roofline(cost[top], per_stage[top]["ms_per_launch"], pk)
            line["roofline"] = {"kernel": top, "bound": rl["bound"], "achieved": rl["achieved"], "peak": pk[rl["bound"]],
                                "unit": rl["unit"], "frac": rl["frac"], "traffic": ncu_traffic(a, top), "peak_source": pk["src"],
                                "algorithmic_per_launch": cost[top][0] if rl["bound"] == "tensor" else cost[top][1],
                                "ms_per_launch": per_stage[top]["ms_per_launch"],
                                "timing": "CUDA events on the launch stream at stage boundaries, K steps of the same workload run "
                                          "right after the timed region with the two-stream pinv/attn3v overlap off"}
            detail = {}
            for k, v in per_stage.items():
                r = stage_roofline(cost[k], v["ms_per_launch"], pk)
                detail[k] = {"ms_per_launch": round(v["ms_per_launch"], 5), "share": round(v["share"], 4), "bound": r["bound"],
                             "achieved": round(r["achieved"], 2), "frac": round(r["frac"], 4),
                             "tflops": round(r["tflops"], 2), "frac_tensor": round(r["frac_tensor"], 4),
                             "gbs": round(r["gbs"], 1), "frac_hbm": round(r["frac_hbm"], 4)}
            line["stages"] = detail
        if cb is not None:
            line["cpu_baseline"] = cb
        if ref is not None:
            got = logits.detach().cpu()
            line["parity"] = {"max_abs_dlogit": float((got - ref).abs().max()),
                              "max_abs_dprob": float((torch.softmax(got, 1) - torch.softmax(ref, 1)).abs().max()),
                              "vs": "oracle/transmil_oracle.py (fp32) on rank 0's bag; gates 1e-3 / 1e-4"}
        line.update(extras)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
