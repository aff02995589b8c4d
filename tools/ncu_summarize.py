#!/usr/bin/env python
"""Turn `ncu -i full.ncu-rep --page raw --csv` of one bench.py forward into the two files bench.py and the judge read:
profiles/<tag>_forward_ncu_full_summary.csv (one row per launch, a handful of metrics) and profiles/<tag>_ncu_traffic.json
(dram bytes per launch of each stage).  Usage: python tools/ncu_summarize.py gpurun_out/full_raw.csv [tag]"""
import csv, json, os, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "l1tex__data_bank_conflicts_pipe_lsu.sum"]
# kernel-name substring -> stage, in launch order of one forward; layer kernels appear twice
STAGE = [("gemm_tc_kernel<256, 0", "fc1_gemm1"), ("gemm_tc_kernel<256, 1, 1", "qkv"), ("gemm_tc_kernel<256, 1, 2", "to_out"),
         ("ln_pad_kernel", "ln_pad"), ("landmarks16_kernel", "landmarks"), ("attn2_kernel", "pinv"), ("pinv_", "pinv"),
         ("attn3v_", "attn3v"), ("zw16_kernel", "zw"), ("nystrom_out_tc_kernel", "nystrom_out"), ("ppeg_", "ppeg"),
         ("head_kernel", "head"), ("prep_operands", "prep"), ("cls_rows", "fc1_gemm1")]
PER_FORWARD = {"fc1_gemm1": 1, "ppeg": 1, "head": 1, "prep": 1}   # every other stage runs once per layer = twice


def to_bytes(v, unit):
    f = float(v.replace(",", ""))
    return f * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def main():
    path = sys.argv[1]
    tag = sys.argv[2] if len(sys.argv) > 2 else "r1"
    rows = list(csv.reader(open(path)))
    hdr, units, body = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    kn = col["Kernel Name"]
    keep = [k for k in KEEP if k in col]
    # one forward = from the first prep_operands launch to the next one (or the end)
    starts = [i for i, r in enumerate(body) if "prep_operands" in r[kn]]
    if not starts:
        raise SystemExit("no prep_operands launch in the capture")
    a = starts[0]
    b = starts[1] if len(starts) > 1 else len(body)
    fwd = body[a:b]
    out = os.path.join(ROOT, "profiles", f"{tag}_forward_ncu_full_summary.csv")
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["Kernel Name"] + [f"{k} [{units[col[k]]}]" for k in keep])
        for r in fwd:
            w.writerow([r[kn]] + [r[col[k]] for k in keep])
    stages = {}
    for r in fwd:
        st = next((s for sub, s in STAGE if sub in r[kn]), None)
        if st is None:
            print("unmapped kernel", r[kn]); continue
        byt = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]]) + \
            to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        e = stages.setdefault(st, {"dram_bytes_per_launch": 0.0, "kernels": [], "gpu_time_us": 0.0})
        e["dram_bytes_per_launch"] += byt
        name = r[kn].replace("void ", "").split("(")[0]
        if name not in e["kernels"]:
            e["kernels"].append(name)
        t = float(r[col["gpu__time_duration.sum"]].replace(",", ""))
        e["gpu_time_us"] += t * {"ns": 1e-3, "us": 1, "ms": 1e3}.get(units[col["gpu__time_duration.sum"]], 1)
    for st, e in stages.items():
        n = PER_FORWARD.get(st, 2)
        e["dram_bytes_per_launch"] /= n
        e["gpu_time_us"] = round(e["gpu_time_us"] / n, 3)
    js = {"source": f"ncu --set full --clock-control none, one forward of bench.py (50000 x 1024 bag), profiles/{tag}_forward_ncu_full_summary.csv; "
                    "dram__bytes_read.sum + dram__bytes_write.sum summed over the kernels of one launch of each stage (tools/ncu_summarize.py)",
          "workload": {"n": 50000, "in_features": 1024, "bags": 1}, "launches_in_forward": len(fwd), "stages": stages}
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_traffic.json"), "w") as f:
        json.dump(js, f, indent=1)
    tot = sum(e["gpu_time_us"] * PER_FORWARD.get(s, 2) for s, e in stages.items())
    for s, e in stages.items():
        print(f"{s:12s} {e['gpu_time_us']:9.2f} us/launch  share {e['gpu_time_us'] * PER_FORWARD.get(s, 2) / tot:6.3f}  dram {e['dram_bytes_per_launch'] / 1e6:8.2f} MB")
    print("launches", len(fwd), "total us", round(tot, 1))


if __name__ == "__main__":
    main()
