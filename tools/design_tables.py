#!/usr/bin/env python
"""Prints the two measured tables of DESIGN.md (section 3: per-stage roofline, section 5: launch list vs event pass) from a
bench.py line, the ncu launch list of the same build and the --set full traffic summary.
    python tools/design_tables.py profiles/r2/bench_1gpu_v3.json profiles/r2_forward_launches.csv profiles/r2_ncu_traffic.json"""
import csv, json, re, sys

STAGE = [("gemm_tc_kernel<256, 0", "fc1_gemm1"), ("gemm_tc_kernel<256, 1, 1", "qkv"), ("gemm_tc_kernel<256, 1, 2", "to_out"),
         ("ln_pad_kernel", "ln_pad"), ("landmarks16_kernel", "landmarks"), ("attn2_kernel", "pinv"), ("pinv_", "pinv"),
         ("attn3v_", "attn3v"), ("zw16_kernel", "zw"), ("nystrom_out_tc_kernel", "nystrom_out"), ("ppeg_", "ppeg"),
         ("head_kernel", "head"), ("prep_operands", "prep"), ("cls_rows", "fc1_tail")]


def main():
    line = json.loads([x for x in open(sys.argv[1]) if x.startswith("{")][-1])
    traffic = json.load(open(sys.argv[3]))["stages"] if len(sys.argv) > 3 else {}
    print("| stage | µs per launch (event pass) | bound | achieved | of measured peak | tensor / HBM fraction | dram MB per launch (ncu) |")
    print("|---|---|---|---|---|---|---|")
    for k, v in line["stages"].items():
        unit = "TFLOP/s" if v["bound"] == "tensor" else "GB/s"
        mb = traffic.get(k, {}).get("dram_bytes_per_launch")
        print(f"| {k} | {v['ms_per_launch'] * 1e3:.1f} | {v['bound']} | {v['achieved']:.0f} {unit} | {v['frac']:.2f} | "
              f"{v['frac_tensor']:.2f} / {v['frac_hbm']:.2f} | {'' if mb is None else round(mb / 1e6)} |")
    if len(sys.argv) > 2:
        rows = [r for r in csv.reader(l for l in open(sys.argv[2]) if l.startswith('"'))]
        col = {h: i for i, h in enumerate(rows[0])}
        body = [r for r in rows[1:] if r[col["Metric Name"]] == "gpu__time_duration.sum"]
        starts = [i for i, r in enumerate(body) if "prep_operands" in r[col["Kernel Name"]]]
        fwd = body[starts[-2]:starts[-1]] if len(starts) > 1 else body[starts[-1]:]
        agg = {}
        for r in fwd:
            st = next((s for sub, s in STAGE if sub in r[col["Kernel Name"]]), "other")
            agg[st] = agg.get(st, 0.0) + float(r[col["Metric Value"]].replace(",", "")) / 1e3
        tot = sum(agg.values())
        ev = {k: v["ms_per_launch"] * 1e3 * (1 if k in ("fc1_gemm1", "fc1_tail", "ppeg", "head") else 2) for k, v in line["stages"].items()}
        evt = sum(ev.values())
        print(f"\n{len(fwd)} launches in one forward of the launch list")
        print("| stage | launch list µs / forward | share | stage pass µs / forward | share |")
        print("|---|---|---|---|---|")
        for k in ["prep"] + list(line["stages"].keys()):
            a, e = agg.get(k), ev.get(k)
            print(f"| {k} | {'' if a is None else f'{a:.1f}'} | {'' if a is None else f'{a / tot:.3f}'} | {'' if e is None else f'{e:.1f}'} | {'' if e is None else f'{e / evt:.3f}'} |")
        print(f"| sum | {tot:.0f} | | {evt:.0f} | |")


if __name__ == "__main__":
    main()
