"""DRAM traffic per launch of k_bounce from an `ncu --set full` raw page (CSV):

    python benchmarks/ncu_traffic.py gpurun_out/r5_k_bounce_raw.csv profiles/k_bounce_traffic.json [KERNEL_REGEX [GROUP]]

KERNEL_REGEX (default k_bounce) selects the kernels; GROUP (default 1) consecutive selected kernels make one "launch" --
the scenes whose bounce is two kernels (closest hit with dynamic ray fetch + shading: synthetic_4096) use
'k_isect_dyn|k_bounce' 2, which is the unit bench.py's bounce_kernel_launches counts.

The capture (benchmarks/profile_round.sh) holds the consecutive launches of ONE whole bench.py step -- every
batch of its 48 iterations, bounce 0 to depth-1 each -- so their mean is the per-launch figure bench.py's
roofline.achieved is quoted per (algorithmic bytes / launch vs dram__bytes_read.sum + dram__bytes_write.sum)."""
import csv
import json
import sys


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main(src, dst, regex="k_bounce", group=1):
    import re
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    per = []
    for r in rows[2:]:
        if not re.search(regex, r[col["Kernel Name"]]):
            continue
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        num = lambda name: float(r[col[name]].replace(",", "")) if name in col and r[col[name]] else None
        per.append({"read": rd, "write": wr, "us": num("gpu__time_duration.sum"),
                    "warp_inst": num("smsp__inst_executed.sum"),
                    "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "threads_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active")})
    if group > 1:                      # one unit = `group` consecutive kernels (sums; percentages time- / instruction-weighted)
        units_ = []
        for i in range(0, len(per) - group + 1, group):
            g = per[i:i + group]
            us = sum(q["us"] for q in g)
            wi = sum(q["warp_inst"] or 0 for q in g)
            units_.append({"read": sum(q["read"] for q in g), "write": sum(q["write"] for q in g), "us": us, "warp_inst": wi,
                           "issue_active_pct": sum((q["issue_active_pct"] or 0) * q["us"] for q in g) / us,
                           "threads_per_inst": sum((q["threads_per_inst"] or 0) * (q["warp_inst"] or 0) for q in g) / max(1.0, wi),
                           "warps_active_pct": sum((q["warps_active_pct"] or 0) * q["us"] for q in g) / us})
        per = units_
    out = {"kernel": regex if group > 1 or regex != "k_bounce" else "k_bounce", "launches": len(per),
           "bytes_per_launch": sum(p["read"] + p["write"] for p in per) / len(per),
           "warp_inst_per_launch": sum(p["warp_inst"] or 0 for p in per) / len(per),
           "issue_active_pct_time_weighted": sum((p["issue_active_pct"] or 0) * p["us"] for p in per) / sum(p["us"] for p in per),
           "threads_per_inst_inst_weighted": sum((p["threads_per_inst"] or 0) * (p["warp_inst"] or 0) for p in per)
                                             / max(1.0, sum(p["warp_inst"] or 0 for p in per)),
           "per_launch": per, "source": src,
           "how": "ncu --set full --clock-control none, dram__bytes_read.sum + dram__bytes_write.sum, mean over the "
                  "consecutive launches of one bench.py step (all batches, bounce 0..depth-1)"}
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps({k: out[k] for k in ("launches", "bytes_per_launch", "warp_inst_per_launch",
                                          "issue_active_pct_time_weighted", "threads_per_inst_inst_weighted")}))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "k_bounce", int(sys.argv[4]) if len(sys.argv) > 4 else 1)
