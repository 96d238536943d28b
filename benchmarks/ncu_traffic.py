"""DRAM traffic per launch of k_bounce from an `ncu --set full` raw page (CSV):

    python benchmarks/ncu_traffic.py gpurun_out/r5_k_bounce_raw.csv profiles/k_bounce_traffic.json

The capture (benchmarks/profile_round.sh) holds the consecutive launches of ONE whole bench.py step -- every
batch of its 48 iterations, bounce 0 to depth-1 each -- so their mean is the per-launch figure bench.py's
roofline.achieved is quoted per (algorithmic bytes / launch vs dram__bytes_read.sum + dram__bytes_write.sum)."""
import csv
import json
import sys


def to_bytes(v, unit):
    return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def main(src, dst):
    rows = list(csv.reader(open(src)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    per = []
    for r in rows[2:]:
        if "k_bounce" not in r[col["Kernel Name"]]:
            continue
        rd = to_bytes(r[col["dram__bytes_read.sum"]], units[col["dram__bytes_read.sum"]])
        wr = to_bytes(r[col["dram__bytes_write.sum"]], units[col["dram__bytes_write.sum"]])
        num = lambda name: float(r[col[name]].replace(",", "")) if name in col and r[col[name]] else None
        per.append({"read": rd, "write": wr, "us": num("gpu__time_duration.sum"),
                    "warp_inst": num("smsp__inst_executed.sum"),
                    "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                    "threads_per_inst": num("smsp__thread_inst_executed_per_inst_executed.ratio"),
                    "warps_active_pct": num("sm__warps_active.avg.pct_of_peak_sustained_active")})
    out = {"kernel": "k_bounce", "launches": len(per),
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
    main(sys.argv[1], sys.argv[2])
