"""Turns an ncu report (.ncu-rep, read here with `ncu -i ... --page raw --csv`) into the small
tracked summaries under profiles/: a per-kernel metrics JSON and profiles/traffic.json, which
bench.py reads for roofline.traffic.  Usage: python profiles/summarize_ncu.py REPORT OUT_PREFIX"""
import csv
import io
import json
import subprocess
import sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
        "lts__t_bytes.sum", "l1tex__t_bytes.sum", "sm__cycles_elapsed.avg.per_second",
        "dram__cycles_elapsed.avg.per_second", "smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__warp_issue_stalled_barrier_per_warp_active.pct", "smsp__inst_executed.sum"]
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    rep, prefix = sys.argv[1], sys.argv[2]
    raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], stderr=subprocess.DEVNULL).decode()
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    out = []
    for r in rows[2:]:
        d = {"kernel": r[hdr.index("Kernel Name")], "grid": r[hdr.index("Grid Size")], "block": r[hdr.index("Block Size")]}
        for k in KEEP:
            if k in hdr:
                i = hdr.index(k)
                d[k] = {"value": r[i], "unit": units[i]}
        out.append(d)
    json.dump(out, open(prefix + "_metrics.json", "w"), indent=1)
    fused = [d for d in out if "::ga_" in d["kernel"] and "drain" not in d["kernel"]]
    if fused:
        def b(d, k):
            return float(d[k]["value"].replace(",", "")) * SCALE.get(d[k]["unit"], 1)
        per = [b(d, "dram__bytes_read.sum") + b(d, "dram__bytes_write.sum") for d in fused]
        json.dump({"dram_bytes_per_launch": sum(per) / len(per), "launches": len(per), "kernel": fused[0]["kernel"],
                   "grid": fused[0]["grid"], "source": rep.split("/")[-1],
                   "note": "dram__bytes_read.sum + dram__bytes_write.sum, ncu --set full, one launch each"},
                  open("profiles/traffic.json", "w"), indent=1)
    print(json.dumps(out[:1], indent=1)[:1500])


if __name__ == "__main__":
    main()
