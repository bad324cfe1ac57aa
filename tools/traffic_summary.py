"""profiles/ncu_traffic.json + a compact per-launch table from the one-step ncu metrics pass of tools/profile_step.sh.
usage: python tools/traffic_summary.py gpurun_out/<tag>_conv_traffic.csv <tag>"""
import collections
import csv
import json
import re
import sys


def main():
    path, tag = sys.argv[1], sys.argv[2]
    per_step = int(sys.argv[3]) if len(sys.argv) > 3 else 0     # conv-kernel launches of ONE step: trim the capture to
    rows = list(csv.DictReader([l for l in open(path) if l.startswith('"')]))   # one period (any cyclic window is one step)
    by = collections.OrderedDict()
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1.0, "us": 1e3, "ms": 1e6}
    for r in rows:
        e = by.setdefault(r["ID"], {"name": re.sub(r"\(.*", "", r["Kernel Name"]).replace("void ", "").replace("ubr::", ""),
                                    "grid": r["Grid Size"]})
        e[r["Metric Name"]] = float(r["Metric Value"].replace(",", "")) * mult.get(r["Metric Unit"], 1.0)
    if per_step and len(by) > per_step:
        by = collections.OrderedDict(list(by.items())[:per_step])
    out_csv = "profiles/%s_conv_traffic_one_step.csv" % tag
    agg = collections.OrderedDict()
    with open(out_csv, "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["id", "kernel", "grid", "time_us", "dram_read_MB", "dram_write_MB", "dram_pct_of_peak",
                    "tensor_pipe_active_pct", "sm_throughput_pct", "l2_hit_pct"])
        for i, e in by.items():
            t = e["gpu__time_duration.sum"]
            rd, wr = e["dram__bytes_read.sum"], e["dram__bytes_write.sum"]
            tp = e["sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"]
            w.writerow([i, e["name"], e["grid"], round(t / 1e3, 2), round(rd / 1e6, 2), round(wr / 1e6, 2),
                        e["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"], tp,
                        e["sm__throughput.avg.pct_of_peak_sustained_elapsed"], e["lts__t_sector_hit_rate.pct"]])
            a = agg.setdefault(re.sub(r"<.*", "", e["name"]), {"n": 0, "bytes": 0.0, "ns": 0.0, "tp": 0.0})
            a["n"] += 1
            a["bytes"] += rd + wr
            a["ns"] += t
            a["tp"] += tp * t
    out = {}
    for k, a in agg.items():
        out[k] = {"launches_per_step": a["n"], "dram_bytes_per_launch": a["bytes"] / a["n"], "dram_bytes_per_step": a["bytes"],
                  "ncu_time_us_per_step": a["ns"] / 1e3, "dram_gbs_under_ncu": a["bytes"] / a["ns"],
                  "tensor_pipe_active_pct_time_weighted": a["tp"] / a["ns"],
                  "source": "%s (ncu dram__bytes_read.sum + dram__bytes_write.sum, every conv-kernel launch of one batch-32 "
                            "training step)" % out_csv}
        print(k, {kk: (round(v, 2) if isinstance(v, float) else v) for kk, v in out[k].items() if kk != "source"})
    json.dump(out, open("profiles/ncu_traffic.json", "w"), indent=1)


if __name__ == "__main__":
    main()
