"""Turns an ncu --csv launch list of `bench.py` into a per-kernel summary of ONE step and the traffic file bench.py quotes.

    python scratch/ncu_step_summary.py gpurun_out/r02h_metrics_rubik.csv rubik profiles/r02h

Expects the metrics gpu__time_duration.sum, dram__bytes_read.sum, dram__bytes_write.sum, lts__t_bytes.sum,
smsp__inst_executed.sum, smsp__thread_inst_executed.sum, sm__cycles_elapsed.max.  A step = the launches from one
clear_accum_kernel (the reset dispatch that opens every bench step) to the next.  Writes
  <prefix>_step_<workload>.txt       per kernel: launches, time, share, DRAM / L2 bytes, lanes per instruction,
                                     issue-slot utilisation = warp instructions / (cycles x 592 schedulers),
                                     lane-issue utilisation = thread instructions / (cycles x 592 x 32)
  profiles/r02_traffic_<workload>.json   dram bytes per step and per ray-answering launch (bench.py roofline.traffic)
Times under ncu are serialised and cold-cache: shares are what they are good for, not absolute step time.
"""
import csv
import json
import os
import sys
from collections import OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RAY_KERNELS = ("wf_init_kernel", "wf_find_kernel", "wf_replay_kernel", "wf_trace_kernel", "wf_tail_kernel", "pathtrace")


def unit_scale(unit):
    u = unit.strip().lower()
    table = {"ns": 1e-9, "us": 1e-6, "usecond": 1e-6, "ms": 1e-3, "msecond": 1e-3, "nsecond": 1e-9, "s": 1.0, "second": 1.0,
             "byte": 1.0, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9, "tbyte": 1e12, "inst": 1.0, "cycle": 1.0, "": 1.0}
    return table.get(u, 1.0)


def main():
    path, workload, prefix = sys.argv[1], sys.argv[2], sys.argv[3]
    with open(path) as f:
        lines = [ln for ln in f if ln.startswith('"')]
    launches = OrderedDict()  # id -> {name, metric: value}
    for row in csv.DictReader(lines):
        k = row["ID"]
        d = launches.setdefault(k, {"name": row["Kernel Name"]})
        try:
            v = float(row["Metric Value"].replace(",", ""))
        except ValueError:
            continue
        d[row["Metric Name"]] = v * unit_scale(row["Metric Unit"])
    seq = list(launches.values())
    clears = [i for i, d in enumerate(seq) if "clear_accum_kernel" in d["name"]]
    if len(clears) < 3:
        raise SystemExit("fewer than three steps in %s" % path)
    # the last complete step of the list that ran the default path (bench.py's counting passes at the end make waves to the
    # end with counting kernels instead of the tail kernel); any step when the workload has no tail kernel at all
    a = b = None
    for i in range(len(clears) - 1, 0, -1):
        names = " ".join(d["name"] for d in seq[clears[i - 1]:clears[i]])
        if "wf_tail_kernel" in names:
            a, b = clears[i - 1], clears[i]
            break
    if a is None:
        a, b = clears[-2], clears[-1]
    step = seq[a:b]

    def short(n):
        n = n.split("(")[0].replace("void ", "").replace("srt::", "")
        return n

    agg = OrderedDict()
    for d in step:
        e = agg.setdefault(short(d["name"]), {"n": 0, "t": 0.0, "dram": 0.0, "l2": 0.0, "wi": 0.0, "ti": 0.0, "cyc": 0.0})
        e["n"] += 1
        e["t"] += d.get("gpu__time_duration.sum", 0.0)
        e["dram"] += d.get("dram__bytes_read.sum", 0.0) + d.get("dram__bytes_write.sum", 0.0)
        e["l2"] += d.get("lts__t_bytes.sum", 0.0)
        e["wi"] += d.get("smsp__inst_executed.sum", 0.0)
        e["ti"] += d.get("smsp__thread_inst_executed.sum", 0.0)
        e["cyc"] += d.get("sm__cycles_elapsed.max", 0.0)
    total_t = sum(e["t"] for e in agg.values())
    total_dram = sum(e["dram"] for e in agg.values())
    out = ["one step of `bench.py --workload %s` under ncu (launches %d..%d of %s; serialised, cold cache: read the shares)" % (
        workload, a, b - 1, os.path.basename(path)),
        "%-34s %4s %9s %6s %10s %10s %6s %7s %7s" % ("kernel", "n", "ms", "share", "DRAM MB", "L2 MB", "lanes", "issue", "lane-u")]
    ray_launches, ray_dram = 0, 0.0
    for name, e in sorted(agg.items(), key=lambda kv: -kv[1]["t"]):
        lanes = e["ti"] / e["wi"] if e["wi"] else 0.0
        issue = e["wi"] / (e["cyc"] * 592.0) if e["cyc"] else 0.0
        laneu = e["ti"] / (e["cyc"] * 592.0 * 32.0) if e["cyc"] else 0.0
        out.append("%-34s %4d %9.3f %5.1f%% %10.1f %10.1f %6.1f %7.3f %7.3f" % (
            name[:34], e["n"], e["t"] * 1e3, 100.0 * e["t"] / total_t, e["dram"] / 1e6, e["l2"] / 1e6, lanes, issue, laneu))
        if any(k in name for k in RAY_KERNELS):
            ray_launches += e["n"]
            ray_dram += e["dram"]
    out.append("%-34s %4d %9.3f %5.1f%% %10.1f" % ("TOTAL", len(step), total_t * 1e3, 100.0, total_dram / 1e6))
    out.append("ray-answering kernels: %d launches, %.1f MB of DRAM traffic (%.1f%% of the step's)" % (
        ray_launches, ray_dram / 1e6, 100.0 * ray_dram / max(total_dram, 1.0)))
    text = "\n".join(out) + "\n"
    with open("%s_step_%s.txt" % (prefix, workload), "w") as f:
        f.write(text)
    print(text)
    # bench.py counts as ray-answering "launches" what SRT_OPT_TIME_KERNELS brackets: one per wave + init + tail
    n_brackets = sum(e["n"] for n, e in agg.items() if "wf_logic_kernel" in n) + sum(
        e["n"] for n, e in agg.items() if "wf_init_kernel" in n or "wf_tail_kernel" in n)
    n_brackets = max(n_brackets, 1)
    with open(os.path.join(ROOT, "profiles", "r02_traffic_%s.json" % workload), "w") as f:
        json.dump({"build": os.path.basename(prefix), "source": os.path.basename(path), "launches_in_step": len(step),
                   "dram_bytes_per_step": total_dram, "ray_kernels_dram_bytes_per_step": ray_dram,
                   "bracketed_launches_per_step": n_brackets, "dram_bytes_per_launch": ray_dram / n_brackets,
                   "l2_bytes_per_step": sum(e["l2"] for e in agg.values())}, f, indent=1)


if __name__ == "__main__":
    main()
