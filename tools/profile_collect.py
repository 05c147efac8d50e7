#!/usr/bin/env python
"""Turn what tools/profile_round.sh left in gpurun_out/<tag>/ into the tracked files under profiles/ (no GPU needed).

    python tools/profile_collect.py r1

  bench_*.json                 -> profiles/<tag>_bench_*.json (reference arms keep their arm name)
  launches_navigation.csv      -> profiles/<tag>_launches_navigation.txt (per kernel: launches, total, share of GPU time)
  step_kernel_<wl>.ncu-rep     -> profiles/<tag>_step_kernel_<wl>.txt (raw-page summary) and <tag>_<wl>_lines.txt
                                  (per-source-line cost), and the kernel's DRAM bytes into profiles/<tag>_traffic.json
"""
import collections
import csv
import glob
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "vectorizedmultiagentsimulator_b200", "libvmasb200.so")


def launches(path, out):
    rows = [r for r in csv.reader(open(path, errors="replace")) if r]
    hi = next(i for i, r in enumerate(rows) if r[0] == "ID")
    H = rows[hi]
    kn, mn, mv = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[hi + 1:]:
        if len(r) <= mv or r[mn] != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r[kn])[:90]
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += float(r[mv].replace(",", ""))
    tot = sum(a[1] for a in agg.values()) or 1.0
    with open(out, "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none -c 400: python bench.py --steps 20 --warmup 3\n")
        f.write("# (cold-cache, serialised launches: read the SHARE, not the absolute)\n")
        f.write(f"{'kernel':<92s} {'launches':>8s} {'total us':>10s} {'avg us':>8s} {'share':>6s}\n")
        for name, (n, ns) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
            f.write(f"{name:<92s} {n:>8d} {ns / 1e3:>10.1f} {ns / 1e3 / n:>8.2f} {100 * ns / tot:>5.1f}%\n")


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = [r for r in csv.reader(out.splitlines()) if r]
    hi = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
    H, U, D = rows[hi], rows[hi + 1], rows[hi + 2]
    return {h: (D[i], U[i]) for i, h in enumerate(H)}


def to_bytes(v, unit):
    mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]
    return float(v.replace(",", "")) * mult


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r1"
    src = os.path.join(ROOT, "gpurun_out", tag)
    dst = os.path.join(ROOT, "profiles")
    for p in sorted(glob.glob(os.path.join(src, "bench_*.json"))):
        name = os.path.basename(p)[len("bench_"):-len(".json")]
        line = open(p).read().strip()
        try:
            json.loads(line)
        except ValueError:
            print("skipping (not a JSON line):", p, line[:120])
            continue
        arm = name.startswith(("reference", "dropin", "generic"))
        with open(os.path.join(dst, f"{tag}_{name}.json" if arm else f"{tag}_bench_{name}.json"), "w") as f:
            f.write(line + "\n")
    lp = os.path.join(src, "launches_navigation.csv")
    if os.path.exists(lp):
        launches(lp, os.path.join(dst, f"{tag}_launches_navigation.txt"))
    # tools/profile_lines.sh reduced every capture on the box: step_kernel_<wl>.txt (raw-page summary, first line = kernel
    # name) and <wl>_lines.txt (per-source-line cost).  The kernel's DRAM bytes / warp instructions per launch go to
    # <tag>_traffic.json, which bench.py reads for roofline.traffic / roofline.issue.
    traffic = {}
    for summ in sorted(glob.glob(os.path.join(src, "step_kernel_*.txt"))):
        wl = os.path.basename(summ)[len("step_kernel_"):-len(".txt")]
        text = open(summ).read()
        kname = text.splitlines()[0].strip("# ") if text else ""
        with open(os.path.join(dst, f"{tag}_step_kernel_{wl}.txt"), "w") as f:
            f.write(f"# ncu --set full --import-source on --clock-control none -k 'regex:step_kernel|vmasb200_step_world' -s 6 -c 1 (bench.py, workload {wl})\n" + text)
        lp = os.path.join(src, f"{wl}_lines.txt")
        if os.path.exists(lp):
            with open(os.path.join(dst, f"{tag}_{wl}_lines.txt"), "w") as f:
                f.write(f"# tools/ncu_lines.py on the same capture ({kname})\n" + open(lp).read())

        def metric(name):
            m = re.search(r"^" + re.escape(name) + r" \[(\w*)\]: ([\d.,]+)", text, re.M)
            return (m.group(2), m.group(1)) if m else None

        rd, wr, inst, dur, grid = (metric(k) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
                                                       "gpu__time_duration.sum", "launch__grid_size"))
        if not (rd and wr and inst):
            continue
        m = re.match(r"(.*)_(\d+)$", wl)
        scen, n = (m.group(1), int(m.group(2))) if m else (wl, 32768)
        rdb, wrb = to_bytes(*rd), to_bytes(*wr)
        fl = [metric(f"smsp__sass_thread_inst_executed_op_{k}_pred_on.sum") for k in ("fadd", "fmul", "ffma")]
        flops = None
        if all(fl):
            fa, fm, ff = (float(x[0].replace(",", "")) for x in fl)
            flops = fa + fm + 2 * ff  # FP32 operations the launch executed (thread level)
        traffic[f"{scen}@{n}"] = {"fp32_flops": flops, "n_envs": n, "dram_bytes": rdb + wrb, "dram_read": rdb, "dram_write": wrb,
                                  "grid": int(float(grid[0].replace(",", ""))) if grid else None,
                                  "warp_inst": float(inst[0].replace(",", "")), "duration": " ".join(dur) if dur else None,
                                  "kernel": kname, "source": f"profiles/{tag}_step_kernel_{wl}.txt"}
    if traffic:
        with open(os.path.join(dst, f"{tag}_traffic.json"), "w") as f:
            json.dump(traffic, f, indent=1, sort_keys=True)
            f.write("\n")
    print("\n".join(sorted(os.listdir(dst))))


if __name__ == "__main__":
    main()
