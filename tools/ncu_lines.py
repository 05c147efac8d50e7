#!/usr/bin/env python
"""Per-source-line cost of a kernel from an .ncu-rep captured with --import-source on (needs -lineinfo).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep path/to/lib.so [kernel-substring] [launch-index]

Joins the SASS page of the report (executed-instruction and stall-sample counts per instruction) with the line table
nvdisasm prints for the same cubin, and prints the hottest source lines (inlined callee lines are attributed to the
innermost line, which is what nvdisasm reports)."""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile


def line_table(so, kernel):
    tmp = tempfile.mkdtemp()
    if so.endswith(".cubin"):  # a world-compiled kernel from the JIT cache (vectorizedmultiagentsimulator_b200/_jit)
        import shutil

        shutil.copy(so, tmp)
    else:
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, capture_output=True)
    table = {}
    for f in os.listdir(tmp):
        if not f.endswith(".cubin"):
            continue
        out = subprocess.run(["nvdisasm", "-gi", "-c", os.path.join(tmp, f)], capture_output=True, text=True).stdout
        sect, cur, chain, fresh = None, None, [], True
        for ln in out.splitlines():
            m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
            if m:
                sect = m.group(1)
                continue
            if sect is None or kernel not in sect:
                continue
            m = re.match(r'\s*//## File "(.*?)", line (\d+)', ln)
            if m:
                if fresh:
                    chain, fresh = [], False
                chain.append((os.path.basename(m.group(1)), int(m.group(2))))  # innermost first, outermost last
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
            if m:
                fresh = True
                table[(sect, int(m.group(1), 16))] = (tuple(chain), m.group(2).strip())
    return table


def main():
    rep, so = sys.argv[1], sys.argv[2]
    kernel = sys.argv[3] if len(sys.argv) > 3 else "step_kernel"
    idx = sys.argv[4] if len(sys.argv) > 4 else "1"
    top = int(os.environ.get("NCU_LINES_TOP", "22"))  # rows per frame table
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f":::{idx}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    H = rows[hi]
    ci = {n: H.index(n) for n in ("Address", "Source", "# Samples", "Instructions Executed", "Thread Instructions Executed")}
    table = line_table(so, kernel)
    sects = sorted({k[0] for k in table})
    # the report lists the entry function first, then its non-inlined callees; offsets restart per function
    by_line = collections.defaultdict(lambda: [0, 0, 0])
    by_op = collections.defaultdict(lambda: [0, 0])
    by_frame = {l: collections.defaultdict(lambda: [0, 0, 0]) for l in (1, 2, 3)}
    total = [0, 0, 0]
    base, prev, si = None, None, 0
    dump, dumped = os.environ.get("NCU_LINES_DUMP"), []  # "file.cuh:123": list every SASS instruction of that line
    entry = [s for s in sects if "step_kernel" in s or kernel in s]
    order = entry[:1] + [s for s in sects if s not in entry[:1]]
    for r in rows[hi + 1:]:
        if r and r[0] == "Kernel Name":
            break  # the next launch of the capture
        if len(r) <= ci["Thread Instructions Executed"]:
            continue
        addr = int(r[ci["Address"]], 16)
        if base is None:
            base = addr
        off = addr - base
        key = None
        for s in order:
            t = table.get((s, off))
            if t and t[1].split()[0] == r[ci["Source"]].strip().split()[0]:
                key = t
                break
        smp, ins, tins = int(r[ci["# Samples"]] or 0), int(r[ci["Instructions Executed"]] or 0), int(r[ci["Thread Instructions Executed"]] or 0)
        total[0] += smp; total[1] += ins; total[2] += tins
        chain = key[0] if key and key[0] else (("?", 0),)
        loc = chain[0]
        if dump and f"{loc[0]}:{loc[1]}" == dump:
            dumped.append((off, r[ci["Source"]].strip(), ins, tins, smp))
        by_line[loc][0] += smp; by_line[loc][1] += ins; by_line[loc][2] += tins
        for lvl in (1, 2, 3):  # frames counted from the kernel body inwards
            fr = chain[-lvl] if len(chain) >= lvl else chain[0]
            by_frame[lvl][fr][0] += smp; by_frame[lvl][fr][1] += ins; by_frame[lvl][fr][2] += tins
        op = r[ci["Source"]].strip().split()[0].split(".")[0]
        by_op[op][0] += smp; by_op[op][1] += ins
    print(f"total: samples {total[0]}  warp-inst {total[1]}  thread-inst {total[2]}")
    print("--- top source lines by warp instructions executed")
    for loc, v in sorted(by_line.items(), key=lambda kv: -kv[1][1])[:45]:
        print(f"{loc[0]}:{loc[1]:<5d} inst {v[1]:>10d} ({100*v[1]/max(total[1],1):5.1f}%)  samples {v[0]:>7d} ({100*v[0]/max(total[0],1):5.1f}%)  thr/inst {v[2]/max(v[1],1):5.1f}")
    for lvl, name in ((1, "kernel body line"), (2, "first inlined frame"), (3, "second inlined frame")):
        print(f"--- by {name} (inclusive of everything inlined below it)")
        for loc, v in sorted(by_frame[lvl].items(), key=lambda kv: -kv[1][1])[:top]:
            print(f"{loc[0]}:{loc[1]:<5d} inst {v[1]:>10d} ({100*v[1]/max(total[1],1):5.1f}%)  samples {v[0]:>7d} ({100*v[0]/max(total[0],1):5.1f}%)  thr/inst {v[2]/max(v[1],1):5.1f}")
    print("--- top source lines by stall samples")
    for loc, v in sorted(by_line.items(), key=lambda kv: -kv[1][0])[:25]:
        print(f"{loc[0]}:{loc[1]:<5d} samples {v[0]:>7d} ({100*v[0]/max(total[0],1):5.1f}%)  inst {v[1]:>10d}")
    print("--- opcodes by warp instructions")
    for op, v in sorted(by_op.items(), key=lambda kv: -kv[1][1])[:25]:
        print(f"{op:<12s} inst {v[1]:>10d} ({100*v[1]/max(total[1],1):5.1f}%)  samples {v[0]:>7d}")
    if dump:
        print(f"--- SASS attributed to {dump} (offset, instruction, warp-inst, thread-inst, samples)")
        for off, src, ins, tins, smp in dumped:
            print(f"{off:06x}  {src[:90]:<90s} {ins:>9d} {tins:>11d} {smp:>5d}")


if __name__ == "__main__":
    main()
