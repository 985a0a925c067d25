#!/usr/bin/env python
"""Turns an .ncu-rep (captured on the GPU box with `ncu --set full --import-source on`)
into a short markdown summary: headline metrics, pipe utilisation, stall reasons,
instruction mix and the hottest SASS lines.  Runs without a GPU (`ncu -i`)."""
import collections
import csv
import io
import subprocess
import sys

HEAD = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
    "launch__registers_per_thread", "launch__block_size", "launch__grid_size", "launch__shared_mem_per_block_dynamic",
    "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "smsp__inst_executed.sum",
]


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    raw = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, rows = raw[0], raw[1], raw[2:]
    idx = {h: i for i, h in enumerate(hdr)}
    print(f"# ncu summary of `{rep.split('/')[-1]}`\n")
    seen = set()
    for n, r in enumerate(rows):
        name = r[idx["Kernel Name"]].split("(")[0]
        if name in seen:
            continue
        seen.add(name)
        print(f"## {name}\n")
        print("| metric | value |\n|---|---|")
        for k in HEAD:
            if k in idx:
                print(f"| {k} | {r[idx[k]]} {units[idx[k]]} |")
        pipes = []
        for h, i in idx.items():
            if h.startswith("sm__inst_executed_pipe_") and h.endswith(".avg.pct_of_peak_sustained_active"):
                try:
                    pipes.append((float(r[i]), h.split("pipe_")[1].split(".")[0]))
                except ValueError:
                    pass
        pipes.sort(reverse=True)
        print("\npipes (% of peak while active): " + ", ".join(f"{p} {v:.1f}" for v, p in pipes[:6]))
        stalls = []
        for h, i in idx.items():
            if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h:
                try:
                    stalls.append((float(r[i]), h.split("issue_stalled_")[1].split("_per_issue")[0]))
                except ValueError:
                    pass
        stalls.sort(reverse=True)
        print("\nstalled warps per issue: " + ", ".join(f"{s} {v:.2f}" for v, s in stalls[:6]))
        src = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv", "--launch-skip", str(n),
                                               "--launch-count", "1"]))))
        if len(src) > 2:
            h = src[1]
            sidx = {c: i for i, c in enumerate(h)}
            secs = [i for i, row in enumerate(src) if row and row[0] == "Address"]
            sass = src[secs[-1] + 1:]
            mix, hot, tot, ts = collections.Counter(), [], 0, 0
            for row in sass:
                if len(row) < len(h):
                    continue
                try:
                    s, e = int(row[sidx["# Samples"]]), int(row[sidx["Instructions Executed"]])
                except ValueError:
                    continue
                text = row[sidx["Source"]].strip()
                op = (text.split()[1] if text.startswith("@") else text.split()[0]).split(".")[0]
                mix[op] += e
                tot += e
                ts += s
                hot.append((s, text))
            if tot:
                print("\ninstruction mix (warp instructions): " +
                      ", ".join(f"{op} {100 * c / tot:.1f}%" for op, c in mix.most_common(10)) + f" (total {tot})")
            hot.sort(reverse=True)
            if ts:
                print("\nhottest SASS by samples:\n")
                for s, text in hot[:8]:
                    print(f"- {100 * s / ts:.1f}% `{text[:90]}`")
        print()


if __name__ == "__main__":
    main()
