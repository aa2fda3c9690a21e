#!/usr/bin/env python3
"""Per-CUDA-source-line instruction and stall-sample shares of one kernel in an .ncu-rep (needs -lineinfo and
--import-source on).  usage: ncu_lines.py rep.ncu-rep kernel-regex [min_share]"""
import csv, io, subprocess, sys


def main():
    rep, k = sys.argv[1], sys.argv[2]
    min_share = float(sys.argv[3]) if len(sys.argv) > 3 else 0.01
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda",
                          "--kernel-name", "regex:" + k], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = next(r for r in rows if r and r[0] == "Line No")
    iE, iS = hdr.index("Instructions Executed"), hdr.index("# Samples")
    lines = [r for r in rows if len(r) > iE and r[2] == "-" and r[0].isdigit()]   # CUDA lines (SASS rows carry an address)
    tot = sum(int(r[iE]) for r in lines) or 1
    ts = sum(int(r[iS]) for r in lines) or 1
    print(f"{k}: {tot} warp instructions, {ts} samples")
    for r in lines:
        e, s = int(r[iE]), int(r[iS])
        if e >= min_share * tot or s >= min_share * ts:
            print(f"{int(r[0]):5d} inst {e / tot:6.3f} samp {s / ts:6.3f}  {r[1].strip()[:110]}")


if __name__ == "__main__":
    main()
