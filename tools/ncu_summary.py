#!/usr/bin/env python3
"""Summarise an .ncu-rep (read here, no GPU): per-kernel headline counters, and for chosen kernels the SASS hot
segments and stall reasons.  usage: ncu_summary.py rep.ncu-rep [kernel-regex ...]"""
import csv, io, subprocess, sys

WANT = ['gpu__time_duration.sum', 'launch__grid_size', 'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem',
        'launch__occupancy_limit_registers', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'dram__bytes_read.sum',
        'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'lts__t_sector_hit_rate.pct']


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    rep = sys.argv[1]
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units = rows[0], rows[1]
    print("kernel," + ",".join(WANT))
    for r in rows[2:]:
        name = r[hdr.index('Kernel Name')].split('(')[0]
        print('"' + name + '",' + ",".join(r[hdr.index(w)] if w in hdr else "" for w in WANT))   # template names hold commas
    for k in sys.argv[2:]:
        rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + k]))))
        if len(rows) < 3:
            print(f"\n== {k}: no kernel matches in {rep}")
            continue
        hdr = rows[1]
        iS, iE, iSm = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
        data = [r for r in rows[2:] if len(r) > iE and r[iE].isdigit()]   # several instances repeat the header
        tot = sum(int(r[iE]) for r in data); ts = max(1, sum(int(r[iSm]) for r in data))
        print(f"\n== {k}: {tot} warp instructions, {len(data)} SASS lines, {ts} samples")
        stall = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
        agg = {}
        for r in data:
            for i in stall:
                agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i])
        print("stalls:", sorted(agg.items(), key=lambda x: -x[1])[:8])
        segs, cur = [], None
        for kk, r in enumerate(data):
            e = int(r[iE]); s = r[iS].strip()
            if cur is None or not (0.7 * cur['e'] <= e <= 1.4 * cur['e']):
                cur = {'start': kk, 'e': e, 'sum': 0, 'n': 0, 'samples': 0, 'ops': {}}; segs.append(cur)
            cur['sum'] += e; cur['n'] += 1; cur['samples'] += int(r[iSm])
            op = s.split()[0] if not s.startswith('@') else s.split()[1]
            cur['ops'][op] = cur['ops'].get(op, 0) + 1
        for sg in segs:
            if sg['sum'] > 0.02 * tot:
                top = sorted(sg['ops'].items(), key=lambda x: -x[1])[:6]
                print(f"  sass[{sg['start']}:{sg['start'] + sg['n']}] exec/inst={sg['e']} inst_share={sg['sum'] / tot:.3f} "
                      f"sample_share={sg['samples'] / ts:.3f} ops={top}")


if __name__ == "__main__":
    main()
