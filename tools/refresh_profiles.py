#!/usr/bin/env python3
"""Rebuild the round's ncu artefacts under profiles/ from the latest gpurun_out captures:
   refresh_profiles.py <launch-list.csv> <extract.ncu-rep> [<hamming.ncu-rep>] [--round r2] [--batch 256]
   -> profiles/<round>_launches_bench_steps2.csv, <round>_ncu_full_summary.csv, <round>_ncu_hamming_summary.csv,
      <round>_traffic.json (per stage: DRAM bytes per frame, ALU-pipe %, warp instructions per frame, ncu time)"""
import csv, io, json, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = [a for a in sys.argv[1:] if not a.startswith("--")]
opt = dict(zip(sys.argv[1:], sys.argv[2:]))
rnd, batch = opt.get("--round", "r2"), int(opt.get("--batch", 256))
launches, rep = args[0], args[1]
ham = args[2] if len(args) > 2 else None
P = lambda name: os.path.join(ROOT, "profiles", f"{rnd}_{name}")
with open(launches) as f, open(P("launches_bench_steps2.csv"), "w") as o:
    o.writelines(l for l in f if not l.startswith("=="))
summ = os.path.join(ROOT, "tools/ncu_summary.py")
with open(P("ncu_full_summary.csv"), "w") as o:
    o.write(subprocess.run([sys.executable, summ, rep, "fastStrip", "gauss9", "describe", "octree", "resizeLinearStaged"],
                           capture_output=True, text=True).stdout)
if ham:
    with open(P("ncu_hamming_summary.csv"), "w") as o:
        o.write(subprocess.run([sys.executable, summ, ham, "hammingBest2"], capture_output=True, text=True).stdout)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); h, u = rows[0], rows[1]
col = lambda k: h.index(k)
iN, iR, iW, iT = col("Kernel Name"), col("dram__bytes_read.sum"), col("dram__bytes_write.sum"), col("gpu__time_duration.sum")
iA, iI = col("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"), col("smsp__inst_executed.sum")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-3, "us": 1, "ms": 1e3, "usecond": 1, "nsecond": 1e-3, "msecond": 1e3}
stage = {"resizeLinear": "pyramid", "fastStrip": "fast", "octree": "octree", "gauss9": "blur", "describe": "describe"}
num = lambda s: float(s.replace(",", ""))
tot, tim, alu, inst = {}, {}, {}, {}
for r in rows[2:]:
    for k, s in stage.items():
        if k in r[iN]:
            t = num(r[iT]) * tscale[u[iT]]
            tot[s] = tot.get(s, 0) + num(r[iR]) * scale[u[iR]] + num(r[iW]) * scale[u[iW]]
            tim[s] = tim.get(s, 0) + t
            alu[s] = alu.get(s, 0) + num(r[iA]) * t                       # time-weighted over a stage's launches
            inst[s] = inst.get(s, 0) + num(r[iI])
res = {"note": f"ncu --set full capture of one step of `python bench.py --steps 2 --warmup 3 --batch {batch}` "
               f"(profiles/{rnd}_ncu_full_summary.csv); per stage over its launches (pyramid = 7): dram__bytes_read.sum + "
               "dram__bytes_write.sum per frame, sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active (time-weighted), "
               "smsp__inst_executed.sum (warp instructions) per frame, gpu__time_duration.sum",
       "bytes_per_frame": {k: v / batch for k, v in tot.items()},
       "alu_pipe_pct": {k: alu[k] / tim[k] for k in alu},
       "warp_instructions_per_frame": {k: v / batch for k, v in inst.items()},
       f"ncu_us_per_{batch}_frames": tim}
json.dump(res, open(P("traffic.json"), "w"), indent=1)
print(json.dumps({k: {kk: round(vv) for kk, vv in v.items()} for k, v in res.items() if isinstance(v, dict)}))
