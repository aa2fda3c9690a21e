#!/usr/bin/env python3
"""Rebuild profiles/ from the latest gpurun_out captures:
   refresh_profiles.py <launch-list.csv> <extract.ncu-rep> <hamming.ncu-rep> <bench_default.json>"""
import csv, io, json, shutil, subprocess, sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
launches, rep, ham, bench = sys.argv[1:5]
with open(launches) as f, open(os.path.join(ROOT, "profiles/r1_launches_bench_steps2.csv"), "w") as o:
    o.writelines(l for l in f if not l.startswith("=="))
summ = os.path.join(ROOT, "tools/ncu_summary.py")
with open(os.path.join(ROOT, "profiles/r1_ncu_full_summary.csv"), "w") as o:
    o.write(subprocess.run([sys.executable, summ, rep, "fastStrip", "gauss9", "describe", "octree", "resizeLinearWord"],
                           capture_output=True, text=True).stdout)
with open(os.path.join(ROOT, "profiles/r1_ncu_hamming_summary.csv"), "w") as o:
    o.write(subprocess.run([sys.executable, summ, ham, "hammingBest2"], capture_output=True, text=True).stdout)
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out))); h, u = rows[0], rows[1]
iN, iR, iW, iT = (h.index(k) for k in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
tscale = {"ns": 1e-3, "us": 1, "ms": 1e3}
stage = {"resizeLinear": "pyramid", "fastStrip": "fast", "octree": "octree", "gauss9": "blur", "describe": "describe"}
tot, tim = {}, {}
for r in rows[2:]:
    for k, s in stage.items():
        if k in r[iN]:
            tot[s] = tot.get(s, 0) + float(r[iR]) * scale[u[iR]] + float(r[iW]) * scale[u[iW]]
            tim[s] = tim.get(s, 0) + float(r[iT]) * tscale[u[iT]]
json.dump({"note": "dram__bytes_read.sum + dram__bytes_write.sum per launch / 256 frames, ncu --set full capture of one step of "
                   "`python bench.py --steps 2 --warmup 3 --batch 256` (profiles/r1_ncu_full_summary.csv); pyramid = sum of its 7 launches",
           "bytes_per_frame": {k: v / 256 for k, v in tot.items()}, "ncu_us_per_256_frames": tim},
          open(os.path.join(ROOT, "profiles/r1_traffic.json"), "w"), indent=1)
shutil.copy(bench, os.path.join(ROOT, "profiles/r1_bench_default.json"))
print(json.dumps({"bytes_per_frame": {k: round(v / 256) for k, v in tot.items()}, "us": tim}))
