# One GPU-box pass for the round's artefacts: full GPU test suite, one bench line per config, the reference arm, the ncu
# launch list and one --set full capture of the extraction step.  Everything lands in gpurun_out/.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r2_pytest_gpu.log
for C in tum kitti euroc 1080p match; do
  timeout 300 python bench.py --config $C > gpurun_out/r2_bench_$C.json 2> gpurun_out/r2_bench_$C.err; echo "bench $C rc=$?"
done
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2_bench_reference.json 2> gpurun_out/r2_bench_reference.err; echo "reference rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches.csv \
    python bench.py --steps 2 --warmup 3 --batch 256 --value-only > gpurun_out/ncu_launch.log 2>&1; echo "launch list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:'fastStrip|gauss9|describe|octree|resizeLinear' \
    --launch-skip 33 --launch-count 11 -o gpurun_out/r2_extract -f python bench.py --steps 2 --warmup 3 --batch 256 --value-only > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep 2>/dev/null
python - <<'PY'
import json,glob
for f in sorted(glob.glob("gpurun_out/r2_bench_*.json")):
    try:
        j=json.loads(open(f).read().strip().splitlines()[-1])
        print(f, j.get("metric"), round(j.get("value",0),1), "e2e", round(j.get("e2e",{}).get("value",0),1), j.get("roofline",{}).get("stage_ms_per_step"), "cpu", j.get("cpu_baseline",{}).get("value"))
    except Exception as e: print(f, "unreadable", e)
PY
