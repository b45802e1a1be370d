# Lean final evidence (fits a few GPU-minutes): gpu tests, both bench arms, ncu launch list + one --set full capture of
# k_extend (each ncu pass after the same command exited 0 without ncu)
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-v20}
(timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -2) | tee gpurun_out/pytest_gpu_$T.log
timeout 200 python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
timeout 120 python bench.py --impl reference > gpurun_out/bench_reference_$T.json 2> /dev/null; echo "reference rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --e2e-steps 1 --no-cpu"
$CMD > gpurun_out/plain_$T.json 2>/dev/null && \
timeout 120 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$T.csv $CMD > gpurun_out/ncu_launches_$T.log 2>&1; echo "launch list rc=$?"
timeout 150 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 1 -f -o gpurun_out/extend_$T $CMD > gpurun_out/ncu_full_$T.log 2>&1; echo "full capture rc=$?"
ncu -i gpurun_out/extend_$T.ncu-rep --page raw --csv > gpurun_out/extend_${T}_raw.csv 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$T.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"Mrays/s %.1f"%d["mrays_per_s"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%r["frac"],"share ext %.3f"%r["share_of_step"], "cpu", d.get("cpu_baseline",{}).get("value"), d["clocks"], d["gpu_ms_each_step"])
print(open("gpurun_out/bench_reference_$T.json").read()[:300])
PY
