# round 2, call A: the widened parity suite + a short bench on the same build
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-r2a}
(timeout 900 python -m pytest tests -m gpu -q -rA 2>&1 | grep -v "^PASSED\|^$" | tail -60) | tee gpurun_out/pytest_gpu_$T.log
timeout 200 python bench.py --steps 4 --warmup 3 --e2e-steps 2 > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$T.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%r["frac"], d["clocks"])
PY
