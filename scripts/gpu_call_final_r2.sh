# round 2, last call (1 GPU): what the driver runs at round end, on the final tree - gpu suite with -x, smoke(), default bench
export PB_MESH_CACHE=/tmp/meshcache
(timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4) | tee gpurun_out/pytest_gpu_final_r2.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -3 | tee gpurun_out/smoke_final_r2.log
timeout 400 python bench.py > gpurun_out/bench_final_r2.json 2> gpurun_out/bench_final_r2.err; echo "bench rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_final_r2.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"frac %.3f"%r["frac"],"launches",d["gpu_launches"],d["clocks"]["reasons"])
PY
