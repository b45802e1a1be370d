export PB_MESH_CACHE=/tmp/meshcache
(timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo pytest rc=$?; tail -3 gpurun_out/pytest_gpu.log)
for r in 16 20 24; do
  PB_REFILL=$r timeout 300 python bench.py --steps 3 --warmup 2 --e2e-steps 1 --no-cpu 2>/dev/null > gpurun_out/sweep_refill_$r.json
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_refill_$r.json"))
print("refill",$r,"Msamples/s %.1f"%d["value"],"Mrays/s %.1f"%d["mrays_per_s"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%d["roofline"]["frac"],"connect %.1f"%d["mrays_per_s_any_kernel"], "share %.2f"%d["roofline"]["share_of_step"])
PY
done
CMD="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --e2e-steps 1 --no-cpu"; timeout 600 $CMD > gpurun_out/plain.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 20 -c 2 -o gpurun_out/prof_extend_v2 $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu rc=$?
