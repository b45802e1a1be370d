# sweep PB_PRIM_ITERS x PB_PRIM_BATCH on C4 (one B200)
export PB_MESH_CACHE=/tmp/meshcache
for it in 1 2 3 999; do for pb in 6 8 12; do
  PB_PRIM_ITERS=$it PB_PRIM_BATCH=$pb python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/pi_${it}_${pb}.json 2>/dev/null
done; done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/pi_*.json")):
    j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
    print(f.split("pi_")[1][:-5].ljust(8), "%.1f Ms/s  step %.1f ms  ext %.1f  con %.1f  nodes/ray %.2f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, r["nodes_per_ray"]))
PY
