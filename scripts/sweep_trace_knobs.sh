# sweep of the traversal knobs on C4 (one B200): PB_PRIM_BATCH x PB_REFILL
export PB_MESH_CACHE=/tmp/meshcache
for pb in 6 8 10 12 16; do for rf in 20; do
  PB_PRIM_BATCH=$pb PB_REFILL=$rf python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/sw_${pb}_${rf}.json 2>/dev/null
done; done
for rf in 12 16 24 28; do
  PB_PRIM_BATCH=8 PB_REFILL=$rf python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/sw_8_${rf}.json 2>/dev/null
done
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/sw_*.json")):
    j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
    print(f.split("sw_")[1][:-5].ljust(8), "%.1f Ms/s  step %.1f ms  ext %.1f  con %.1f  shade %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms))
PY
