export PB_MESH_CACHE=/tmp/meshcache
run() { python bench.py --no-cpu --e2e-steps 1 > gpurun_out/ab_$1.json 2>/dev/null; }
run base; run base2
for v in "PB_CONNECT_RED=0" "PB_TRACE_PREFETCH=0" "PB_SHADE_PREFETCH=0"; do
  PB_NVCC_EXTRA="-D$v" python pupiloptixlab_b200/build.py --force > gpurun_out/build_$v.log 2>&1
  run $v
done
PB_NVCC_EXTRA="-DPB_SHADE_PREFETCH=0 -DPB_CONNECT_RED=0 -DPB_TRACE_PREFETCH=0" python pupiloptixlab_b200/build.py --force > /dev/null 2>&1
run alloff
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/ab_*.json")):
    j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
    print(f.split("ab_")[1][:-5].ljust(22), "%.1f Ms/s  step %.1f ms  ext %.1f  con %.1f  shade %.1f  other %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, ms*(1-r["share_of_step"]-j["roofline_connect"]["share_of_step"]-j["shade_share_of_step"])))
PY
