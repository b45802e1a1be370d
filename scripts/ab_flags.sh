# usage: bash scripts/ab_flags.sh "<nvcc -D flags of variant 1>" "<variant 2>" ...   (first run = default build as shipped)
export PB_MESH_CACHE=/tmp/meshcache
run() { python bench.py --no-cpu --e2e-steps 2 > "gpurun_out/ab_$1.json" 2>/dev/null; }
run base; run base2
n=0
for v in "$@"; do
  n=$((n+1))
  PB_NVCC_EXTRA="$v" python pupiloptixlab_b200/build.py --force > gpurun_out/build_$n.log 2>&1
  run "v$n"
done
python - "$@" <<PY
import json,glob,sys
names = {"v%d" % (i + 1): a for i, a in enumerate(sys.argv[1:])}
for f in sorted(glob.glob("gpurun_out/ab_*.json")):
    k = f.split("ab_")[1][:-5]
    j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
    print(k.ljust(6), names.get(k, "").ljust(28), "%.1f Ms/s  step %.1f ms  ext %.1f  con %.1f  shade %.1f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
PY
