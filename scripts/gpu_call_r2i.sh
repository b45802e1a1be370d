# round 2, call I (1 GPU): tree-shape knobs re-measured now that sphere tests are cheap (results never depend on the tree)
export PB_MESH_CACHE=/tmp/meshcache
run() {  # name [env...]
  name=$1; shift
  env "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/tr_$name.json 2>gpurun_out/tr_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/tr_$name.err; }
}
rm -f gpurun_out/tr_*.json
run default X=1
run leaf3 PB_LEAF_MAX=3
run leaf3_iters3 PB_LEAF_MAX=3 PB_PRIM_ITERS=3
run leaf3_bottom16 PB_LEAF_MAX=3 PB_BOTTOM_MAX=16
run leaf3_bottom24 PB_LEAF_MAX=3 PB_BOTTOM_MAX=24
run leaf2_bottom16 PB_LEAF_MAX=2 PB_BOTTOM_MAX=16
run leaf1 PB_LEAF_MAX=1
run default2 X=1
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/tr_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]; k=r["kernel_counters"]
        print(f.split("tr_")[1][:-5].ljust(18), "%.1f Ms/s  step %.1f ms  ext %.2f  con %.2f  shade %.2f | nodes/ray %.2f tris %.2f spheres %.2f  bvh nodes %d build %.0f ms"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, k["nodes_per_ray"], k["tris_per_ray"], k["spheres_per_ray"], j["bvh_nodes"], j["bvh_build_ms"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
