# round 2, call G (1 GPU): gpu suite on the current build, then a same-box sweep of the runtime traversal knobs
# (closest-hit and any-hit separately since round 2)
export PB_MESH_CACHE=/tmp/meshcache
(timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4) | tee gpurun_out/pytest_gpu_r2g.log
run() {  # name [env...]
  name=$1; shift
  env "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/sw_$name.json 2>gpurun_out/sw_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/sw_$name.err; }
}
rm -f gpurun_out/sw_*.json
run default X=1
run iters3 PB_PRIM_ITERS=3
run batch12 PB_PRIM_BATCH=12
run refill16 PB_REFILL=16
run refill22 PB_REFILL=22
run any_iters1 PB_PRIM_ITERS_ANY=1
run any_batch4 PB_PRIM_BATCH_ANY=4
run any_batch4_iters1 PB_PRIM_BATCH_ANY=4 PB_PRIM_ITERS_ANY=1
run any_batch12 PB_PRIM_BATCH_ANY=12
run any_refill26 PB_REFILL_ANY=26
run any_refill14 PB_REFILL_ANY=14
run default2 X=1
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/sw_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("sw_")[1][:-5].ljust(20), "%.1f Ms/s  graph %.1f  step %.1f ms  ext %.2f  con %.2f  shade %.2f"%(j["value"], j["launch_path"]["graph"]["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms))
    except Exception as e:
        print(f, "unreadable", e)
PY
