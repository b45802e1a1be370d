# round 2, call K (1 GPU): per-material occupancy of the shade kernels + traversal knobs re-measured on the cold-smem build
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/k_$name.json 2>gpurun_out/k_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/k_$name.err; }
}
rm -f gpurun_out/k_*.json
run base a_base
run sm9 b_smooth9
run sm10 c_smooth10
run sh6 d_rough6
run all9 e_all9
run base f_base2
run base g_iters3 PB_PRIM_ITERS=3
run base h_batch10 PB_PRIM_BATCH=10
run base i_batch6 PB_PRIM_BATCH=6
run base j_refill16 PB_REFILL=16
run base k_refill24 PB_REFILL=24
run base l_anyiters1 PB_PRIM_ITERS_ANY=1
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/k_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/k_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
