# round 2, call U (1 GPU): the first primitive record a lane will test is requested when the node visit finds it (prefetch.global.L1 / .L2), same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/u_$name.json 2>gpurun_out/u_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/u_$name.err; }
}
rm -f gpurun_out/u_*.json
run base a_base
run pp1 b_ppL1
run pp2 c_ppL2
run base d_base2
run pp1 e_ppL1_2
run pp2 f_ppL2_2
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/u_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/u_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
