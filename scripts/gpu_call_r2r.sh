# round 2, call R (1 GPU): shade kernels request the next iteration's path record one iteration ahead (prefetch.global.L2 / .L1), same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/r_$name.json 2>gpurun_out/r_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/r_$name.err; }
}
rm -f gpurun_out/r_*.json
run base a_base
run pf2 b_pfL2
run pf1 c_pfL1
run base d_base2
run pf2 e_pfL2_2
run pf1 f_pfL1_2
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/r_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/r_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
