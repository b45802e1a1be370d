# round 2, call Q (1 GPU): shared-memory stack depth 8 / 12 (shipped) / 16 entries per lane, same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/q_$name.json 2>gpurun_out/q_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/q_$name.err; }
}
rm -f gpurun_out/q_*.json
run base a_base
run st8 b_st8
run st16 c_st16
run base d_base2
run st8 e_st8_2
run st16 f_st16_2
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/q_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/q_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
