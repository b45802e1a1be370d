# round 2, call O (1 GPU): triangle test first for closest-hit rays only, and the early "behind the best hit" reject, same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/o_$name.json 2>gpurun_out/o_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/o_$name.err; }
}
rm -f gpurun_out/o_*.json
run base a_base
run tf b_tf
run tftb c_tftb
run base d_base2
run tf e_tf2
run tftb f_tftb2
for v in tf tftb; do
  cp scratch/lib_$v.so $LIB
  timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/o_pytest_$v.log
  echo "$v pytest: $(tail -1 gpurun_out/o_pytest_$v.log)"
done
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/o_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/o_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
