# round 2, call J (1 GPU): cold ray state in shared memory (PB_COLD_SMEM) at 8 and 9 CTAs/SM against the default build, same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # name
  cp scratch/lib_$1.so $LIB
  python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/cold_$2.json 2>gpurun_out/cold_$2.err || { echo "$2 FAILED"; tail -3 gpurun_out/cold_$2.err; }
}
rm -f gpurun_out/cold_*.json
run final a_default
run cold8 b_cold8
run cold9 c_cold9
run final d_default2
run cold8 e_cold8_2
run cold9 f_cold9_2
for v in cold8 cold9; do
  cp scratch/lib_$v.so $LIB
  timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/cold_pytest_$v.log
  echo "$v pytest: $(tail -1 gpurun_out/cold_pytest_$v.log)"
done
cp scratch/lib_final.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/cold_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("cold_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
