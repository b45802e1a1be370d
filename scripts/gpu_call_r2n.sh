# round 2, call N (1 GPU): triangle test before its slab (PB_TRI_TEST_FIRST) against the shipped build, same box
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # lib name [env...]
  lib=$1; name=$2; shift; shift
  cp scratch/lib_$lib.so $LIB
  env X=1 "$@" python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/n_$name.json 2>gpurun_out/n_$name.err || { echo "$name FAILED"; tail -3 gpurun_out/n_$name.err; }
}
rm -f gpurun_out/n_*.json
run base a_base
run trifirst b_trifirst
run base c_base2
run trifirst d_trifirst2
cp scratch/lib_trifirst.so $LIB
timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q -m gpu 2>&1 | tail -3 > gpurun_out/n_pytest_trifirst.log
echo "trifirst pytest: $(tail -1 gpurun_out/n_pytest_trifirst.log)"
timeout 300 python raycast_bench.py --mesh 10M --rays 100000000 --incoherent 100000000 --spheres 10000 --check 2000000 --brute 1000 > gpurun_out/n_c3_trifirst.jsonl 2> gpurun_out/n_c3_trifirst.err; echo "c3 rc=$?"
cp scratch/lib_base.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/n_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("/n_")[1][:-5].ljust(14), "%.1f Ms/s  step %.2f ms  ext %.2f  con %.2f  shade %.2f  e2e %.1f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, j["e2e"]["value"]))
    except Exception as e:
        print(f, "unreadable", e)
for l in open("gpurun_out/n_c3_trifirst.jsonl"):
    j=json.loads(l)
    if "set" in j: print(j["set"], round(j["mrays_per_s"]))
    elif "check" in j: print(j["check"], {k:v for k,v in j.items() if "mism" in k or k=="rays"})
PY
