# same-box A/B of the node-test variants on C4 (one B200): baseline library built from the previous commit
# (scratch/libpupil_b200_base.so), then rebuilds of the working tree with different switches
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # name
  python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/ab_$1.json 2>gpurun_out/ab_$1.err || { echo "$1 FAILED"; tail -3 gpurun_out/ab_$1.err; }
}
cp $LIB /tmp/new_default.so
timeout 600 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py -x -q 2>&1 | tail -3
run new_fma2_mb8
cp scratch/libpupil_b200_base.so $LIB; run base
cp scratch/libpupil_b200_base.so $LIB; run base2
PB_NVCC_EXTRA="-DPB_TRACE_MIN_BLOCKS=7" python -m pupiloptixlab_b200.build --force > /dev/null 2>&1; grep -A2 "k_extendILb0" pupiloptixlab_b200/csrc/pb_wavefront.ptxas.log | grep -E "Used|spill"; run new_fma2_mb7
PB_NVCC_EXTRA="-DPB_NODE_FMA2=0" python -m pupiloptixlab_b200.build --force > /dev/null 2>&1; grep -A2 "k_extendILb0" pupiloptixlab_b200/csrc/pb_wavefront.ptxas.log | grep -E "Used|spill"; run new_nofma2_mb8
cp /tmp/new_default.so $LIB; run new_fma2_mb8_again
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/ab_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("ab_")[1][:-5].ljust(20), "%.1f Ms/s  step %.1f ms  ext %.1f  con %.1f  shade %.1f  nodes/ray %.2f"%(j["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, r["nodes_per_ray"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
