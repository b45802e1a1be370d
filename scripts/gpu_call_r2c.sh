# round 2, call C: same-box A/B of the traversal changes (libraries prebuilt in scratch/ by scratch/build_variant.sh),
# parity tests on the new build, ncu source-level captures of k_extend and two shade kernels
export PB_MESH_CACHE=/tmp/meshcache
LIB=pupiloptixlab_b200/csrc/libpupil_b200.so
run() {  # name
  python bench.py --no-cpu --steps 4 --e2e-steps 1 > gpurun_out/ab_$1.json 2>gpurun_out/ab_$1.err || { echo "$1 FAILED"; tail -3 gpurun_out/ab_$1.err; }
}
cp scratch/lib_sph_defer.so $LIB
(timeout 900 python -m pytest tests/test_gpu_trace.py tests/test_gpu_render.py tests/test_gpu_edge.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -4) | tee gpurun_out/pytest_r2c.log
run sph_defer
cp scratch/lib_base.so $LIB; run base
cp scratch/lib_sph_nodefer.so $LIB; run sph_nodefer
cp scratch/lib_sph_defer.so $LIB; run sph_defer2
for v in $EXTRA_VARIANTS; do cp scratch/lib_$v.so $LIB; run $v; done
cp scratch/lib_base.so $LIB; run base2
cp scratch/lib_sph_defer.so $LIB
python - <<PY
import json,glob
for f in sorted(glob.glob("gpurun_out/ab_*.json")):
    try:
        j=json.load(open(f)); r=j["roofline"]; ms=j["ms_per_step"]
        print(f.split("ab_")[1][:-5].ljust(20), "%.1f Ms/s  graph %.1f  step %.1f ms  ext %.2f  con %.2f  shade %.2f  kernel nodes/ray %.2f"%(j["value"], j["launch_path"]["graph"]["value"], ms, r["share_of_step"]*ms, j["roofline_connect"]["share_of_step"]*ms, j["shade_share_of_step"]*ms, r["kernel_counters"]["nodes_per_ray"]))
    except Exception as e:
        print(f, "unreadable", e)
PY
CMD="python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 1 -f -o gpurun_out/extend_r2c $CMD > gpurun_out/ncu_extend_r2c.log 2>&1; echo "extend capture rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_shade" -s 7 -c 7 -f -o gpurun_out/shade_r2c $CMD > gpurun_out/ncu_shade_r2c.log 2>&1; echo "shade capture rc=$?"
ls -la gpurun_out/*.ncu-rep
