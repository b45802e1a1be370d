# round 2, call E: no-alloc test alone, whole gpu suite, bench (both arms), compute-sanitizer, ncu launch list + captures, C3
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-r2e}
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q 2>&1 | tail -3) | tee gpurun_out/pytest_noalloc_alone_$T.log
(timeout 1500 python -m pytest tests -m gpu -q -rA 2>&1 | grep -v "^PASSED\|^$" | tail -120) > gpurun_out/pytest_gpu_$T.log; tail -8 gpurun_out/pytest_gpu_$T.log
timeout 400 python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
timeout 200 python bench.py --impl reference > gpurun_out/bench_reference_$T.json 2> /dev/null; echo "reference rc=$?"
# compute-sanitizer is closed on this pool (scripts/sanitize_gpu.sh exits 86 there): see DESIGN.md §1
CMD="python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu"
$CMD > gpurun_out/plain_$T.json 2>/dev/null && \
timeout 300 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$T.csv $CMD > gpurun_out/ncu_launches_$T.log 2>&1; echo "launch list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 1 -f -o gpurun_out/extend_$T $CMD > gpurun_out/ncu_extend_$T.log 2>&1; echo "extend capture rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:k_shade -s 7 -c 7 -f -o gpurun_out/shade_$T $CMD > gpurun_out/ncu_shade_$T.log 2>&1; echo "shade capture rc=$?"
ncu -i gpurun_out/extend_$T.ncu-rep --page raw --csv > gpurun_out/extend_${T}_raw.csv 2>/dev/null
ncu -i gpurun_out/extend_$T.ncu-rep --page source --csv > gpurun_out/extend_${T}_src.csv 2>/dev/null
ncu -i gpurun_out/shade_$T.ncu-rep --page raw --csv > gpurun_out/shade_${T}_raw.csv 2>/dev/null
timeout 600 python raycast_bench.py --mesh 10M --rays 100000000 --incoherent 100000000 --spheres 10000 --check 10000000 --brute 2000 > gpurun_out/c3_10m_$T.jsonl 2> gpurun_out/c3_10m_$T.err; echo "c3 rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$T.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%r["frac"],"bmin",r.get("b_min_frac"),"B/ray %.0f"%r["bytes_per_ray"],"cpu",(d.get("cpu_baseline") or {}).get("value"),d["clocks"])
PY
