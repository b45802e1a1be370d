# Final round-1 evidence for the shipped code: gpu tests, both bench arms, C3 at 10 M triangles with the brute-force
# check, ncu launch list + one --set full capture of k_extend (each after the same command exited 0 without ncu)
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-v18}
(timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) | tee gpurun_out/pytest_gpu_$T.log
timeout 600 python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/bench_reference_$T.json 2> gpurun_out/bench_reference_$T.err; echo "reference rc=$?"
timeout 900 python raycast_bench.py --mesh 10M --rays 100000000 --incoherent 100000000 --check 200000 --reps 3 > gpurun_out/c3_10m_$T.jsonl 2> gpurun_out/c3_10m_$T.err; echo "c3 rc=$?"
CMD="python bench.py --steps 1 --warmup 1 --spp-per-step 4 --e2e-steps 1 --no-cpu"
$CMD > gpurun_out/plain_$T.json 2>/dev/null && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_$T.csv $CMD > gpurun_out/ncu_launches_$T.log 2>&1; echo "launch list rc=$?"
CMD2="python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu"
$CMD2 > gpurun_out/plain2_$T.json 2>/dev/null && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_extend -s 1 -c 1 -f -o gpurun_out/extend_$T $CMD2 > gpurun_out/ncu_full_$T.log 2>&1; echo "full capture rc=$?"
ncu -i gpurun_out/extend_$T.ncu-rep --page raw --csv > gpurun_out/extend_${T}_raw.csv 2>/dev/null
python - <<PY
import json
d=json.load(open("gpurun_out/bench_$T.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"Mrays/s %.1f"%d["mrays_per_s"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%r["frac"],"share ext %.3f"%r["share_of_step"],"n/ray %.2f"%r["nodes_per_ray"], "cpu", d.get("cpu_baseline",{}).get("value"), d["clocks"])
print(open("gpurun_out/bench_reference_$T.json").read()[:600])
for l in open("gpurun_out/c3_10m_$T.jsonl"):
    j=json.loads(l)
    if 'set' in j: print(j['set'], '%.0f Mrays/s'%j['mrays_per_s'], 'frac %.3f'%j['roofline']['frac'], 'nodes/ray %.2f'%j['nodes_per_ray'])
    else: print(j)
PY
