# GPU call A: gpu tests, full C3 ray-cast microbench (1M / 10M / 100M triangles, 1e8 + 1e8 rays), per-kernel ncu metrics of one C4 wave
export PB_MESH_CACHE=/tmp/meshcache
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3) | tee gpurun_out/pytest_gpu_a.log
timeout 900 python raycast_bench.py --mesh 1M,10M --rays 100000000 --incoherent 100000000 --check 200000 --reps 3 > gpurun_out/c3_1m_10m.jsonl 2> gpurun_out/c3_1m_10m.err; echo "c3 small rc=$?"
timeout 1200 python raycast_bench.py --mesh 100M --rays 100000000 --incoherent 100000000 --reps 3 --no-sorted > gpurun_out/c3_100m.jsonl 2> gpurun_out/c3_100m.err; echo "c3 100M rc=$?"
tail -3 gpurun_out/c3_100m.err
python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu > gpurun_out/pk_plain.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum \
  --clock-control none -k regex:'k_extend|k_connect|k_shade|k_miss|k_generate|k_accumulate|k_after|k_prepare' -c 200 --csv --log-file gpurun_out/r01_per_kernel.csv \
  python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu > gpurun_out/pk_ncu.log 2>&1; echo "ncu rc=$?"
python profiles/per_kernel.py gpurun_out/r01_per_kernel.csv | tee gpurun_out/r01_per_kernel.md
cat gpurun_out/c3_1m_10m.jsonl gpurun_out/c3_100m.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'set' in d: print(d['mesh'], d['set'], '%.0f Mrays/s'%d['mrays_per_s'], 'frac %.3f'%d['roofline']['frac'], 'nodes/ray %.2f'%d['nodes_per_ray'])
    else: print(d)
"
