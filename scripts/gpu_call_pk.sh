# per-kernel ncu metrics of one C4 wave (shipped build), folded by profiles/per_kernel.py
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-v18}
python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu > gpurun_out/pk_plain_$T.json 2>/dev/null && \
timeout 900 ncu --metrics gpu__time_duration.sum,smsp__thread_inst_executed_per_inst_executed.ratio,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,l1tex__t_sector_hit_rate.pct,lts__t_sector_hit_rate.pct,smsp__inst_executed.sum \
  --clock-control none -k regex:'k_extend|k_connect|k_shade|k_miss|k_generate|k_accumulate|k_after|k_prepare' -c 200 --csv --log-file gpurun_out/per_kernel_$T.csv \
  python bench.py --steps 1 --warmup 1 --e2e-steps 1 --no-cpu > gpurun_out/pk_ncu_$T.log 2>&1; echo "ncu rc=$?"
python profiles/per_kernel.py gpurun_out/per_kernel_$T.csv | tee gpurun_out/per_kernel_$T.md
