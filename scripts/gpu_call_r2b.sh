# round 2, call B (2 GPUs): whole gpu suite incl. the multi-GPU tests, 1-GPU bench with the oracle-derived roofline,
# 2-GPU bench (library reduce + parity_check), C5 film on 2 GPUs
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-r2b}
(timeout 1200 python -m pytest tests -m gpu -q -rA 2>&1 | grep -v "^PASSED\|^$" | tail -70) | tee gpurun_out/pytest_gpu_$T.log
timeout 400 python bench.py > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; echo "bench rc=$?"
timeout 400 python bench.py --gpus 2 --steps 4 --e2e-steps 2 > gpurun_out/scale_n2_$T.json 2> gpurun_out/scale_n2_$T.err; echo "bench n2 rc=$?"
timeout 400 python bench.py --gpus 2 --film 4k --spp-per-step 128 --steps 2 --warmup 1 --e2e-steps 1 > gpurun_out/c5_n2_$T.json 2> gpurun_out/c5_n2_$T.err; echo "c5 n2 rc=$?"
python - <<PY
import json
for f in ("bench_$T","scale_n2_$T","c5_n2_$T"):
    try:
        d=json.load(open("gpurun_out/"+f+".json")); r=d["roofline"]
        print(f,"Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"frac %.3f"%r["frac"],"bmin",r.get("b_min_frac"),"dram",r.get("dram_frac"),"issue",(r.get("issue") or {}).get("frac"),"B/ray %.0f"%r["bytes_per_ray"], "kernel B/ray %.0f"%r["kernel_counters"]["bytes_per_ray"], "parity",d.get("parity_check"))
    except Exception as e:
        print(f,"failed",e)
PY
tail -3 gpurun_out/*_$T.err
