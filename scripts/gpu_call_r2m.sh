# round 2, call M (gpurun --gpus N): the shipped build on N GPUs - multi-GPU tests (Python + the C++ host), default bench, C5 as configured
export PB_MESH_CACHE=/tmp/meshcache
N=${1:-8}; T=${2:-r2m}
(timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_zz_no_alloc.py -m gpu -q -rA 2>&1 | grep -v "^$" | tail -12) | tee gpurun_out/pytest_multi_n${N}_$T.log
timeout 500 python bench.py --gpus $N > gpurun_out/scale_n${N}_$T.json 2> gpurun_out/scale_n${N}_$T.err; echo "bench n$N rc=$?"
STEPS=$((8 / N)); [ $STEPS -lt 1 ] && STEPS=1
timeout 600 python bench.py --gpus $N --film 4k --spp-per-step 128 --steps $STEPS --warmup 1 --e2e-steps 1 > gpurun_out/c5_n${N}_$T.json 2> gpurun_out/c5_n${N}_$T.err; echo "c5 n$N rc=$?"
python - <<PY
import json
for f in ("scale_n${N}_$T","c5_n${N}_$T"):
    try:
        d=json.load(open("gpurun_out/"+f+".json"))
        print(f,"Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"ms/step %.2f"%d["ms_per_step"],"parity",d.get("parity_check"))
    except Exception as e:
        print(f,"failed",e)
PY
