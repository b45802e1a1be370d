# round 2, call F (gpurun --gpus N): multi-GPU tests, the default bench on N GPUs, C5 as configured (3840x2160, 1024 spp in
# total, sample-sharded, 133 MB reduce per step) and a small-work point (4 spp per GPU per step) where the reduce and the
# launch overhead become visible
export PB_MESH_CACHE=/tmp/meshcache
N=${1:-2}; T=${2:-r2f}
[ "$3" = "notest" ] || (timeout 900 python -m pytest tests/test_gpu_multi.py tests/test_gpu_render.py -m gpu -q -rA 2>&1 | grep -v "^$" | tail -15) | tee gpurun_out/pytest_multi_n${N}_$T.log
timeout 500 python bench.py --gpus $N > gpurun_out/scale_n${N}_$T.json 2> gpurun_out/scale_n${N}_$T.err; echo "bench n$N rc=$?"
STEPS=$((8 / N)); [ $STEPS -lt 1 ] && STEPS=1
timeout 600 python bench.py --gpus $N --film 4k --spp-per-step 128 --steps $STEPS --warmup 1 --e2e-steps 1 > gpurun_out/c5_n${N}_$T.json 2> gpurun_out/c5_n${N}_$T.err; echo "c5 n$N rc=$?"
timeout 400 python bench.py --gpus $N --spp-per-step 4 --steps 16 --warmup 3 --e2e-steps 4 > gpurun_out/small_n${N}_$T.json 2> gpurun_out/small_n${N}_$T.err; echo "small n$N rc=$?"
python - <<PY
import json
for f in ("scale_n${N}_$T","c5_n${N}_$T","small_n${N}_$T"):
    try:
        d=json.load(open("gpurun_out/"+f+".json"))
        print(f,"Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"ms/step %.2f"%d["ms_per_step"],"parity",d.get("parity_check"))
    except Exception as e:
        print(f,"failed",e)
PY
