# round 2, call T (gpurun --gpus N): the default bench of the shipped build on N GPUs (scaling series 2 / 4 / 8)
export PB_MESH_CACHE=/tmp/meshcache
N=${1:-2}; T=${2:-r2t}
timeout 500 python bench.py --gpus $N > gpurun_out/scale_n${N}_$T.json 2> gpurun_out/scale_n${N}_$T.err; echo "bench n$N rc=$?"
python - <<PY
import json
d=json.load(open("gpurun_out/scale_n${N}_$T.json"))
print("n$N","Msamples/s %.1f"%d["value"],"e2e %.1f"%d["e2e"]["value"],"graph %.1f"%d["launch_path"]["graph"]["value"],"ms/step %.2f"%d["ms_per_step"],"parity",d.get("parity_check",{}).get("max_rel"))
PY
