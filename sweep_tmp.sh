export PB_MESH_CACHE=/tmp/meshcache
(timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2)
timeout 600 python bench.py --no-cpu 2>gpurun_out/bench_v5.err > gpurun_out/bench_v5.json; python - <<PY
import json
d=json.load(open("gpurun_out/bench_v5.json")); r=d["roofline"]
print("Msamples/s %.1f"%d["value"],"Mrays/s %.1f"%d["mrays_per_s"],"extend Mrays/s %.1f"%d["mrays_per_s_closest_kernel"],"frac %.3f"%r["frac"],"connect %.1f"%d["mrays_per_s_any_kernel"],"share ext %.3f con %.3f"%(r["share_of_step"], d["roofline_connect"]["share_of_step"]),"n/ray %.2f"%r["nodes_per_ray"], "prims/ray %.2f"%(r["tris_per_ray"]+r["spheres_per_ray"]))
PY
timeout 600 python raycast_bench.py --check 100000 2>&1 | tee gpurun_out/raycast_v5.log | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'set' in d: print(d['set'], '%.0f Mrays/s'%d['mrays_per_s'], 'frac %.3f'%d['roofline']['frac'], 'nodes/ray %.2f'%d['nodes_per_ray'])
    else: print(d)
"
