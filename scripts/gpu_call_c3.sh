# C3 at BASELINE's named sizes for the shipped build: 1 M / 10 M / 100 M triangles, 1e8 coherent + 1e8 incoherent rays
export PB_MESH_CACHE=/tmp/meshcache
T=${1:-v18}
timeout 900 python raycast_bench.py --mesh 1M,10M --rays 100000000 --incoherent 100000000 --check 200000 --reps 3 > gpurun_out/c3_1m_10m_$T.jsonl 2> gpurun_out/c3_1m_10m_$T.err; echo "c3 small rc=$?"
timeout 1500 python raycast_bench.py --mesh 100M --rays 100000000 --incoherent 100000000 --reps 3 --no-sorted > gpurun_out/c3_100m_$T.jsonl 2> gpurun_out/c3_100m_$T.err; echo "c3 100M rc=$?"
tail -3 gpurun_out/c3_100m_$T.err
cat gpurun_out/c3_1m_10m_$T.jsonl gpurun_out/c3_100m_$T.jsonl | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l)
    if 'set' in d: print(d['mesh'], d['set'], '%.0f Mrays/s'%d['mrays_per_s'], 'frac %.3f'%d['roofline']['frac'], 'nodes/ray %.2f'%d['nodes_per_ray'])
    else: print(d)
"
