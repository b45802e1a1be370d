# round 2, call S (the shipped build) (1 GPU): C3 at 100 M triangles with an oracle check on a ray subset (the oracle's binary BVH over 10^8
# primitives needs ~20 GB of host memory: only attempted when the box has >= 96 GB available)
export PB_MESH_CACHE=/tmp/meshcache
AVAIL=$(awk '/MemAvailable/ {print int($2/1048576)}' /proc/meminfo)
echo "host memory available: ${AVAIL} GB, cores: $(nproc)"
if [ "$AVAIL" -ge 96 ]; then CHECK="--check 200000 --brute 100"; else CHECK=""; echo "skipping the oracle check"; fi
timeout 1500 python raycast_bench.py --mesh 100M --rays 20000000 --incoherent 20000000 --no-sorted $CHECK > gpurun_out/c3_100m_r2s.jsonl 2> gpurun_out/c3_100m_r2s.err; echo "c3 100M rc=$?"
cut -c1-330 gpurun_out/c3_100m_r2s.jsonl; tail -3 gpurun_out/c3_100m_r2s.err
