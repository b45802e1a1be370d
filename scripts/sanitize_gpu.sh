# compute-sanitizer over the product's kernels (SURVEY §5 sanitizer targets): Cornell with all materials (every shade kernel,
# AOVs), refit of a sphere and of a light, the edge scene, env map, textures, ray casts - under memcheck / racecheck /
# initcheck / synccheck.  Usage on the GPU box: bash scripts/sanitize_gpu.sh [out_dir]
OUT=${1:-gpurun_out}
mkdir -p $OUT
cat > /tmp/san_case.py <<'PY'
import numpy as np, sys
sys.path.insert(0, ".")
from pupiloptixlab_b200 import capi, scenes
api = capi.load()
g = capi.Context(api, 0)
info = scenes.cornell_box(g, 64, 64, "materials")
g.render_ex(64, 64, 6, 0, 2)
m = np.eye(4, dtype=np.float32); m[:3, :3] *= 0.2; m[:3, 3] = (0.1, 0.8, 0.2)
g.update_instance(info["objects"]["sphere1"], m[:3])
g.update_instance(info["objects"]["bulb"], (scenes.translate(0.5, 1.5, 0.3) @ scenes.scale(0.05, 0.05, 0.05))[:3])
g.render(64, 64, 6, 2, 2)
scenes.edge_scene(g, 48, 36)
g.render_ex(48, 36, 8, 0, 2)
scenes.cornell_box(g, 40, 40, "envmap"); g.render(40, 40, 5, 0, 2)
scenes.cornell_box(g, 40, 40, "textured"); g.render(40, 40, 5, 0, 2)
rays = np.zeros(5000, capi.RAY_DTYPE)
rng = np.random.RandomState(0)
d = rng.normal(size=(5000, 3)); d /= np.linalg.norm(d, axis=1, keepdims=True)
rays["ox"], rays["oy"], rays["oz"] = 0, 1, 0
rays["dx"], rays["dy"], rays["dz"] = d.T.astype(np.float32); rays["tmin"], rays["tmax"] = 1e-3, 1e16
g.trace_closest(rays); g.trace_any(rays); g.trace_closest(rays, counted=True)
print("san_case ok")
PY
for tool in memcheck racecheck initcheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 9 --print-limit 20 python /tmp/san_case.py > $OUT/sanitizer_$tool.log 2>&1
  rc=$?
  echo "$tool rc=$rc: $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $OUT/sanitizer_$tool.log | tail -1)"
done
