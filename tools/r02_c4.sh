# C4 (render only, 6 M Gaussians, 3840x2160 = 32 400 tiles) with the radix and with the counting tile sort
set -u
mkdir -p gpurun_out
for MODE in radix count; do
  GSR_TILE_SORT=$MODE timeout 400 python bench.py --workload c4-render --steps 10 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_c4_$MODE.json 2> gpurun_out/r02_c4_$MODE.err
  python - <<PY
import json
j = json.loads(open("gpurun_out/r02_c4_$MODE.json").read().strip().splitlines()[-1])
print("$MODE: ms/frame %.4f" % j["ms_per_step"], {k: v["ms"] for k, v in j.get("stages", {}).items()})
PY
done
