# one bench line per environment setting: bash tools/r02_var.sh TAG "ENV=1 ENV2=2" ...
set -u
mkdir -p gpurun_out
while [ $# -ge 2 ]; do
  TAG=$1; ENVS=$2; shift 2
  env $ENVS timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r02_var_$TAG.json 2> gpurun_out/r02_var_$TAG.err
  python - <<PY
import json
j = json.loads(open("gpurun_out/r02_var_$TAG.json").read().strip().splitlines()[-1])
print("$TAG: ms/step %.4f" % j["ms_per_step"], {k: v["ms"] for k, v in j["stages"].items() if k in ("depth_sort", "tile_sort", "preprocess", "preprocess_bwd")})
PY
done
