set -u
mkdir -p gpurun_out
TAG=${1:-n8}
shift || true
env "$@" GSR_MV_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r02_n8_$TAG.json 2> gpurun_out/r02_n8_$TAG.err
python - <<PY
import json
j = json.loads(open("gpurun_out/r02_n8_$TAG.json").read().strip().splitlines()[-1])
print("$TAG", "ms/step %.4f value %.1f e2e %.1f" % (j["ms_per_step"], j["value"], j["e2e"]["value"]), "check", j.get("exchange_check", {}).get("ok"), j.get("exchange_check", {}).get("max_rel_l2"))
PY
grep "exchange timeline" gpurun_out/r02_n8_$TAG.err | tail -1
