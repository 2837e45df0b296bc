# 8 GPUs: the C5 step under a few NCCL settings (does the all-reduce / all-gather use NVLS? does forcing it help?)
set -u
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" GSR_MV_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 8 --steps 20 --warmup 5 --no-extras > gpurun_out/r02_n8_$tag.json 2> gpurun_out/r02_n8_$tag.err
  python - <<PY
import json
try:
    j = json.loads(open("gpurun_out/r02_n8_$tag.json").read().strip().splitlines()[-1])
    print("$tag", "ms/step %.4f value %.1f e2e %.1f" % (j["ms_per_step"], j["value"], j["e2e"]["value"]), "check", j.get("exchange_check", {}).get("ok"))
except Exception as e:
    print("$tag failed", e)
PY
  grep "exchange timeline" gpurun_out/r02_n8_$tag.err | tail -1
}
run default X=1
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT,COLL timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 8 --steps 2 --warmup 3 --no-extras --no-exchange-check 2>&1 | grep -i -E "nvls|algo|channels|multicast" | sort | uniq -c | sort -rn | head -20 > gpurun_out/r02_n8_nccl_info.txt; head -20 gpurun_out/r02_n8_nccl_info.txt
run nvls NCCL_ALGO=NVLS
run ring NCCL_ALGO=Ring
run tree_nvlstree NCCL_ALGO=NVLSTree
