# 2-GPU check of the multicast transport: the parity test, then a short bench with either transport
set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multiview_2gpu.py -x -q -s > gpurun_out/r02_mc2_test.log 2>&1; echo "test exit $?"
tail -5 gpurun_out/r02_mc2_test.log
for MODE in allgather multicast; do
  GSR_MULTIVIEW_MODE=$MODE GSR_MV_TIMELINE=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 --no-extras > gpurun_out/r02_mc2_$MODE.json 2> gpurun_out/r02_mc2_$MODE.err
  python - <<PY
import json
j = json.loads(open("gpurun_out/r02_mc2_$MODE.json").read().strip().splitlines()[-1])
print("$MODE", "ms/step %.4f value %.1f e2e %.1f" % (j["ms_per_step"], j["value"], j["e2e"]["value"]), "check", j.get("exchange_check", {}).get("ok"), j.get("exchange_check", {}).get("max_rel_l2"), j.get("exchange_check", {}).get("transport"))
PY
  grep "exchange timeline" gpurun_out/r02_mc2_$MODE.err | tail -1
  grep -i "unavailable" gpurun_out/r02_mc2_$MODE.err | head -2
done
