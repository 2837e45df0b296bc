set -u
mkdir -p gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/r02_bench_c5_n4_v14.json 2> gpurun_out/r02_bench_c5_n4_v14.err
python - <<PY
import json
j = json.loads(open("gpurun_out/r02_bench_c5_n4_v14.json").read().strip().splitlines()[-1])
print("N=4: ms/step %.4f value %.1f e2e %.1f" % (j["ms_per_step"], j["value"], j["e2e"]["value"]), "check", j.get("exchange_check", {}).get("ok"), "c4", j.get("c4_render", {}).get("ms_per_frame"), j.get("c4_render", {}).get("speedup_vs_n1"))
PY
