# 2-GPU check: the exchange parity test (all transports), then a short C5 bench line at N = 2
set -u
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_multiview_2gpu.py -x -q > gpurun_out/r02_pytest_2gpu_v14.log 2>&1; echo "test exit $?"
tail -3 gpurun_out/r02_pytest_2gpu_v14.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r02_bench_c5_n2_v14.json 2> gpurun_out/r02_bench_c5_n2_v14.err
python - <<PY
import json
j = json.loads(open("gpurun_out/r02_bench_c5_n2_v14.json").read().strip().splitlines()[-1])
print("N=2: ms/step %.4f value %.1f e2e %.1f" % (j["ms_per_step"], j["value"], j["e2e"]["value"]), "check", j.get("exchange_check", {}).get("ok"), j.get("exchange_check", {}).get("max_rel_l2"), "c4", j.get("c4_render", {}).get("ms_per_frame"), j.get("c4_render", {}).get("speedup_vs_n1"), j.get("c4_render", {}).get("bands_bit_identical_to_full_frame"))
PY
