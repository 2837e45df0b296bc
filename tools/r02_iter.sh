# Round 2 kernel iteration: rasterizer parity on the GPU, then one bench line (no CPU baseline leg).  TAG names the outputs.
set -u
TAG=${1:-iter}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02_pytest_parity_$TAG.log
timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras > gpurun_out/r02_bench_c2_$TAG.json 2> gpurun_out/r02_bench_c2_$TAG.err; echo "bench rc=$?"
python - <<PY
import json
j = json.loads(open("gpurun_out/r02_bench_c2_$TAG.json").read().strip().splitlines()[-1])
print("ms/step %.4f fwd %.4f bwd %.4f e2e %.4f" % (j["ms_per_step"], j["fwd_ms"], j["bwd_ms"], j["e2e"]["ms_per_step"]))
print({k: v["ms"] for k, v in j["stages"].items()})
PY
