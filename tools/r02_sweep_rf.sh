# Sweep the GSR_RF_VARIANT choices of the blending forward: parity vs the reference build + stage times.
set -u
mkdir -p gpurun_out
for v in ${VARIANTS:-0 1 2}; do
  echo "== GSR_RF_VARIANT=$v"
  GSR_RF_VARIANT=$v timeout 300 python -m pytest tests/test_parity_gpu.py -m gpu -x -q 2>&1 | tail -1
  GSR_RF_VARIANT=$v timeout 200 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-extras 2> gpurun_out/exp.err | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['ms_per_step'],4), j.get('fwd_ms'), j.get('bwd_ms'), {k:round(v['ms'],4) for k,v in j['stages'].items() if k in ('render','render_bwd','preprocess_bwd')})"
done 2>&1 | tee gpurun_out/r02_sweep_rf.log
