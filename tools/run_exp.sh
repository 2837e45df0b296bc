timeout 200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { echo "== $*"; env "$@" timeout 120 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/exp.err | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['ms_per_step'],4), j.get('fwd_ms'), j.get('bwd_ms'), {k:round(v['ms'],4) for k,v in j['stages'].items()})"; }
run GSR_FUSED_SCAN=0
run GSR_RF_MINB=8 GSR_RB_MINB=8
run GSR_RF_MINB=5 GSR_RB_MINB=6
