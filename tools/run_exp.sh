timeout 300 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
run() { echo "== $*"; env "$@" python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/exp.err | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(j['ms_per_step'],4), j.get('fwd_ms'), j.get('bwd_ms'), {k:round(v['ms'],4) for k,v in j['stages'].items()}); print(j['e2e'], j['roofline'])"; }
run GSR_X=0
