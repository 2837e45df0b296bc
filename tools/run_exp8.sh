N=$1
shift
for mode in "$@"; do
GSR_MULTIVIEW_SMALL=$mode GSR_MV_TIMELINE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/n${N}_$mode.err | tail -1 > gpurun_out/bench_n${N}_$mode.json
python -c "
import json,sys
j=json.loads(open('gpurun_out/bench_n${N}_$mode.json').read().strip().splitlines()[-1]); print('N=$N $mode', round(j['ms_per_step'],4), round(j['value'],1), j['e2e']['ms_per_step'])"
grep -i "timeline\|unavailable" gpurun_out/n${N}_$mode.err | head -2
done
