for mode in multimem; do
GSR_MULTIVIEW_SMALL=$mode GSR_MV_TIMELINE=1 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 --no-cpu-baseline 2>gpurun_out/n2_$mode.err | tail -1 | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('N=2 $mode', round(j['ms_per_step'],4), round(j['value'],1), j['e2e']['ms_per_step'])"
grep -i "timeline\|unavailable" gpurun_out/n2_$mode.err | head -2
done
