# ncu --set full capture of ONE launch of one kernel of the C2 bench step; summaries land in gpurun_out/.
# usage: bash tools/r02_ncu_kernel.sh <kernel-regex> <tag>
set -u
K=$1; TAG=$2
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1; echo "plain run rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"$K" -s 3 -c 1 -o /tmp/prof_$TAG -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_ncu_$TAG.log 2>&1; echo "ncu rc=$?"
ncu -i /tmp/prof_$TAG.ncu-rep --page raw --csv > gpurun_out/r02_ncu_${TAG}_raw.csv 2>/dev/null
ncu -i /tmp/prof_$TAG.ncu-rep --page source --csv > gpurun_out/r02_ncu_${TAG}_source.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_ncu_${TAG}_raw.csv pipe_fma pipe_alu pipe_xu inst_executed_pipe smsp__warps_eligible smsp__warps_active lsu_mem_shared > gpurun_out/r02_ncu_${TAG}_summary.txt
wc -c gpurun_out/r02_ncu_${TAG}_*
