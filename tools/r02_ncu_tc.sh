# launch times of the counting tile sort kernels + one full capture of the scatter kernel
set -u
mkdir -p gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1; echo "plain run rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"tc_" -s 40 -c 24 --csv --log-file gpurun_out/r02_tc_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras > /dev/null 2>&1; echo "launch list rc=$?"
grep -o '"[a-z_:]*\(tc_[a-z]*\|duplicate[a-z_]*\|rs_[a-z_]*\)[^"]*".*' gpurun_out/r02_tc_launches.csv | awk -F'","' '{print $1, $NF}' | sort | uniq -c | head -30
bash tools/r02_ncu_kernel.sh tc_scatter_kernel tc_scatter
