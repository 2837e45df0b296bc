# Round-end measurement on ONE B200: bench line, ncu launch list and one full capture (exported to CSV on the box:
# the .ncu-rep with sources is larger than what gpurun copies back).
set -u
mkdir -p gpurun_out
timeout 300 python bench.py > gpurun_out/bench_b200_v12.json 2> gpurun_out/bench_b200_v12.err; echo "bench rc=$?"
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_v12.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches_v12.log 2>&1; echo "ncu list rc=$?"
timeout 400 ncu --set full --clock-control none --import-source on -c 44 -o /tmp/prof_v12 -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full_v12.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/prof_v12.ncu-rep --page raw --csv > gpurun_out/prof_v12_raw.csv 2>/dev/null
ncu -i /tmp/prof_v12.ncu-rep --page source --csv --kernel-name regex:render_bwd_kernel > gpurun_out/src_bwd_v12.csv 2>/dev/null
ncu -i /tmp/prof_v12.ncu-rep --page source --csv --kernel-name regex:render_fwd_kernel > gpurun_out/src_fwd_v12.csv 2>/dev/null
ls -la gpurun_out/ | head -20; du -sh gpurun_out
