# Round-2 end measurement on ONE B200: whole gpu suite, default bench on both arms, ncu launch list and one full capture of a
# step (exported to CSV on the box: the .ncu-rep with sources is larger than what gpurun copies back).
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -4 | tee gpurun_out/r02_final_pytest_gpu.log
timeout 600 python bench.py > gpurun_out/r02_final_bench_b200.json 2> gpurun_out/r02_final_bench_b200.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r02_final_bench_reference.json 2> gpurun_out/r02_final_bench_reference.err; echo "reference rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_final_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_final_ncu_launches.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"preprocess_|rs_|scan_|gather_|duplicate_|tc_|tile_ranges|render_" -s 17 -c 17 -o /tmp/prof_final -f python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras > gpurun_out/r02_final_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/prof_final.ncu-rep --page raw --csv > gpurun_out/r02_final_prof_raw.csv 2>/dev/null
ncu -i /tmp/prof_final.ncu-rep --page source --csv --kernel-name regex:render_bwd_kernel > gpurun_out/r02_final_src_bwd.csv 2>/dev/null
ncu -i /tmp/prof_final.ncu-rep --page source --csv --kernel-name regex:render_fwd_kernel > gpurun_out/r02_final_src_fwd.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_final_prof_raw.csv pipe_fma_cycles pipe_alu_cycles smsp__warps_eligible > gpurun_out/r02_final_ncu_summary.txt
python tools/mask_stats.py > gpurun_out/r02_final_mask_stats.json 2>/dev/null
ls -la gpurun_out | grep r02_final
