# the two default bench lines (b200 arm, reference arm) of the round-end measurement
set -u
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02_final_bench_b200.json 2> gpurun_out/r02_final_bench_b200.err; echo "bench rc=$?"
timeout 900 python bench.py --impl reference > gpurun_out/r02_final_bench_reference.json 2> gpurun_out/r02_final_bench_reference.err; echo "reference rc=$?"
