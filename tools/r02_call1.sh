# Round 2, GPU call 1: the whole gpu suite with nothing soft-marked, the no-alloc test alone in a fresh process, the
# reference-build k-NN fixture, the round's first bench lines, the micro-benchmarks of the 8(f) rows, C3 on both arms.
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader
timeout 900 python -m pytest tests -m gpu -x -q -rxXs 2>&1 | tail -25 | tee gpurun_out/r02_pytest_gpu_call1.log
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q 2>&1 | tail -5 | tee gpurun_out/r02_pytest_no_alloc_alone.log
python tests/golden/make_golden_knn_gpu.py > gpurun_out/r02_make_golden_knn.log 2>&1; echo "knn golden rc=$?"; tail -2 gpurun_out/r02_make_golden_knn.log
timeout 300 python tools/mask_stats.py > gpurun_out/r02_mask_stats.json 2> gpurun_out/r02_mask_stats.err; echo "mask stats rc=$?"; cat gpurun_out/r02_mask_stats.json
timeout 400 python bench.py --steps 20 --warmup 5 > gpurun_out/r02_bench_c2_call1.json 2> gpurun_out/r02_bench_c2_call1.err; echo "bench rc=$?"; cut -c1-600 gpurun_out/r02_bench_c2_call1.json
timeout 420 python tools/bench_next.py all > gpurun_out/r02_bench_next.jsonl 2> gpurun_out/r02_bench_next.err; echo "bench_next rc=$?"; cut -c1-500 gpurun_out/r02_bench_next.jsonl
timeout 400 python tools/train_c3.py --impl b200 > gpurun_out/r02_c3_b200.json 2> gpurun_out/r02_c3_b200.err; echo "c3 b200 rc=$?"; tail -3 gpurun_out/r02_c3_b200.err
timeout 500 python tools/train_c3.py --impl reference > gpurun_out/r02_c3_reference.json 2> gpurun_out/r02_c3_reference.err; echo "c3 reference rc=$?"; tail -3 gpurun_out/r02_c3_reference.err
python - <<'PY'
import json
for impl in ("b200", "reference"):
    try:
        j = json.loads(open("gpurun_out/r02_c3_%s.json" % impl).read().strip().splitlines()[-1])
        print(impl, "mean ms/iter %.3f final P %d" % (j["value"], j["final_P"]), [(b["P_from"], round(b["iter_ms"], 3)) for b in j["buckets"]])
    except Exception as e:
        print(impl, "no result:", e)
PY
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r02_launches_next.csv python tools/bench_next.py all --steps 2 --warmup 1 > gpurun_out/r02_ncu_next.log 2>&1; echo "ncu list rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"adam_kernel|knn_search_kernel|knn_gather_kernel|densify_apply_kernel|densify_classify_kernel|densify_stats" -c 8 -o /tmp/prof_next -f python tools/bench_next.py all --steps 1 --warmup 1 > gpurun_out/r02_ncu_full_next.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/prof_next.ncu-rep --page raw --csv > gpurun_out/r02_prof_next_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_prof_next_raw.csv > gpurun_out/r02_ncu_next_summary.txt 2>/dev/null
ls -la gpurun_out | head -40
