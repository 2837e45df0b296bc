# First GPU call for the SURVEY 8(f) rows built without a GPU.  One B200, about 10 minutes:
#   gpurun --timeout 900 -- 'bash tools/first_gpu_run_next_rows.sh'
# Order: the established suite first (must stay green), then the new rows' tests, their micro-benchmarks, the C3 loop on
# both arms, launch lists.  Everything lands in gpurun_out/; copy what is to be judged into profiles/.
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -rxX -k "not zz_next" 2>&1 | tail -5 | tee gpurun_out/r02_pytest_gpu_established.log
timeout 600 python -m pytest tests/test_zz_next_rows_gpu.py -m gpu -q -rxX 2>&1 | tail -40 | tee gpurun_out/r02_pytest_gpu_next_rows.log
# (compute-sanitizer is closed on this GPU pool; the memory check of these kernels is tools/emu_sanitize.sh on the host)
python tests/golden/make_golden_knn_gpu.py > gpurun_out/r02_make_golden_knn.log 2>&1; echo "knn golden rc=$?"
timeout 420 python tools/bench_next.py all > gpurun_out/r02_bench_next.jsonl 2> gpurun_out/r02_bench_next.err; echo "bench_next rc=$?"; cat gpurun_out/r02_bench_next.jsonl | cut -c1-400
timeout 300 python tools/train_c3.py --impl b200 > gpurun_out/r02_c3_b200.json 2> gpurun_out/r02_c3_b200.err; echo "c3 b200 rc=$?"
timeout 400 python tools/train_c3.py --impl reference > gpurun_out/r02_c3_reference.json 2> gpurun_out/r02_c3_reference.err; echo "c3 reference rc=$?"
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
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"adam_kernel|knn_search_kernel|knn_gather_kernel|densify_apply_kernel|densify_classify_kernel" -c 8 -o /tmp/prof_next -f python tools/bench_next.py all --steps 1 --warmup 1 > gpurun_out/r02_ncu_full_next.log 2>&1; echo "ncu full rc=$?"
ncu -i /tmp/prof_next.ncu-rep --page raw --csv > gpurun_out/r02_prof_next_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/r02_prof_next_raw.csv > gpurun_out/r02_ncu_next_summary.txt 2>/dev/null
ls -la gpurun_out | head -30
