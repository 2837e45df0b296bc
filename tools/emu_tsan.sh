# The host-emulated kernel tests (tests/test_kernels_emulated.py) under ThreadSanitizer: data races between the emulated threads of a block (a missing __syncthreads) in
# the kernels.  CPU only, about 2 minutes; prints the number of reports (0 = clean).
#   bash tools/emu_tsan.sh
set -eu
cd "$(dirname "$0")/.."
D=/tmp/tsan_emu; mkdir -p $D
for f in emu_next emu_fused emu_preprocess; do
  g++ -std=c++20 -O1 -g -fsanitize=thread -pthread -fPIC -shared -ffp-contract=off -mfma -w \
      -I/usr/local/cuda/include tests/cuda_emu/$f.cpp -o $D/lib$f.so
done
sed -e "s#os.path.join(EMU_DIR, \"libemu_next.so\")#\"$D/libemu_next.so\"#" -e "s#os.path.join(EMU_DIR, \"libemu_fused.so\")#\"$D/libemu_fused.so\"#" \
    -e "s#os.path.join(EMU_DIR, \"libemu_preprocess.so\")#\"$D/libemu_preprocess.so\"#" tests/test_kernels_emulated.py > tests/_tsan_emulated_tmp.py
trap 'rm -f tests/_tsan_emulated_tmp.py' EXIT
LD_PRELOAD=$(gcc -print-file-name=libtsan.so) TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 exitcode=0" \
  python -m pytest tests/_tsan_emulated_tmp.py -q -p no:cacheprovider > /tmp/tsan_emu/run.log 2>&1 || true
tail -2 /tmp/tsan_emu/run.log
echo "ThreadSanitizer reports: $(grep -c 'WARNING: ThreadSanitizer' /tmp/tsan_emu/run.log || true)"
