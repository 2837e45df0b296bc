# The host-emulated kernel tests (tests/test_kernels_emulated.py) under AddressSanitizer: out-of-bounds reads / writes of
# the kernels on host arrays (numpy allocations get red zones through the preloaded runtime).  CPU only, about 90 s.
#   bash tools/emu_asan.sh
set -eu
cd "$(dirname "$0")/.."
D=/tmp/asan_emu; mkdir -p $D
for f in emu_next emu_fused emu_preprocess; do
  g++ -std=c++20 -O1 -g -fsanitize=address -fno-omit-frame-pointer -pthread -fPIC -shared -ffp-contract=off -mfma -w \
      -I/usr/local/cuda/include tests/cuda_emu/$f.cpp -o $D/lib$f.so
done
sed -e "s#os.path.join(EMU_DIR, \"libemu_next.so\")#\"$D/libemu_next.so\"#" -e "s#os.path.join(EMU_DIR, \"libemu_fused.so\")#\"$D/libemu_fused.so\"#" \
    -e "s#os.path.join(EMU_DIR, \"libemu_preprocess.so\")#\"$D/libemu_preprocess.so\"#" tests/test_kernels_emulated.py > tests/_asan_emulated_tmp.py
trap 'rm -f tests/_asan_emulated_tmp.py' EXIT
LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 \
  python -m pytest tests/_asan_emulated_tmp.py -x -q -p no:cacheprovider
