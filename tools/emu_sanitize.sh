# The host-emulated kernel and ABI tests (tests/test_kernels_emulated.py, tests/test_abi_emulated.py, tests/test_densify_cpu.py) under a sanitizer:
#   bash tools/emu_sanitize.sh address   out-of-bounds reads / writes of the kernels on host arrays (numpy's buffers get
#                                        red zones through the preloaded runtime) -- the host stand-in for memcheck
#   bash tools/emu_sanitize.sh thread    data races between the emulated threads of a block, i.e. a missing
#                                        __syncthreads() / __syncwarp() -- the host stand-in for racecheck
# CPU only, several minutes.  Prints the pytest summary and the number of sanitizer reports (0 = clean).
set -eu
cd "$(dirname "$0")/.."
KIND=${1:-address}
D=/tmp/emu_$KIND; mkdir -p $D
python - "$KIND" "$D" <<'PY'
import sys
sys.path.insert(0, "tests/cuda_emu")
import build_emu
kind, d = sys.argv[1], sys.argv[2]
flags = ["-g", "-fsanitize=" + kind] + (["-fno-omit-frame-pointer"] if kind == "address" else [])
for n in ("emu_next", "emu_fused", "emu_preprocess", "emu_abi"):
    print(build_emu.build(n, extra_flags=flags, out="%s/lib%s.so" % (d, n)))
PY
rm -f $D/san.*
if [ "$KIND" = address ]; then
  export LD_PRELOAD=$(gcc -print-file-name=libasan.so) ASAN_OPTIONS=detect_leaks=0:halt_on_error=0:log_path=$D/san
else
  export LD_PRELOAD=$(gcc -print-file-name=libtsan.so) TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 exitcode=0 log_path=$D/san"
fi
# reports go to $D/san.<pid> (pytest captures stderr); torch's own OpenMP workers are not what is under test
OMP_NUM_THREADS=1 GSR_EMU_SO_DIR=$D GSR_EMU_ABI_SO=$D/libemu_abi.so python -m pytest tests/test_kernels_emulated.py tests/test_abi_emulated.py tests/test_densify_cpu.py -q -p no:cacheprovider > $D/run.log 2>&1 || true
unset LD_PRELOAD
tail -3 $D/run.log
python3 - "$D" <<'PY'
import glob, sys
total = ours = 0
for f in glob.glob(sys.argv[1] + "/san.*"):
    for block in open(f, errors="ignore").read().split("=================="):
        if "Sanitizer" in block:
            total += 1
            ours += "libemu" in block          # a frame inside the emulated kernels / ABI
print("sanitizer reports: %d in the emulated code (%d in total, the rest inside torch / OpenMP)" % (ours, total))
PY
