# C3 training loop, b200 arm, standalone: run-to-run spread with the default and with the expandable-segments allocator
set -u
mkdir -p gpurun_out
for TAG in default1 default2 expand1 expand2; do
  case $TAG in expand*) export PYTORCH_CUDA_ALLOC_CONF=expandable_segments:True;; *) unset PYTORCH_CUDA_ALLOC_CONF;; esac
  timeout 600 python tools/train_c3.py --impl b200 > gpurun_out/r02_c3_$TAG.json 2> gpurun_out/r02_c3_$TAG.err
  python - <<PY
import json
j = json.loads(open("gpurun_out/r02_c3_$TAG.json").read().strip().splitlines()[-1])
print("$TAG: ms/iter %.4f final_P %d wall %.1f s" % (j["value"], j["final_P"], j["wall_s"]), [round(b["iter_ms"], 2) for b in j["buckets"]], "densify", [round(b["densify_stats_ms"], 2) for b in j["buckets"]])
PY
done
