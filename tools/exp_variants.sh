#!/bin/bash
# tools/exp_variants.sh "<flags 1>" "<flags 2>" ... -- on the GPU box: rebuild the batch kernel with each set of
# -D flags and print the c5 bench value (3 steps) + the checksum of the results (must not change).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for flags in "$@"; do
  touch qpdo_b200/csrc/batch.cu
  EXTRA_NVFLAGS="$flags" ./build.sh > /dev/null 2>&1 || { echo "[$flags] build failed"; continue; }
  python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/variant.json 2> gpurun_out/variant.err || { echo "[$flags] bench failed"; tail -3 gpurun_out/variant.err; continue; }
  python - "$flags" <<'EOF'
import json, sys
d = json.load(open("gpurun_out/variant.json"))
r = d["result"]
print("[%s] %.0f QP/s device, %.0f e2e, kernel %.2f ms, solved %d pinf %d objsum %.10f match %s" % (
    sys.argv[1], d["value"], d["e2e"]["value"], d["kernel_ms_per_step"], r["solved"], r["primal_infeasible"], r["objective_sum"], r["host_leg_matches_device_leg"]))
EOF
done
