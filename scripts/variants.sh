#!/bin/bash
# Kernel-variant sweep for the headline workload: each line of flags is one build of the Robertson library.
# usage: scripts/variants.sh build   (here, no GPU: compiles every variant so the .so files travel)
#        scripts/variants.sh run     (on the GPU box: times each variant, one JSON line each)
mode=${1:-run}
while IFS= read -r flags; do
  export PYODESYS_CUDA_NVCC_FLAGS="$flags"
  if [ "$mode" = build ]; then
    python -c "from pyodesys_b200 import systems; s=systems.robertson(compile=True); from oracle import build_oracle; build_oracle(s._native); print(repr('$flags'), s._native.tag)"
  else
    C3_METHODS=${C3_METHODS:-rosenbrock4} python scripts/bench_configs.py c3 --reps 3
  fi
done < "${VARIANTS_FILE:-scripts/variants.txt}"
