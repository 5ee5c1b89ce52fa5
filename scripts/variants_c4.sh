#!/bin/bash
# Kernel-family / group-size sweep for C4 (radiolysis ny = 14): one line "flags|kernel" per variant.
# usage: scripts/variants_c4.sh build | run
mode=${1:-run}
while IFS='|' read -r flags kern; do
  export PYODESYS_CUDA_NVCC_FLAGS="$flags"
  if [ "$mode" = build ]; then
    python -c "from pyodesys_b200 import systems; s=systems.radiolysis(compile=True); from oracle import build_oracle; build_oracle(s._native); print(repr('$flags'), s._native.tag)"
  else
    echo "## flags='$flags' kernel=$kern"
    C4_KERNEL=$kern python scripts/bench_configs.py c4 --reps 2 | cut -c1-330
  fi
done < "${VARIANTS_FILE:-scripts/variants_c4.txt}"
