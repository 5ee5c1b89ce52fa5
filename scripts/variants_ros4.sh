#!/bin/bash
# Build-flag sweep for the thread-per-system Rosenbrock-4 kernel on Robertson (C3): one line of nvcc flags per variant.
# usage: scripts/variants_ros4.sh build | run
mode=${1:-run}
while read -r flags; do
  export PYODESYS_CUDA_NVCC_FLAGS="$flags"
  if [ "$mode" = build ]; then
    python -c "from pyodesys_b200 import systems; s=systems.robertson(compile=True); print(repr('$flags'), s._native.tag)"
    python scripts/regs.py robertson 2>/dev/null | grep "integrate_kernel<2, 0, 128, false"
  else
    echo "## flags='$flags'"
    REPS=3 python scripts/run_c4.py robertson 1000000 auto rosenbrock4 | tail -2
  fi
done < "${VARIANTS_FILE:-scripts/variants_ros4.txt}"
