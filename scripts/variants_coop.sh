#!/bin/bash
# Occupancy / group-size sweep for the block-cooperative kernel on radiolysis (ny = 14): "flags|kernel" per line.
# usage: scripts/variants_coop.sh build | run      (METHOD=rosenbrock4|bdf)
mode=${1:-run}
method=${METHOD:-rosenbrock4}
while IFS='|' read -r flags kern; do
  export PYODESYS_CUDA_NVCC_FLAGS="$flags"
  if [ "$mode" = build ]; then
    python -c "from pyodesys_b200 import systems; s=systems.radiolysis(compile=True); print(repr('$flags'), s._native.tag)"
    python scripts/regs.py radiolysis coop 2>/dev/null | grep ", 0>"
  else
    echo "## flags='$flags' kernel=$kern method=$method"
    REPS=3 python scripts/run_c4.py radiolysis 100000 $kern $method | tail -1
  fi
done < "${VARIANTS_FILE:-scripts/variants_c4c.txt}"
