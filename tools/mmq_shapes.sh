#!/bin/bash
# per-shape timing of the tensor-core path under a few grid / tile settings
cd "$(dirname "$0")/.."
for shp in "4096 4096" "14336 4096" "4096 14336" "1024 4096"; do
  for cfg in "" "B200_MMQ_GRID=112" "B200_MMQ_BN=128" "B200_MMQ_BN=256"; do
    echo -n "[$cfg] "; env $cfg python tools/prof_mmq.py q4_K $shp 512 4 2>&1 | tail -1
  done
done
for n in 1024 2048; do echo -n "[n=$n] "; python tools/prof_mmq.py q4_K 14336 4096 $n 4 2>&1 | tail -1; echo -n "[n=$n] "; python tools/prof_mmq.py q4_K 4096 4096 $n 4 2>&1 | tail -1; done
