#!/bin/bash
cd "$(dirname "$0")/.."
for shp in "4096 4096" "14336 4096" "4096 14336" "1024 4096"; do
  for cfg in "" "B200_MMQ_BN=128" "B200_MMQ_BN=256"; do
    echo -n "[$cfg] "; env $cfg timeout 120 python tools/prof_mmq.py q4_K $shp 512 4 2>&1 | tail -1
  done
done
for n in 1024 2048; do for cfg in "" "B200_MMQ_BN=128"; do echo -n "[n=$n $cfg] "; env $cfg timeout 120 python tools/prof_mmq.py q4_K 14336 4096 $n 4 2>&1 | tail -1; done; done
