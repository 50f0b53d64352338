#!/bin/bash
# sweep of the DMMA profile kernel's tile height / warps per CTA per model (development aid; run on the GPU box)
for m in "$@"; do
  for ms in 1 2; do for w in 4 8; do
    FAIRIFY_B200_MMA_MS=$ms FAIRIFY_B200_MMA_WARPS=$w python tools/time_profile.py $m 20000 2>&1 | tail -1 | cut -c1-150 | sed "s/^/ms=$ms warps=$w /"
  done; done
done
