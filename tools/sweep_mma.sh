#!/bin/bash
# DMMA profile kernel against the FMA kernels per model, for MS = 1, 2, 4 sample blocks per warp step (development aid)
for m in "$@"; do
  FAIRIFY_B200_MMA_PROFILE=0 python tools/time_profile.py $m 200000 2>&1 | tail -1 | cut -c1-120 | sed "s/^/fma       /"
  for ms in 1 2 4; do
    FAIRIFY_B200_MMA_MS=$ms python tools/time_profile.py $m 200000 2>&1 | tail -1 | cut -c1-120 | sed "s/^/dmma ms=$ms /"
  done
done
