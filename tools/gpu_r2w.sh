#!/bin/bash
set -u
for S in "cfg4_open 3840 2160 8" "cfg4_open 3840 2160 16" "cfg2_open 1280 720 8" "cfg1_open 640 360 4" "cfg4_open 1920 1080 8"; do
  for V in pixel persistent chain samples; do
    python tools/prof_case.py $S $V 3 1 | tail -2 | tr '\n' ' '; echo
  done
done
