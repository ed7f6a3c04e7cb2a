#!/bin/bash
set -u
for S in "cfg3_open 1920 1080 16" "cfg2_open 1280 720 8" "cfg3_box 1920 1080 16"; do
  for V in auto persistent chain samples pixel; do
    python tools/prof_case.py $S $V 3 1 | tail -2 | tr '\n' ' '; echo
  done
done
