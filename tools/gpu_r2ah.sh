#!/bin/bash
set -u
for T in 16,1,32,0 8,1,32,0 12,1,32,0 24,1,32,0 32,1,32,0 16,1,16,0 16,1,8,0; do
  RT_CHAIN_TUNE=$T python tools/prof_case.py cfg4_open 3840 2160 64 chain 4 1 | tail -2 | head -1 | tr '\n' ' '; echo " tune=$T"
done
