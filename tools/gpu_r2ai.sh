#!/bin/bash
set -u
RT_NODE_ORDER=dfspairs timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "golden or image_ids" 2>&1 | tail -2
for R in 1 2; do
for O in bfs dfs dfspairs; do
  RT_NODE_ORDER=$O python tools/prof_case.py cfg4_open 3840 2160 64 auto 4 1 | tail -2 | head -1 | tr '\n' ' '; echo " order=$O"
  RT_NODE_ORDER=$O python tools/prof_case.py cfg4_box 3840 2160 8 auto 3 1 | tail -2 | head -1 | tr '\n' ' '; echo " order=$O"
done
done
for O in bfs dfs dfspairs; do
  RT_NODE_ORDER=$O python tools/prof_case.py cfg3_box 1920 1080 16 auto 3 1 | tail -2 | head -1 | tr '\n' ' '; echo " order=$O"
  RT_NODE_ORDER=$O python tools/prof_case.py cfg3_open 1920 1080 16 auto 3 1 | tail -2 | head -1 | tr '\n' ' '; echo " order=$O"
  RT_NODE_ORDER=$O python tools/prof_case.py cfg2_box 1280 720 8 auto 3 1 | tail -2 | head -1 | tr '\n' ' '; echo " order=$O"
done
