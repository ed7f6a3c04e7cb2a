python -m pytest tests -q -x -m gpu 2>&1 | tail -3
for C in "cfg3_box 1920 1080 16" "cfg3_open 1920 1080 16" "cfg2_open 1280 720 8"; do python tools/prof_case.py $C auto 3 1 2>&1 | tail -1; done
for B in 2 8; do python tools/prof_case.py cfg4_box 3840 2160 64 auto 2 $B 2>&1 | tail -1; python tools/prof_case.py cfg4_open 3840 2160 64 auto 3 $B 2>&1 | tail -1; done
