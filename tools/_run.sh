for T in "0,0,0,0" "0,0,0,2" "0,0,0,3" "0,0,0,4" "0,0,0,5"; do
for C in "cfg4_open 3840 2160 64" "cfg4_box 3840 2160 8"; do
RT_VOTE_TUNE=$T python tools/prof_case.py $C persistent 4 1 2>&1 | awk '{printf "%s %s %s %s %s | ", $1, $3,$4,$5,$6} END {print ""}'
done; done
