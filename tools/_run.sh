for B in 1 8; do
for T in tools tools/_old/tools; do
echo "TREE=$T B=$B"
python $T/prof_case.py cfg4_open 3840 2160 64 samples 3 $B 2>&1 | tail -1
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/spec_ab.csv python $T/prof_case.py cfg4_open 3840 2160 64 samples 2 $B > /dev/null 2>&1
python tools/spec_launches.py gpurun_out/spec_ab.csv
done; done
