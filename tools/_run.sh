PYTEST_K="samples" VARIANTS=samples bash tools/gpu_quick.sh
for B in 1 8; do
python tools/prof_case.py cfg4_open 3840 2160 64 samples 2 $B > /dev/null 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/spec_launches_$B.csv python tools/prof_case.py cfg4_open 3840 2160 64 samples 2 $B > /dev/null 2>&1
grep -v "^==" gpurun_out/spec_launches_$B.csv | awk -F'","' 'NR>1{print $5, $NF}' | tail -3
done
