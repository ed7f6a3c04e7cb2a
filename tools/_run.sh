for L in "" tools/_libs/librt_m7.so tools/_libs/librt_m8.so; do
for C in "cfg4_open 3840 2160 64" "cfg4_box 3840 2160 8"; do
echo -n "lib=$L "; RT_B200_LIB=$L python tools/prof_case.py $C persistent 5 1 2>&1 | tail -1
done; done
