mkdir -p gpurun_out; rm -f gpurun_out/rc.txt
for L in gen.cnx0.pw1 stft16; do
STTS_PROFILE_LAYER=$L timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -o /tmp/r1d_$L python tools/profile_step.py f16 > gpurun_out/ncu_$L.log 2>&1; echo "ncu_$L rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1d_$L.ncu-rep --page raw --csv > gpurun_out/r1d_$L.raw.csv 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page details > gpurun_out/r1d_$L.details.txt 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page source --csv > gpurun_out/r1d_$L.source.csv 2>/dev/null
done
L=flow3.wn1
STTS_PROFILE_LAYER=$L timeout 600 ncu --set full --import-source on --clock-control none --profile-from-start off -c 1 -o /tmp/r1d_$L python tools/profile_step.py f16 > gpurun_out/ncu_$L.log 2>&1; echo "ncu_$L rc=$?" >> gpurun_out/rc.txt
ncu -i /tmp/r1d_$L.ncu-rep --page raw --csv > gpurun_out/r1d_$L.raw.csv 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page details > gpurun_out/r1d_$L.details.txt 2>/dev/null
ncu -i /tmp/r1d_$L.ncu-rep --page source --csv > gpurun_out/r1d_$L.source.csv 2>/dev/null
cp /tmp/r1d_$L.ncu-rep gpurun_out/r1_full_$L.ncu-rep
cat gpurun_out/rc.txt
