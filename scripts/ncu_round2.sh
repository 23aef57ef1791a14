set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-extras"
$CMD > gpurun_out/r2f_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2f_launches.csv $CMD > gpurun_out/r2f_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:decoder_persistent -s 3 -c 1 -o gpurun_out/r2f_dec $CMD > gpurun_out/r2f_ncu2.log 2>&1
echo "full decoder rc=$?"
ncu --set full --clock-control none --import-source on -k regex:conv_tc_kernel -s 9 -c 3 -o gpurun_out/r2f_conv $CMD > gpurun_out/r2f_ncu3.log 2>&1
echo "full conv rc=$?"
ls -la gpurun_out/ | tail -8
