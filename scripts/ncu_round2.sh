set -x
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-parity --no-extras"
$CMD > gpurun_out/r2i_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2i_launches.csv $CMD > gpurun_out/r2i_ncu1.log 2>&1
echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:decoder_persistent -s 3 -c 1 -o gpurun_out/r2i_dec $CMD > gpurun_out/r2i_ncu2.log 2>&1
echo "full decoder rc=$?"
ls -la gpurun_out/ | tail -8
