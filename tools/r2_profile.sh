# round-2 profile artefacts of the config-2 bench (run under gpurun): launch list of the bench command, then one
# ncu --set full capture of the fused kernel
set -x
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-spinup --no-e2e --no-secondary"
$CMD > gpurun_out/r2_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches.csv $CMD > gpurun_out/r2_ncu1.log 2>&1
python tools/prof_run.py 4 > gpurun_out/plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:"fused_kernel" -s 2 -c 1 -o gpurun_out/r2_fused python tools/prof_run.py 4 > gpurun_out/r2_ncu2.log 2>&1
ls -la gpurun_out | grep r2_
