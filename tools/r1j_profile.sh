set -x
python -m pytest tests -m gpu -q > gpurun_out/r1j_pytest.log 2>&1; tail -2 gpurun_out/r1j_pytest.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1j_bench_ref.json 2> gpurun_out/r1j_bench_ref.err
python bench.py > gpurun_out/r1j_bench.json 2> gpurun_out/r1j_bench.err; cat gpurun_out/r1j_bench.json
python bench.py --train --steps 100 > gpurun_out/r1j_bench_train.json 2>> gpurun_out/r1j_bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1j_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-spinup --no-e2e > gpurun_out/r1j_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'finalize_kernel|products_kernel' -s 4 -c 2 -o gpurun_out/r1j_full python tools/prof_run.py 4 > gpurun_out/r1j_ncu3.log 2>&1
ls -la gpurun_out/ | grep r1j
