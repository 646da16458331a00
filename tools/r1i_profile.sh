set -x
python -m pytest tests -m gpu -q > gpurun_out/r1i_pytest.log 2>&1; tail -2 gpurun_out/r1i_pytest.log
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1i_bench_ref.json 2> gpurun_out/r1i_bench_ref.err
python bench.py > gpurun_out/r1i_bench.json 2> gpurun_out/r1i_bench.err; cat gpurun_out/r1i_bench.json
python bench.py --train --steps 100 > gpurun_out/r1i_bench_train.json 2>> gpurun_out/r1i_bench.err
python tools/grad_time.py > gpurun_out/r1i_grad_time.txt 2>&1
python tools/viterbi_time.py > gpurun_out/r1i_viterbi_time.txt 2>&1
python tools/compact_time.py > gpurun_out/r1i_compact_time.txt 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1i_launches.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-spinup --no-e2e > gpurun_out/r1i_ncu1.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1i_gradvit_launches.csv python tools/prof_grad_vit.py > gpurun_out/r1i_ncu2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:'grad_kernel|viterbi_kernel|finalize_kernel|products_kernel' -s 6 -c 5 -o gpurun_out/r1i_full python tools/prof_grad_vit.py > gpurun_out/r1i_ncu3.log 2>&1
ls -la gpurun_out/ | tail -15
