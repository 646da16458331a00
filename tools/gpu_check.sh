python -m pytest tests -m gpu -x -q > gpurun_out/$1_pytest.log 2>&1; tail -3 gpurun_out/$1_pytest.log
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/$1_bench.json 2> gpurun_out/$1_bench.err
python -c "
import json; d=json.load(open('gpurun_out/$1_bench.json')); print('ms_per_step', d['ms_per_step'], d['roofline']['kernel_ms'], 'e2e', d['e2e']['ms_per_step'])"
python tools/phase_times.py 35 2>&1 | tee gpurun_out/$1_phase35.txt
