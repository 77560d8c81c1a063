#!/bin/bash
# GPU trip: tests, smoke, bench (with the evaluation leg), launch list, ncu full captures of the evaluation and the
# large-negative training kernels (each after its command exited 0 without ncu)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; cat gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
python tools/profile_eval.py amazon-book 1 > gpurun_out/eval_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:eval_kernel -c 1 -f -o gpurun_out/eval_prof \
    python tools/profile_eval.py amazon-book 1 > gpurun_out/eval_ncu.log 2>&1
python tools/bench_configs.py amazon-book-k1000 --samples 100000 --epochs 2 > gpurun_out/bigk_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:bigk -s 1 -c 1 -f -o gpurun_out/bigk_prof \
    python tools/bench_configs.py amazon-book-k1000 --samples 100000 --epochs 2 > gpurun_out/bigk_ncu.log 2>&1
tail -2 gpurun_out/bigk_ncu.log
