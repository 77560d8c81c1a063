#!/bin/bash
# GPU trip (round 1, session 11): tests, smoke, bench (both arms), per-config rows incl. evaluation, launch list
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; cat gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err; cat gpurun_out/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cat gpurun_out/bench_ref.json
python tools/bench_configs.py gowalla yelp18 amazon-book --eval > gpurun_out/configs.json 2> gpurun_out/configs.err; cat gpurun_out/configs.json
timeout 600 python tools/bench_configs.py amazon-book-k1000 --samples 400000 > gpurun_out/configs_k1000.json 2> gpurun_out/configs_k1000.err; cat gpurun_out/configs_k1000.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
