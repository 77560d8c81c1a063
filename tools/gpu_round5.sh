#!/bin/bash
# GPU trip: all GPU tests, smoke, full bench line, reference arm, launch list, ncu full capture of the CTA kernel,
# loss / Recall parity of the library defaults on the three BASELINE shapes
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -5 > gpurun_out/pytest_gpu.log; cat gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -3 gpurun_out/bench.err; cut -c1-300 gpurun_out/bench.json
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; cut -c1-300 gpurun_out/bench_ref.json
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_list.log 2>&1
python tools/profile_train.py amazon-book 2 3 > gpurun_out/cta_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:train_cta_kernel -s 1 -c 1 -f -o gpurun_out/cta_prof \
    python tools/profile_train.py amazon-book 2 3 > gpurun_out/cta_ncu.log 2>&1
tail -2 gpurun_out/cta_ncu.log
for s in amazon-book yelp18 gowalla; do python tools/recall_parity.py $s 5 auto 2>&1 | grep -v "^Test in" | tee gpurun_out/recall_$s.log; done
