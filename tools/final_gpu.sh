#!/bin/bash
# Round-end validation on one B200: GPU test suite, smoke, the default bench line, then the ncu evidence for
# profiles/ (launch list + one --set full capture per configuration, summarised on the box).
python -m pytest tests -x -q -m gpu > gpurun_out/final_gpu_tests.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/final_gpu_tests.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/final_smoke.log
python bench.py > gpurun_out/final_bench_1gpu.json 2> gpurun_out/final_bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/final_bench_reference.json 2>/dev/null; echo "reference arm rc=$?"
python bench.py --config cfg3 --no-cpu-baseline --steps 50 > gpurun_out/final_bench_cfg3.json 2>/dev/null
for c in "0.02 1" "0.05 1" "0.1 1" "0.05 10"; do set -- $c
  python bench.py --config cfg5 --segment-len $1 --log-every $2 --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null > gpurun_out/final_bench_cfg5_$1_$2.json; done
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02b_launches.csv \
    python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02b_launches.log 2>&1
tools/ncu_capture.sh r02b_cfg2 4096000
tools/ncu_capture.sh r02b_cfg3 16384000 --config cfg3
tools/ncu_capture.sh r02b_cfg5_2200_l1 1024000 --config cfg5 --segment-len 0.05 --log-every 1
tools/ncu_capture.sh r02b_cfg5_5500_l1 1024000 --config cfg5 --segment-len 0.02 --log-every 1
rm -f gpurun_out/r02b_cfg3_src.csv.gz gpurun_out/r02b_cfg5_*_src.csv.gz
