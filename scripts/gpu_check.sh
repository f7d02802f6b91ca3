#!/bin/bash
# One gpurun call: kernel tests, engine tests, short A/B benches.  Every step runs under its own timeout so a hung
# kernel cannot hold the box; logs go to gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
O=gpurun_out
python build_native.py > $O/build.log 2>&1 || { tail -20 $O/build.log; exit 1; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
for step in "$@"; do
  case $step in
    conv)    timeout 600 python -m pytest tests/test_gpu_conv.py -x -q > $O/t_conv.log 2>&1; echo "conv rc=$?"; tail -15 $O/t_conv.log;;
    conv1)   B200_CONV_IMPL=1 timeout 600 python -m pytest tests/test_gpu_conv.py -q > $O/t_conv_v1.log 2>&1; echo "conv(v1) rc=$?"; tail -8 $O/t_conv_v1.log;;
    ops)     timeout 600 python -m pytest tests/test_gpu_ops.py -x -q > $O/t_ops.log 2>&1; echo "ops rc=$?"; tail -8 $O/t_ops.log;;
    engine)  timeout 900 python -m pytest tests/test_gpu_engine.py tests/test_gpu_e2e.py tests/test_gpu_vgg.py -x -q > $O/t_engine.log 2>&1; echo "engine rc=$?"; tail -15 $O/t_engine.log;;
    all)     timeout 1500 python -m pytest tests -m gpu -x -q > $O/t_all.log 2>&1; echo "all rc=$?"; tail -15 $O/t_all.log;;
    bench8)  timeout 300 python bench.py --batch 8 --steps 30 --warmup 5 > $O/bench_b8.json 2> $O/bench_b8.err; echo "bench8 rc=$?"; cat $O/bench_b8.json; tail -3 $O/bench_b8.err;;
    bench8v1) B200_CONV_IMPL=1 timeout 300 python bench.py --batch 8 --steps 30 --warmup 5 --no-e2e > $O/bench_b8_v1.json 2> $O/bench_b8_v1.err; echo "bench8v1 rc=$?"; cat $O/bench_b8_v1.json;;
    bench16) timeout 400 python bench.py --batch 16 --steps 50 --warmup 5 > $O/bench_b16.json 2> $O/bench_b16.err; echo "bench16 rc=$?"; cat $O/bench_b16.json; tail -3 $O/bench_b16.err;;
    bench1)  timeout 300 python bench.py --batch 1 --steps 100 --warmup 10 > $O/bench_b1.json 2> $O/bench_b1.err; echo "bench1 rc=$?"; cat $O/bench_b1.json;;
    launches) B200_GRAPH=0 timeout 300 python bench.py --batch 8 --steps 1 --warmup 3 --no-e2e > $O/plain.log 2>&1 && B200_GRAPH=0 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/launches.csv python bench.py --batch 8 --steps 1 --warmup 3 --no-e2e > $O/ncu_launch.log 2>&1; echo "launches rc=$?"; python scripts/launch_summary.py $O/launches.csv x > $O/launch_summary.txt 2>&1; head -20 $O/launch_summary.txt;;
    *) echo "unknown step $step";;
  esac
done
