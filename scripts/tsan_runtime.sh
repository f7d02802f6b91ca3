#!/bin/bash
# ThreadSanitizer pass over the C++ host runtime: builds csrc/rt with -fsanitize=thread and runs the
# multi-rank CPU collectives worker (cycle thread vs API thread, fusion, timeline, stall check) on it.
# Usage: scripts/tsan_runtime.sh [ranks]   -> report in profiles/tsan_runtime.txt   (no GPU needed)
set -u
cd "$(dirname "$0")/.."
NP=${1:-3}
mkdir -p build/tsan profiles
g++ -O1 -g -fsanitize=thread -std=c++17 -fPIC -pthread -I include -shared csrc/rt/*.cpp -o build/tsan/libb200rt.so -lrt || exit 1
TSAN=$(gcc -print-file-name=libtsan.so)
TL=$(mktemp -d)/timeline.json
export B200RT_LIB=$PWD/build/tsan/libb200rt.so PYTHONPATH=$PWD
# the interpreter itself is not instrumented: only report races whose stack touches the runtime library
export TSAN_OPTIONS="halt_on_error=0 report_signal_unsafe=0 exitcode=0 log_path=$PWD/build/tsan/report"
rm -f build/tsan/report.*
LD_PRELOAD=$TSAN python -m batchai_retinanet_horovod_coco_b200.parallel.launcher -np $NP python tests/workers/rt_collectives.py basic $TL > build/tsan/run.log 2>&1
RC=$?
# allreduce and broadcast released in one response with a delayed non-root rank (the result-buffer race of round 1)
B200RT_FAULT=delay:1:3 LD_PRELOAD=$TSAN python -m batchai_retinanet_horovod_coco_b200.parallel.launcher -np $NP python tests/workers/rt_collectives.py ar_then_bcast >> build/tsan/run.log 2>&1
RC2=$?
{
  echo "# ThreadSanitizer run of libb200rt.so: $NP ranks, tests/workers/rt_collectives.py basic (exit $RC) + ar_then_bcast with B200RT_FAULT=delay:1:3 (exit $RC2)"
  echo "# $(g++ --version | head -1)"
  grep -o WORKER_OK build/tsan/run.log | wc -l | sed 's/^/workers finished: /'
  N=$(cat build/tsan/report.* 2>/dev/null | grep -c "WARNING: ThreadSanitizer")
  echo "ThreadSanitizer warnings: $N"
  cat build/tsan/report.* 2>/dev/null | grep -A12 "WARNING: ThreadSanitizer" | grep -E "WARNING|libb200rt|#0|#1" | head -60
} > profiles/tsan_runtime.txt
cat profiles/tsan_runtime.txt | head -40
