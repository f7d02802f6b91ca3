#!/bin/bash
# AddressSanitizer pass over the HDF5 subset reader: builds csrc/rt with -fsanitize=address and runs the damaged-file
# fuzz worker (tests/workers/h5_fuzz.py) on it.  Report in profiles/asan_h5_fuzz.txt.  No GPU needed.
set -u
cd "$(dirname "$0")/.."
mkdir -p build/asan profiles
g++ -O1 -g -fsanitize=address -fno-omit-frame-pointer -std=c++17 -fPIC -pthread -I include -shared csrc/rt/*.cpp -o build/asan/libb200rt.so -lrt 2> build/asan/build.log || exit 1
ASAN=$(gcc -print-file-name=libasan.so)
B200RT_LIB=$PWD/build/asan/libb200rt.so ASAN_OPTIONS=detect_leaks=0:halt_on_error=1 LD_PRELOAD=$ASAN python tests/workers/h5_fuzz.py ${1:-1000} > build/asan/run.log 2>&1
RC=$?
# the same build under the 3-rank shared-memory collectives worker and the COCOeval / IoU tests
rm -f build/asan/rt.*
B200RT_LIB=$PWD/build/asan/libb200rt.so PYTHONPATH=$PWD ASAN_OPTIONS=detect_leaks=0:halt_on_error=0:log_path=$PWD/build/asan/rt LD_PRELOAD=$ASAN \
  python -m batchai_retinanet_horovod_coco_b200.parallel.launcher -np 3 python tests/workers/rt_collectives.py basic > build/asan/rt_run.log 2>&1
RC2=$?
B200RT_LIB=$PWD/build/asan/libb200rt.so ASAN_OPTIONS=detect_leaks=0:halt_on_error=0:log_path=$PWD/build/asan/rt LD_PRELOAD=$ASAN \
  python -m pytest tests/test_eval.py -q -p no:cacheprovider > build/asan/eval_run.log 2>&1
RC3=$?
{
  echo "# AddressSanitizer run of the HDF5 reader in libb200rt.so over ${1:-1000} damaged copies of tests/fixtures/keras_resnet50_stem_libhdf5.h5 (exit $RC)"
  echo "# $(g++ --version | head -1)"
  grep FUZZ_OK build/asan/run.log
  echo "AddressSanitizer reports: $(grep -c 'ERROR: AddressSanitizer' build/asan/run.log)"
  grep -A12 'ERROR: AddressSanitizer' build/asan/run.log | head -40
  echo "# collectives worker, 3 ranks (exit $RC2): $(grep -o WORKER_OK build/asan/rt_run.log | wc -l) workers finished; tests/test_eval.py (exit $RC3): $(tail -1 build/asan/eval_run.log)"
  echo "AddressSanitizer reports in those runs: $(cat build/asan/rt.* 2>/dev/null | grep -c 'ERROR: AddressSanitizer')"
} > profiles/asan_h5_fuzz.txt
cat profiles/asan_h5_fuzz.txt
