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
{
  echo "# AddressSanitizer run of the HDF5 reader in libb200rt.so over ${1:-1000} damaged copies of tests/fixtures/keras_resnet50_stem_libhdf5.h5 (exit $RC)"
  echo "# $(g++ --version | head -1)"
  grep FUZZ_OK build/asan/run.log
  echo "AddressSanitizer reports: $(grep -c 'ERROR: AddressSanitizer' build/asan/run.log)"
  grep -A12 'ERROR: AddressSanitizer' build/asan/run.log | head -40
} > profiles/asan_h5_fuzz.txt
cat profiles/asan_h5_fuzz.txt
