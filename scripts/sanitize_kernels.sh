#!/bin/bash
# compute-sanitizer pass over the kernel-level GPU tests.  ONE tool per invocation:
#   scripts/sanitize_kernels.sh memcheck|racecheck|synccheck|initcheck
# The tests run once without the tool first; the report lands in gpurun_out/<tool>_kernels.txt.
# On the GPU box: gpurun --timeout 1500 -- 'scripts/sanitize_kernels.sh memcheck'
set -u
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
T="tests/test_gpu_conv.py tests/test_gpu_ops.py tests/test_gpu_vgg.py"
python -m pytest $T -m gpu -x -q -p no:cacheprovider -k "not autograd" > gpurun_out/plain_kernels.log 2>&1 || { tail -5 gpurun_out/plain_kernels.log; exit 1; }
timeout 1200 compute-sanitizer --tool $TOOL --error-exitcode 77 --print-limit 20 --log-file gpurun_out/${TOOL}_kernels.txt \
  python -m pytest $T -m gpu -x -q -p no:cacheprovider -k "not autograd" > gpurun_out/${TOOL}_pytest.log 2>&1
echo "sanitizer_rc=$?"
tail -2 gpurun_out/${TOOL}_pytest.log
tail -4 gpurun_out/${TOOL}_kernels.txt
