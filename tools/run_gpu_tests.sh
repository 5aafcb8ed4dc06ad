#!/bin/bash
# Run on the GPU box via gpurun: the gpu test tier, logs under gpurun_out/.
export PYTHONPATH=$PWD:$PYTHONPATH
mkdir -p gpurun_out
timeout ${TEST_TIMEOUT:-900} python -m pytest tests -m gpu -x -q "$@" 2>&1 | tail -60 > gpurun_out/tests_gpu.log
rc=${PIPESTATUS[0]}
cat gpurun_out/tests_gpu.log
exit $rc
