#!/bin/bash
# the C++ drop-in on all GPUs of the box: one model.inference(batch) of 2^24 leaves == the one-GPU call
O=gpurun_out
mkdir -p $O
timeout 600 python -m pytest tests/test_dropin_cpp.py -m gpu -x -q > $O/r2x_pytest_dropin_8gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2x_pytest_dropin_8gpu.log; tail -3 $O/r2x_pytest_dropin_8gpu.log
timeout 300 ./tests/cpp/_bin/test_dropin --multi 16777216 > $O/r2x_dropin_multi.log 2>&1; echo "rc=$?" >> $O/r2x_dropin_multi.log; tail -3 $O/r2x_dropin_multi.log
timeout 300 ./tests/cpp/_bin/test_dropin --multi 8 >> $O/r2x_dropin_multi.log 2>&1; echo "rc=$?" >> $O/r2x_dropin_multi.log; tail -3 $O/r2x_dropin_multi.log
