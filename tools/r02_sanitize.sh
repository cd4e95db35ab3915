#!/bin/bash
# round 2: compute-sanitizer memcheck over the tests that drive the kernels written or rewritten this round
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 --print-limit 20 python -m pytest tests/test_gpu_parity.py -x -q -m gpu \
  -k "triangle or coalesced_generation or schedule_variants or index_in_chunks or constructor_refuses or bulk_sampler or keyed_mh_mode" > gpurun_out/x_memcheck.log 2>&1
echo "memcheck rc=$?"
grep -E "ERROR SUMMARY|passed|failed|Invalid|out of bounds" gpurun_out/x_memcheck.log | tail -8
