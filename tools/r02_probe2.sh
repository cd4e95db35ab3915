#!/bin/bash
# round 2, GPU call 2: store-bandwidth probe (floor of the corpus store) + the GPU suite on the edited engine
mkdir -p gpurun_out
tools/_build/store_probe 8 > gpurun_out/b_store_probe.txt 2>&1
cat gpurun_out/b_store_probe.txt
timeout 900 python -m pytest tests -x -q -m gpu > gpurun_out/b_pytest.log 2>&1
tail -15 gpurun_out/b_pytest.log
