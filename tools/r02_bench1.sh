#!/bin/bash
# round 2: default bench line (scale-25 headline + blocks) on one GPU, smoke, GPU suite
mkdir -p gpurun_out
( time python bench.py --steps 5 --warmup 3 > gpurun_out/c_bench_default.json 2> gpurun_out/c_bench_default.err ) 2> gpurun_out/c_bench_default.time
echo "bench rc=$?"; tail -5 gpurun_out/c_bench_default.err; cat gpurun_out/c_bench_default.time
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/c_smoke.log 2>&1; tail -2 gpurun_out/c_smoke.log
timeout 1200 python -m pytest tests -x -q -m gpu > gpurun_out/c_pytest.log 2>&1
tail -8 gpurun_out/c_pytest.log
