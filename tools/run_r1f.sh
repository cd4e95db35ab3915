set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; tail -c 600 gpurun_out/f_bench_n1.err
python tools/det_modes.py dw25 2 coalesced > gpurun_out/f_det_dw25.txt 2>&1; cat gpurun_out/f_det_dw25.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/f_launches_dw25.csv python tools/det_modes.py dw25 1 coalesced > gpurun_out/f_ncu_dw25.log 2>&1
tail -2 gpurun_out/f_ncu_dw25.log
