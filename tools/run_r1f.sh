set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py > gpurun_out/k_bench_n2v20.json 2> gpurun_out/k_bench_n2v20.err; tail -c 300 gpurun_out/k_bench_n2v20.err
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/k_bench_reference.json 2> gpurun_out/k_bench_reference.err; cat gpurun_out/k_bench_reference.json | head -c 600
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/k_launches_n2v20.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --updates 0 > gpurun_out/k_ncu1.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:walk_n2v_kernel --launch-skip 3 -c 1 -o gpurun_out/k_walk_n2v python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --updates 0 > gpurun_out/k_ncu2.log 2>&1
WHARF_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:coalesced_emit_kernel --launch-skip 40 -c 1 -o gpurun_out/k_emit python tools/det_modes.py dw25 1 coalesced > gpurun_out/k_ncu3.log 2>&1
WHARF_GRAPH=0 ncu --set full --clock-control none --import-source on -k regex:coalesced_advance_kernel --launch-skip 41 -c 1 -o gpurun_out/k_advance python tools/det_modes.py dw25 1 coalesced > gpurun_out/k_ncu4.log 2>&1
ls -la gpurun_out/*.ncu-rep
