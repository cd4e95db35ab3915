set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/det_modes.py dw25 3 coalesced > gpurun_out/f_det_dw25_graph.txt 2>&1; cat gpurun_out/f_det_dw25_graph.txt
WHARF_GRAPH=0 python tools/det_modes.py dw25 3 coalesced > gpurun_out/f_det_dw25_nograph.txt 2>&1; cat gpurun_out/f_det_dw25_nograph.txt
WHARF_EAGER_INIT=1 python bench.py --no-e2e --updates 0 --no-cpu-baseline > gpurun_out/f_n2v20_eager.json 2> gpurun_out/f_n2v20_eager.err
WHARF_GRAPH=0 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/f_launches_dw25.csv python tools/det_modes.py dw25 1 coalesced > gpurun_out/f_ncu_dw25.log 2>&1
WHARF_EAGER_INIT=1 ncu --set full --clock-control none --import-source on -k regex:n2v_share_init -c 1 -o gpurun_out/f_share_init python bench.py --no-e2e --updates 0 --no-cpu-baseline --steps 1 > gpurun_out/f_ncu_share.log 2>&1
tail -2 gpurun_out/f_ncu_share.log
