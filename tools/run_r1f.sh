set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python tools/det_modes.py dw25 3 coalesced > gpurun_out/f_det_dw25_w1.txt 2>&1; cat gpurun_out/f_det_dw25_w1.txt
for b in 1 2 5 10; do WHARF_BRANCHES=$b python tools/det_modes.py dw25 3 coalesced 0 8 > gpurun_out/f_det_dw25_w8_b$b.txt 2>&1; cat gpurun_out/f_det_dw25_w8_b$b.txt; done
for b in 1 3; do WHARF_BRANCHES=$b python tools/det_modes.py dw25 3 coalesced 0 2 > gpurun_out/f_det_dw25_w2_b$b.txt 2>&1; cat gpurun_out/f_det_dw25_w2_b$b.txt; done
WHARF_BRANCHES=2 python tools/det_modes.py dw25 3 coalesced > gpurun_out/f_det_dw25_w1_b2.txt 2>&1; cat gpurun_out/f_det_dw25_w1_b2.txt
