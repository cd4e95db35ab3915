set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for rw in "0 8" "7 8" "0 4" "3 4"; do python tools/det_modes.py dw25 3 coalesced $rw > gpurun_out/f_det.txt 2>&1; cat gpurun_out/f_det.txt; done
WHARF_BRANCHES=2 python tools/det_modes.py dw25 3 coalesced 0 8 2>&1 | tail -1
