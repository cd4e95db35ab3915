set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python bench.py > gpurun_out/p_bench_n2v20.json 2> gpurun_out/p_bench_n2v20.err; tail -c 300 gpurun_out/p_bench_n2v20.err
python - <<EOF2
import json
d = json.loads([l for l in open("gpurun_out/p_bench_n2v20.json") if l.startswith("{")][-1])
print(d["value"] / 1e9, d["ms_per_step"], d["clocks"], d["index_build_ms"], d["e2e"]["value"] / 1e9, d["gpu_launches"])
EOF2
