# The reference's own test file tests/wharfmh.cpp, compiled unchanged against this library
# (tests/cpp/build_reference_callers.py -> tests/cpp/_build/ref_callers/ref_test_wharfmh), run on the GPU box.  Its fixture
# shells out to ./SNAPtoAdj on data/flickr-graph (a file the reference does not ship): here ./SNAPtoAdj is this package's
# device-backed replacement (graph_io) and data/flickr-graph is cora's edge list (tests/golden/cora_edge_list.npz).
# Not yet a pytest case: it has not run on a GPU (written after the round's GPU budget ended); promote it once seen green.
# usage (GPU box): bash tools/run_reference_tests.sh <tag> [filter]
set -x
tag=${1:-reftests}
root=$(cd "$(dirname "$0")/.." && pwd)
exe=$root/tests/cpp/_build/ref_callers/ref_test_wharfmh
[ -x "$exe" ] || { echo "build it first: python tests/cpp/build_reference_callers.py (needs /root/reference)"; exit 1; }
work=$(mktemp -d)
mkdir -p "$work/data" "$root/gpurun_out"
PYTHONPATH=$root python - "$work" <<'PY'
import sys, numpy as np
pairs = np.load("tests/golden/cora_edge_list.npz")["pairs"]
open(sys.argv[1] + "/data/flickr-graph", "w").write("".join(f"{u} {v}\n" for u, v in pairs.tolist()))
PY
printf '#!/bin/sh\nPYTHONPATH=%s exec python -m dynamicgraphrepresentationlearning_b200.graph_io "$@"\n' "$root" > "$work/SNAPtoAdj"
chmod +x "$work/SNAPtoAdj"
( cd "$work" && MINI_GTEST_FILTER=${2:-} timeout 1500 "$exe" > "$root/gpurun_out/${tag}.out" 2> "$root/gpurun_out/${tag}.err" ; echo "exit $?" >> "$root/gpurun_out/${tag}.err" )
grep -E "^\[ (RUN|  +OK|  FAILED|  TOTAL) " "$root/gpurun_out/${tag}.out" | tail -30
tail -5 "$root/gpurun_out/${tag}.err"
