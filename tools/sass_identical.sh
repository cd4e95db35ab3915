# Are the kernels of the working tree instruction-identical to those of <commit>?  Compiles the kernel translation units
# of <commit> in a temporary worktree and compares the SASS instruction lines (cuobjdump -sass) with the objects of the
# current build, per file.  Used when host code changed after the last GPU run: SAME on every line means the device code
# the GPU tests validated is the device code that ships.
# usage: bash tools/sass_identical.sh <commit>
set -e
commit=${1:?commit}
root=$(git rev-parse --show-toplevel)
wt=$(mktemp -d)
git -C "$root" worktree add -f "$wt" "$commit" -q
trap 'git -C "$root" worktree remove --force "$wt"; git -C "$root" worktree prune' EXIT
make -C "$root/dynamicgraphrepresentationlearning_b200/csrc" -j8 > /dev/null
flags="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fvisibility=hidden,-Wno-deprecated-declarations -diag-suppress 177"
insts() { cuobjdump -sass "$1" | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed 's/_GLOBAL__N__[0-9a-f]*_/_G_/g' | md5sum | cut -c1-16; }
for f in kernels_walk kernels_graph kernels_coalesced kernels_index kernels_pairings; do
  ( cd "$wt/dynamicgraphrepresentationlearning_b200/csrc" && nvcc $flags -c $f.cu -o "$wt/$f.o" ) &
done
wait
for f in kernels_walk kernels_graph kernels_coalesced kernels_index kernels_pairings; do
  a=$(insts "$wt/$f.o"); b=$(insts "$root/dynamicgraphrepresentationlearning_b200/csrc/_build/$f.o")
  echo "$f $a $b $([ "$a" = "$b" ] && echo SAME || echo DIFF)"
done
