// Step-synchronous bulk generation for the deterministic rule (graph/wharfmh.h:275-326 with
// config::deterministic_mode, :289-292,304-309).
//
// In deterministic mode every walk of replica r = walk_id / n consumes the same draw X[r][pos] at position pos,
// so the step is a function of the vertex alone:  F_{r,pos}(u) = N(u)[ X[r][pos] % deg(u) ].  Two walkers of a
// replica that ever stand on the same vertex at the same position are identical from there on.  On RMAT graphs
// the n walkers of a replica collapse quickly (43 % distinct vertices after one step, 4 % after ten, 2.5 % on
// average over 80 positions; measured), so instead of two dependent random DRAM loads per
// walker step this path keeps, per replica, the CLASSES (groups of walkers on the same vertex) and works in
// SEGMENTS of k <= 8 consecutive positions:
//   advance  one thread per class walks its vertex k steps and writes the k vertices class-major:
//            trail[c][0..k) (one 32-byte sector per class) and marks the final vertex in a bitmap;
//   rank     a prefix sum over the bitmap's word popcounts numbers the occupied vertices in ascending order:
//            that number is the class of the next segment (classes that met share it), so the next advance reads
//            descriptors and rows in ascending address order;
//   emit     one thread per walker loads the sector of its class once and stores k corpus cells
//            walks[pos + 1 + j][i] = trail[cls[i]][j]  (coalesced per row), then moves to its merged class.
// Segment lengths grow 1,1,1,1,2,2,4,4,8,8,... because the class count falls fastest at the first positions.
// The corpus is bit-identical to the one-walker-per-thread kernel (tests/test_gpu_parity.py).
#include "engine.hpp"

#include <cub/device/device_scan.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <algorithm>

namespace wharf
{
namespace
{
constexpr int kBlock = 256;
constexpr u32 kMaxSeg = 8;          // positions per segment: 8 x 4 B = one sector per class

struct SegDraws { u64 x[kMaxSeg]; };

__device__ __forceinline__ RowDesc ld_desc(const RowDesc* __restrict__ desc, u32 v)
{
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(desc) + v);
    return RowDesc{t.x, t.y};
}

// cls[i] = i, vertex of class i = start vertex of local walker i, first corpus row of the replica.
__global__ void __launch_bounds__(kBlock) coalesced_init_kernel(u32* __restrict__ cls, u32* __restrict__ cv, u32* __restrict__ row0, u32 n_local,
                                                                u32 rank, u32 world, u32* __restrict__ n_classes)
{
    const u32 i = blockIdx.x * kBlock + threadIdx.x;
    if (i == 0) *n_classes = n_local;
    if (i >= n_local) return;
    const u32 v = i * world + rank;                  // = walk_id % n (wharfmh.h:292)
    cls[i] = i;
    cv[i] = v;
    row0[i] = v;
}

// trail[c][j] = vertex of class c after j + 1 steps from cv[c]; a class on a degree-0 vertex (isolated start or
// sink) turns into kNoVertex and stays there.  The final vertex is marked for the merge.
template <u32 K>
__global__ void __launch_bounds__(kBlock) coalesced_advance_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, const u32* __restrict__ cv,
                                                                   const u32* __restrict__ n_classes, SegDraws draws, u32* __restrict__ trail,
                                                                   u32* __restrict__ bitmap)
{
    const u32 nc = *n_classes;
    for (u32 c = blockIdx.x * kBlock + threadIdx.x; c < nc; c += gridDim.x * kBlock)
    {
        u32 u = cv[c];
        u32 t[K];
#pragma unroll
        for (u32 j = 0; j < K; j++)
        {
            if (u != kNoVertex)
            {
                const RowDesc d = ld_desc(desc, u);
                u = d.degree ? __ldg(pool + d.begin + mod_u64_u32(draws.x[j], d.degree)) : kNoVertex;
            }
            t[j] = u;
        }
        u32* out = trail + (u64) c * K;
        if constexpr (K == 8) { reinterpret_cast<uint4*>(out)[0] = make_uint4(t[0], t[1], t[2], t[3]); reinterpret_cast<uint4*>(out)[1] = make_uint4(t[4], t[5], t[6], t[7]); }
        else if constexpr (K == 4) reinterpret_cast<uint4*>(out)[0] = make_uint4(t[0], t[1], t[2], t[3]);
        else if constexpr (K == 2) reinterpret_cast<uint2*>(out)[0] = make_uint2(t[0], t[1]);
        else out[0] = t[0];
        if (u != kNoVertex) atomicOr(bitmap + (u >> 5), 1u << (u & 31));
    }
}

struct PopcountOf
{
    const u32* bitmap;
    u32 words;
    __device__ u32 operator()(u32 w) const { return w < words ? (u32) __popc(bitmap[w]) : 0u; }
};

// New class of every class = rank of its final vertex among the marked vertices; dead classes share the extra
// class `total`.  Also lays down the vertices of the next classes (ascending) and their count.
template <u32 K>
__global__ void __launch_bounds__(kBlock) coalesced_rank_kernel(const u32* __restrict__ trail, const u32* __restrict__ bitmap, const u32* __restrict__ word_rank,
                                                                u32 words, u32* __restrict__ new_class, u32* __restrict__ cv_next,
                                                                const u32* __restrict__ n_classes, u32* __restrict__ n_classes_next)
{
    const u32 nc = *n_classes;
    const u32 total = word_rank[words];
    for (u32 c = blockIdx.x * kBlock + threadIdx.x; c < nc; c += gridDim.x * kBlock)
    {
        const u32 v = trail[(u64) c * K + (K - 1)];
        u32 r = total;
        if (v != kNoVertex) r = word_rank[v >> 5] + (u32) __popc(bitmap[v >> 5] & ((1u << (v & 31)) - 1u));
        new_class[c] = r;
        cv_next[r] = v;                                // classes that met write the same value
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
    {
        cv_next[total] = kNoVertex;
        *n_classes_next = total + 1;
    }
}

// Corpus cells of the segment: walks[first_row + j][i] = trail[cls[i]][j] for j < stored (the draw at the last
// position is unobservable, wharfmh.h:315-317: it is counted, not stored), then the walker moves to its merged
// class.  Every trail entry that holds a vertex is one step.
template <u32 K, bool kRelabel>
__global__ void __launch_bounds__(kBlock) coalesced_emit_kernel(u32* __restrict__ cls, const u32* __restrict__ trail, const u32* __restrict__ new_class,
                                                                u32* __restrict__ first_row, u64 row_stride, u32 stored, u32 n_local,
                                                                unsigned long long* __restrict__ steps_out)
{
    const u32 i = blockIdx.x * kBlock + threadIdx.x;
    u32 moved = 0;
    if (i < n_local)
    {
        const u32 c = cls[i];
        u32 t[K];
        const u32* src = trail + (u64) c * K;
        if constexpr (K == 8)
        {
            const uint4 a = __ldg(reinterpret_cast<const uint4*>(src)), b = __ldg(reinterpret_cast<const uint4*>(src) + 1);
            t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w;
        }
        else if constexpr (K == 4) { const uint4 a = __ldg(reinterpret_cast<const uint4*>(src)); t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; }
        else if constexpr (K == 2) { const uint2 a = __ldg(reinterpret_cast<const uint2*>(src)); t[0] = a.x; t[1] = a.y; }
        else t[0] = __ldg(src);
        if (kRelabel) cls[i] = __ldg(new_class + c);
        u32* out = first_row + i;
#pragma unroll
        for (u32 j = 0; j < K; j++)
        {
            if (j < stored) __stcs(out + (u64) j * row_stride, t[j]);
            moved += t[j] != kNoVertex;
        }
    }
    block_add_u64(steps_out, moved);
}

// Segment length at advance index a (the advance that produces the vertex of position a + 1).
u32 segment_length(u32 a, u32 remaining)
{
    const u32 want = a < 4 ? 1 : (a < 8 ? 2 : (a < 16 ? 4 : kMaxSeg));
    u32 k = 1;
    while (k * 2 <= want && k * 2 <= remaining) k *= 2;
    return k;
}

struct SegmentCtx
{
    const LaunchCtx& lc;
    const WalkParams& wp;
    u32 words;
    unsigned walker_grid, class_grid;
    u32 *cls, *trail, *new_class, *bitmap, *word_rank;
    void* scan_temp;
    size_t scan_temp_bytes;
    unsigned long long* d_steps;
};

template <u32 K>
void run_segment(const SegmentCtx& s, const SegDraws& draws, const u32* cv, u32* cv_next, const u32* nc, u32* nc_next, u32* first_row, u32 stored, bool merge)
{
    const LaunchCtx& lc = s.lc;
    coalesced_advance_kernel<K><<<s.class_grid, kBlock, 0, lc.stream>>>(s.wp.desc, s.wp.pool, cv, nc, draws, s.trail, s.bitmap);
    lc.count();
    if (merge)
    {
        cub::CountingInputIterator<u32> iota(0);
        cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{s.bitmap, s.words});
        size_t bytes = s.scan_temp_bytes;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(s.scan_temp, bytes, pops, s.word_rank, (int) (s.words + 1), lc.stream));
        lc.count_lib();
        coalesced_rank_kernel<K><<<s.class_grid, kBlock, 0, lc.stream>>>(s.trail, s.bitmap, s.word_rank, s.words, s.new_class, cv_next, nc, nc_next);
        coalesced_emit_kernel<K, true><<<s.walker_grid, kBlock, 0, lc.stream>>>(s.cls, s.trail, s.new_class, first_row, s.wp.w_local, stored, s.wp.n_local, s.d_steps);
        lc.count(2);
        WHARF_CUDA(cudaMemsetAsync(s.bitmap, 0, (u64) s.words * 4, lc.stream));
    }
    else
    {
        coalesced_emit_kernel<K, false><<<s.walker_grid, kBlock, 0, lc.stream>>>(s.cls, s.trail, nullptr, first_row, s.wp.w_local, stored, s.wp.n_local, s.d_steps);
        lc.count();
    }
}
}  // namespace

void walks_generate_coalesced(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps,
                              const std::function<void(u32)>& after_replica)
{
    if (wp.w_local == 0) return;
    const u32 nl = wp.n_local, L = wp.L;
    const u32 words = (u32) ((wp.n + 31) / 32);
    Scratch<u32> cls(lc.stream, nl), cv_a(lc.stream, (u64) nl + 1), cv_b(lc.stream, (u64) nl + 1), new_class(lc.stream, (u64) nl + 1);
    Scratch<u32> trail(lc.stream, ((u64) nl + 1) * kMaxSeg);
    Scratch<u32> bitmap(lc.stream, words), word_rank(lc.stream, (u64) words + 1), counts(lc.stream, 2);
    WHARF_CUDA(cudaMemsetAsync(bitmap.ptr, 0, (u64) words * 4, lc.stream));

    size_t temp_bytes = 0;
    {
        cub::CountingInputIterator<u32> iota(0);
        cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{bitmap.ptr, words});
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, pops, word_rank.ptr, (int) (words + 1), lc.stream));
    }
    Scratch<u8> temp(lc.stream, temp_bytes);

    SegmentCtx s{lc, wp, words, div_up(nl, kBlock), grid_for(nl, kBlock, 8, lc.sm_count), cls.ptr, trail.ptr, new_class.ptr, bitmap.ptr,
                 word_rank.ptr, temp.ptr, temp_bytes, d_steps};
    for (u32 r = 0; r < wp.wpv; r++)
    {
        u32 *cv = cv_a.ptr, *cv_next = cv_b.ptr, *nc = counts.ptr, *nc_next = counts.ptr + 1;
        u32* col = wp.walks + (u64) r * nl;                       // column block of the replica in every corpus row
        coalesced_init_kernel<<<s.walker_grid, kBlock, 0, lc.stream>>>(cls.ptr, cv, col, nl, wp.rank, wp.world, nc);
        lc.count();
        for (u32 a = 0; a < L;)
        {
            const u32 k = segment_length(a, L - a);
            SegDraws draws{};
            for (u32 j = 0; j < k; j++) draws.x[j] = host_draw[(u64) r * L + a + j];
            const u32 stored = std::min(k, L - 1 - a);            // cells a + 1 .. a + stored exist; the last draw is only counted
            const bool merge = a + k < L;
            u32* first_row = col + (u64) (a + 1) * wp.w_local;
            switch (k)
            {
                case 1: run_segment<1>(s, draws, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
                case 2: run_segment<2>(s, draws, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
                case 4: run_segment<4>(s, draws, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
                default: run_segment<8>(s, draws, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
            }
            if (merge) { std::swap(cv, cv_next); std::swap(nc, nc_next); }
            a += k;
        }
        WHARF_CUDA(cudaGetLastError());
        if (after_replica) after_replica(r);
    }
}

}  // namespace wharf
