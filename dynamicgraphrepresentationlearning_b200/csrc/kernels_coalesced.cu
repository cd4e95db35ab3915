// Step-synchronous bulk generation for the deterministic rule (graph/wharfmh.h:275-326 with
// config::deterministic_mode, :289-292,304-309).
//
// In deterministic mode every walk of replica r = walk_id / n consumes the same draw X[r][pos] at position pos,
// so the step is a function of the vertex alone:  F_{r,pos}(u) = N(u)[ X[r][pos] % deg(u) ].  Two walkers of a
// replica that ever stand on the same vertex at the same position are identical from there on.  On RMAT graphs
// the n walkers of a replica collapse quickly (46 % distinct vertices after one step, 5 % after ten, 3 % on
// average over 80 positions; measured), so instead of two dependent random DRAM loads per
// walker step this path keeps, per replica, the CLASSES (groups of walkers on the same vertex) and works in
// SEGMENTS of k <= 7 consecutive positions.  A class owns one RECORD of R = 2, 4 or 8 words: the R - 1 vertices of
// the segment followed by the class it belongs to in the next segment, so a walker needs ONE aligned vector load
// (8, 16 or 32 bytes: never more than one sector) per segment:
//   advance  one thread per class walks its vertex k steps, writes them into the record and marks the final
//            vertex in a bitmap;
//   scan     exclusive prefix sum over the bitmap's word popcounts: the rank of an occupied vertex among the
//            occupied vertices is the class of the next segment (classes that met share it, and class order =
//            vertex order, so the next advance reads descriptors and rows at ascending addresses);
//   expand   one thread per bitmap word writes the vertices of the next classes (ascending), the (rank, bits)
//            table used by relabel, and clears the word;
//   relabel  one thread per class writes the next class into the last word of its record;
//   emit     one thread per walker loads the record of its class and stores the segment's corpus cells
//            walks[pos + 1 + j][i] = record[cls[i]][j]  (coalesced per row), then moves to the next class.
// Segment lengths grow 1,1,1,1,3,3,3,7,7,... because the class count falls fastest at the first positions.
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
constexpr u32 kEmitWalkers = 4;     // walkers per emit thread
constexpr u32 kMaxRec = 8;          // words per class record: 7 vertices + next class = one 32-byte sector

struct SegDraws { u64 x[kMaxRec]; };

__device__ __forceinline__ RowDesc ld_desc(const RowDesc* __restrict__ desc, u32 v)
{
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(desc) + v);
    return RowDesc{t.x, t.y};
}

template <u32 R> struct Record;
template <> struct Record<2> { using vec = uint2; };
template <> struct Record<4> { using vec = uint4; };
template <> struct Record<8> { using vec = uint4; };

template <u32 R>
__device__ __forceinline__ void load_record(const u32* __restrict__ rec, u32 c, u32 (&t)[R])
{
    if constexpr (R == 2) { const uint2 a = __ldg(reinterpret_cast<const uint2*>(rec) + c); t[0] = a.x; t[1] = a.y; }
    else if constexpr (R == 4) { const uint4 a = __ldg(reinterpret_cast<const uint4*>(rec) + c); t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; }
    else
    {
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(rec) + 2 * (u64) c), b = __ldg(reinterpret_cast<const uint4*>(rec) + 2 * (u64) c + 1);
        t[0] = a.x; t[1] = a.y; t[2] = a.z; t[3] = a.w; t[4] = b.x; t[5] = b.y; t[6] = b.z; t[7] = b.w;
    }
}

template <u32 R>
__device__ __forceinline__ void store_record(u32* __restrict__ rec, u32 c, const u32 (&t)[R])
{
    if constexpr (R == 2) reinterpret_cast<uint2*>(rec)[c] = make_uint2(t[0], t[1]);
    else if constexpr (R == 4) reinterpret_cast<uint4*>(rec)[c] = make_uint4(t[0], t[1], t[2], t[3]);
    else
    {
        reinterpret_cast<uint4*>(rec)[2 * (u64) c] = make_uint4(t[0], t[1], t[2], t[3]);
        reinterpret_cast<uint4*>(rec)[2 * (u64) c + 1] = make_uint4(t[4], t[5], t[6], t[7]);
    }
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

// record[c][j] = vertex of class c after j + 1 steps from cv[c], j < k; a class on a degree-0 vertex (isolated
// start or sink) turns into kNoVertex and stays there.  The final vertex is marked for the merge.
template <u32 R>
__global__ void __launch_bounds__(kBlock) coalesced_advance_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, const u32* __restrict__ cv,
                                                                   const u32* __restrict__ n_classes, SegDraws draws, u32 k, u32* __restrict__ rec,
                                                                   u32* __restrict__ bitmap)
{
    const u32 nc = *n_classes;
    for (u32 c = blockIdx.x * kBlock + threadIdx.x; c < nc; c += gridDim.x * kBlock)
    {
        u32 u = cv[c];
        u32 t[R];
#pragma unroll
        for (u32 j = 0; j < R - 1; j++)
        {
            if (j < k && u != kNoVertex)
            {
                const RowDesc d = ld_desc(desc, u);
                u = d.degree ? __ldg(pool + d.begin + mod_u64_u32(draws.x[j], d.degree)) : kNoVertex;
            }
            t[j] = j < k ? u : kNoVertex;
        }
        t[R - 1] = u;                                  // final vertex; relabel turns it into the next class
        store_record<R>(rec, c, t);
        if (bitmap && u != kNoVertex) atomicOr(bitmap + (u >> 5), 1u << (u & 31));
    }
}

struct PopcountOf
{
    const u32* bitmap;
    u32 words;
    __device__ u32 operator()(u32 w) const { return w < words ? (u32) __popc(bitmap[w]) : 0u; }
};

// One thread per bitmap word: the occupied vertices, in ascending order, are the classes of the next segment.
// Leaves table[w] = (rank of the word's first occupied vertex, the word's bits) for relabel and clears the bitmap.
__global__ void __launch_bounds__(kBlock) coalesced_expand_kernel(u32* __restrict__ bitmap, const u32* __restrict__ word_rank, u32 words,
                                                                  uint2* __restrict__ table, u32* __restrict__ cv_next, u32* __restrict__ n_classes_next)
{
    const u32 w = blockIdx.x * kBlock + threadIdx.x;
    if (w == 0)
    {
        const u32 total = word_rank[words];
        cv_next[total] = kNoVertex;                    // the class of every walker that stopped
        *n_classes_next = total + 1;
    }
    if (w >= words) return;
    u32 bits = bitmap[w];
    const u32 r = word_rank[w];
    table[w] = make_uint2(r, bits);
    if (!bits) return;
    bitmap[w] = 0;
    u32* out = cv_next + r;
    while (bits)
    {
        *out++ = (w << 5) + (u32) (__ffs(bits) - 1);
        bits &= bits - 1;
    }
}

// Last word of every record: final vertex -> its rank among the occupied vertices (dead classes share `total`).
template <u32 R>
__global__ void __launch_bounds__(kBlock) coalesced_relabel_kernel(u32* __restrict__ rec, const uint2* __restrict__ table, const u32* __restrict__ word_rank,
                                                                   u32 words, const u32* __restrict__ n_classes)
{
    const u32 nc = *n_classes;
    const u32 total = word_rank[words];
    for (u32 c = blockIdx.x * kBlock + threadIdx.x; c < nc; c += gridDim.x * kBlock)
    {
        u32* slot = rec + (u64) c * R + (R - 1);
        const u32 v = *slot;
        u32 r = total;
        if (v != kNoVertex)
        {
            const uint2 e = __ldg(table + (v >> 5));
            r = e.x + (u32) __popc(e.y & ((1u << (v & 31)) - 1u));
        }
        *slot = r;
    }
}

// Corpus cells of the segment: walks[first_row + j][i] = record[cls[i]][j] for j < stored (the draw at the last
// position is unobservable, wharfmh.h:315-317: it is counted, not stored), then the walker moves to its next
// class.  Every record entry that holds a vertex is one step.
template <u32 R, bool kRelabel>
__global__ void __launch_bounds__(kBlock) coalesced_emit_kernel(u32* __restrict__ cls, const u32* __restrict__ rec, u32* __restrict__ first_row,
                                                                u64 row_stride, u32 stored, u32 n_local, unsigned long long* __restrict__ steps_out)
{
    // kEmitWalkers walkers per thread, kBlock apart (every access stays coalesced): the class loads of all of them are
    // in flight before the first record gather, the gathers before the first store
    const u32 i0 = blockIdx.x * (kBlock * kEmitWalkers) + threadIdx.x;
    u32 c[kEmitWalkers];
#pragma unroll
    for (u32 v = 0; v < kEmitWalkers; v++)
    {
        const u32 i = i0 + v * kBlock;
        c[v] = i < n_local ? __ldcs(cls + i) : 0xFFFFFFFFu;
    }
    u32 t[kEmitWalkers][R];
#pragma unroll
    for (u32 v = 0; v < kEmitWalkers; v++)
        if (c[v] != 0xFFFFFFFFu) load_record<R>(rec, c[v], t[v]);
    u32 moved = 0;
#pragma unroll
    for (u32 v = 0; v < kEmitWalkers; v++)
    {
        const u32 i = i0 + v * kBlock;
        if (c[v] == 0xFFFFFFFFu) continue;
        if (kRelabel) __stcs(cls + i, t[v][R - 1]);
        u32* out = first_row + i;
#pragma unroll
        for (u32 j = 0; j < R - 1; j++)
        {
            if (j < stored) __stcs(out + (u64) j * row_stride, t[v][j]);
            moved += t[v][j] != kNoVertex;
        }
    }
    block_add_u64(steps_out, moved);
}

// Record size (words) of the segment whose first advance index is a: R - 1 positions per segment.
u32 record_words(u32 a) { return a < 4 ? 2 : (a < 13 ? 4 : kMaxRec); }

struct SegmentCtx
{
    const LaunchCtx& lc;
    const WalkParams& wp;
    u32 words;
    unsigned walker_grid, emit_grid, class_grid, word_grid;
    u32 *cls, *rec, *bitmap, *word_rank;
    uint2* table;
    void* scan_temp;
    size_t scan_temp_bytes;
    unsigned long long* d_steps;
};

template <u32 R>
void run_segment(const SegmentCtx& s, const SegDraws& draws, u32 k, const u32* cv, u32* cv_next, const u32* nc, u32* nc_next, u32* first_row, u32 stored,
                 bool merge)
{
    const LaunchCtx& lc = s.lc;
    coalesced_advance_kernel<R><<<s.class_grid, kBlock, 0, lc.stream>>>(s.wp.desc, s.wp.pool, cv, nc, draws, k, s.rec, merge ? s.bitmap : nullptr);
    lc.count();
    if (merge)
    {
        cub::CountingInputIterator<u32> iota(0);
        cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{s.bitmap, s.words});
        size_t bytes = s.scan_temp_bytes;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(s.scan_temp, bytes, pops, s.word_rank, (int) (s.words + 1), lc.stream));
        lc.count_lib();
        coalesced_expand_kernel<<<s.word_grid, kBlock, 0, lc.stream>>>(s.bitmap, s.word_rank, s.words, s.table, cv_next, nc_next);
        coalesced_relabel_kernel<R><<<s.class_grid, kBlock, 0, lc.stream>>>(s.rec, s.table, s.word_rank, s.words, nc);
        coalesced_emit_kernel<R, true><<<s.emit_grid, kBlock, 0, lc.stream>>>(s.cls, s.rec, first_row, s.wp.w_local, stored, s.wp.n_local, s.d_steps);
        lc.count(3);
    }
    else
    {
        coalesced_emit_kernel<R, false><<<s.emit_grid, kBlock, 0, lc.stream>>>(s.cls, s.rec, first_row, s.wp.w_local, stored, s.wp.n_local, s.d_steps);
        lc.count();
    }
}
}  // namespace

// Scratch of the schedule plus, when no per-replica callback is wanted, the whole pass as one CUDA graph: ~2700
// kernel launches per generate (scale-25: most of them a few microseconds long) replayed without host work.
struct CoalescedPlan
{
    Scratch<u32> cls, cv_a, cv_b, rec, bitmap, word_rank, counts;
    Scratch<uint2> table;
    Scratch<u8> temp;
    size_t temp_bytes = 0;
    u32 n_local = 0, words = 0;
    cudaGraphExec_t exec = nullptr;
    u64 graph_launches = 0, graph_lib_calls = 0;
    struct Key
    {
        const void *desc, *pool, *walks, *steps;
        u64 n, w_local, draw_hash;
        u32 n_local, wpv, L, rank, world;
        bool operator==(const Key& o) const
        {
            return desc == o.desc && pool == o.pool && walks == o.walks && steps == o.steps && n == o.n && w_local == o.w_local &&
                   draw_hash == o.draw_hash && n_local == o.n_local && wpv == o.wpv && L == o.L && rank == o.rank && world == o.world;
        }
    } key{};
    void drop_graph()
    {
        if (exec) cudaGraphExecDestroy(exec);
        exec = nullptr;
    }
    ~CoalescedPlan() { drop_graph(); }
};

void coalesced_plan_release(CoalescedPlan*& plan)
{
    delete plan;
    plan = nullptr;
}

namespace
{
void plan_buffers(CoalescedPlan& pl, const LaunchCtx& lc, u32 nl, u32 words)
{
    if (pl.n_local == nl && pl.words == words && pl.cls.ptr) return;
    pl.drop_graph();
    pl.cls.allocate(lc.stream, nl);
    pl.cv_a.allocate(lc.stream, (u64) nl + 1);
    pl.cv_b.allocate(lc.stream, (u64) nl + 1);
    pl.rec.allocate(lc.stream, ((u64) nl + 1) * kMaxRec);
    pl.bitmap.allocate(lc.stream, words);
    pl.word_rank.allocate(lc.stream, (u64) words + 1);
    pl.counts.allocate(lc.stream, 2);
    pl.table.allocate(lc.stream, words);
    cub::CountingInputIterator<u32> iota(0);
    cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{pl.bitmap.ptr, words});
    pl.temp_bytes = 0;
    WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, pl.temp_bytes, pops, pl.word_rank.ptr, (int) (words + 1), lc.stream));
    pl.temp.allocate(lc.stream, pl.temp_bytes);
    pl.n_local = nl;
    pl.words = words;
}

void enqueue_pass(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps, CoalescedPlan& pl,
                  const std::function<void(u32)>& after_replica)
{
    const u32 nl = wp.n_local, L = wp.L, words = pl.words;
    WHARF_CUDA(cudaMemsetAsync(pl.bitmap.ptr, 0, (u64) words * 4, lc.stream));
    SegmentCtx s{lc, wp, words, div_up(nl, kBlock), div_up(nl, kBlock * kEmitWalkers), grid_for(nl, kBlock, 8, lc.sm_count), div_up(words, kBlock), pl.cls.ptr, pl.rec.ptr, pl.bitmap.ptr,
                 pl.word_rank.ptr, pl.table.ptr, pl.temp.ptr, pl.temp_bytes, d_steps};
    for (u32 r = 0; r < wp.wpv; r++)
    {
        u32 *cv = pl.cv_a.ptr, *cv_next = pl.cv_b.ptr, *nc = pl.counts.ptr, *nc_next = pl.counts.ptr + 1;
        u32* col = wp.walks + (u64) r * nl;                       // column block of the replica in every corpus row
        coalesced_init_kernel<<<s.walker_grid, kBlock, 0, lc.stream>>>(pl.cls.ptr, cv, col, nl, wp.rank, wp.world, nc);
        lc.count();
        for (u32 a = 0; a < L;)
        {
            const u32 R = record_words(a);
            const u32 k = std::min(R - 1, L - a);                 // draws a .. a + k - 1
            SegDraws draws{};
            for (u32 j = 0; j < k; j++) draws.x[j] = host_draw[(u64) r * L + a + j];
            const u32 stored = std::min(k, L - 1 - a);            // cells a + 1 .. a + stored exist; the last draw is only counted
            const bool merge = a + k < L;
            u32* first_row = col + (u64) (a + 1) * wp.w_local;
            switch (R)
            {
                case 2: run_segment<2>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
                case 4: run_segment<4>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
                default: run_segment<8>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
            }
            if (merge) { std::swap(cv, cv_next); std::swap(nc, nc_next); }
            a += k;
        }
        WHARF_CUDA(cudaGetLastError());
        if (after_replica) after_replica(r);
    }
}

bool graphs_enabled()
{
    static const bool on = [] { const char* e = std::getenv("WHARF_GRAPH"); return !e || std::atoi(e) != 0; }();
    return on;
}
}  // namespace

void walks_generate_coalesced(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps,
                              const std::function<void(u32)>& after_replica, CoalescedPlan** plan_slot)
{
    if (wp.w_local == 0) return;
    const u32 words = (u32) ((wp.n + 31) / 32);
    CoalescedPlan local;
    CoalescedPlan* pl = &local;
    if (plan_slot)
    {
        if (!*plan_slot) *plan_slot = new CoalescedPlan();
        pl = *plan_slot;
    }
    plan_buffers(*pl, lc, wp.n_local, words);
    if (after_replica || !plan_slot || !graphs_enabled())
    {
        enqueue_pass(lc, wp, host_draw, d_steps, *pl, after_replica);
        return;
    }
    u64 draw_hash = 1469598103934665603ull;
    for (u64 i = 0; i < (u64) wp.wpv * wp.L; i++) draw_hash = (draw_hash ^ host_draw[i]) * 1099511628211ull;
    const CoalescedPlan::Key key{wp.desc, wp.pool, wp.walks, d_steps, wp.n, wp.w_local, draw_hash, wp.n_local, wp.wpv, wp.L, wp.rank, wp.world};
    if (!pl->exec || !(pl->key == key))
    {
        pl->drop_graph();
        u64 launches = 0, lib_calls = 0;
        LaunchCtx cap = lc;
        cap.launches = &launches;
        cap.library_calls = &lib_calls;
        cudaGraph_t graph = nullptr;
        WHARF_CUDA(cudaStreamBeginCapture(lc.stream, cudaStreamCaptureModeThreadLocal));
        try
        {
            enqueue_pass(cap, wp, host_draw, d_steps, *pl, nullptr);
        }
        catch (...)
        {
            cudaStreamEndCapture(lc.stream, &graph);
            if (graph) cudaGraphDestroy(graph);
            cudaGetLastError();
            throw;
        }
        WHARF_CUDA(cudaStreamEndCapture(lc.stream, &graph));
        const cudaError_t err = cudaGraphInstantiate(&pl->exec, graph, 0);
        cudaGraphDestroy(graph);
        if (err != cudaSuccess) { pl->exec = nullptr; WHARF_CUDA(err); }
        pl->key = key;
        pl->graph_launches = launches;
        pl->graph_lib_calls = lib_calls;
    }
    WHARF_CUDA(cudaGraphLaunch(pl->exec, lc.stream));
    lc.count(pl->graph_launches);
    lc.count_lib(pl->graph_lib_calls);
}

}  // namespace wharf
