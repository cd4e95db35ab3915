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

#include <cub/block/block_reduce.cuh>
#include <cub/block/block_scan.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

#include <algorithm>
#include <cstdlib>
#include <vector>

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
        // one 256-bit load (sm_100: LDG.E.ENL2.256): the whole sector in one request - the scattered record gather is
        // bound by L1 wavefronts (one per distinct sector and instruction), and two 128-bit loads would cost two
        asm volatile("ld.global.nc.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                     : "=r"(t[0]), "=r"(t[1]), "=r"(t[2]), "=r"(t[3]), "=r"(t[4]), "=r"(t[5]), "=r"(t[6]), "=r"(t[7])
                     : "l"(rec + 8 * (u64) c));
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
                                                                u32 rank, u32 world, u64 n, u32* __restrict__ n_classes)
{
    const u32 i = blockIdx.x * kBlock + threadIdx.x;
    if (i == 0) *n_classes = n_local;
    if (i >= n_local) return;
    const u32 v = shard_start(i, rank, world, n);    // = walk_id % n (wharfmh.h:292)
    cls[i] = i;
    cv[i] = v;
    row0[i] = v;
}

// First step of every walker of up to kFirstReplicas WHOLE replicas at once: walker i of such a replica starts on vertex
// i, so at position 0 all of them read the same descriptor and the same row - one thread per vertex loads them once and
// writes corpus row 1 of each replica (cell = N(i)[X[r][0] % deg]), which the replica's first advance then reads instead
// of the graph (ten replicas: ten sweeps over the 2.4 GB adjacency become one; scale-25: 71.0 -> 68 ms).
constexpr u32 kFirstReplicas = 16;
struct FirstDraws { u64 x[kFirstReplicas]; };

__global__ void __launch_bounds__(kBlock) coalesced_first_steps_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, u32 n, u32 shift,
                                                                       FirstDraws draws, u32 replicas, u32* __restrict__ row1, u64 replica_stride)
{
    const u32 v = blockIdx.x * kBlock + threadIdx.x;
    if (v >= n) return;
    const RowDesc d = ld_desc(desc, v);
    const double inv = d.degree ? reciprocal_of(d.degree) : 0.0;
#pragma unroll 4
    for (u32 r = 0; r < replicas; r++)
    {
        const u32 u = d.degree ? __ldg(pool + d.begin + ((u64) mod_u64_u32_rcp(draws.x[r], d.degree, inv) << shift)) : kNoVertex;
        __stcs(row1 + (u64) r * replica_stride + v, u);
    }
}

// record[c][j] = vertex of class c after j + 1 steps from cv[c], j < k; a class on a degree-0 vertex (isolated
// start or sink) turns into kNoVertex and stays there.  The final vertex is marked for the merge.
template <u32 R>
__global__ void __launch_bounds__(kBlock) coalesced_advance_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, const u32* __restrict__ cv,
                                                                   const u32* __restrict__ n_classes, SegDraws draws, u32 k, u32* __restrict__ rec,
                                                                   u32* __restrict__ bitmap, u32 shift, const u32* __restrict__ first_step)
{
    const u32 nc = *n_classes;
    for (u32 c = blockIdx.x * kBlock + threadIdx.x; c < nc; c += gridDim.x * kBlock)
    {
        u32 u = cv[c];
        u32 t[R];
#pragma unroll
        for (u32 j = 0; j < R - 1; j++)
        {
            if (j == 0 && first_step) u = __ldcs(first_step + c);          // the replica's first segment: class c = walker c (coalesced_first_steps_kernel)
            else if (j < k && u != kNoVertex)
            {
                const RowDesc d = ld_desc(desc, u);
                u = d.degree ? __ldg(pool + d.begin + ((u64) mod_u64_u32(draws.x[j], d.degree) << shift)) : kNoVertex;
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

// Ranking of the occupied vertices, two launches over the bitmap (n / 32 words):
//   tile_counts  one block per tile of kTileWords words: number of occupied vertices in the tile;
//   expand       the block sums the counts of the tiles before its own (at most kMaxTiles of them), scans its tile and,
//                for every word, writes table[w] = (rank of the word's first occupied vertex, the word's bits), lays
//                down the vertices of the next classes (ascending: the occupied vertices) and clears the word.
// table[words] = (number of occupied vertices, 0): the class shared by every walker that stopped.
constexpr u32 kWordsPerThread = 4;
constexpr u32 kTileWords = kBlock * kWordsPerThread;
constexpr u32 kMaxTiles = 4096;      // larger bitmaps (n > 134 M vertices) take the library scan

__global__ void __launch_bounds__(kBlock) coalesced_tile_counts_kernel(const u32* __restrict__ bitmap, u32 words, u32* __restrict__ tile_count)
{
    using BlockReduce = cub::BlockReduce<u32, kBlock>;
    __shared__ typename BlockReduce::TempStorage temp;
    const u32 w0 = blockIdx.x * kTileWords + threadIdx.x * kWordsPerThread;
    u32 c = 0;
    if (w0 + kWordsPerThread <= words)
    {
#pragma unroll
        for (u32 q = 0; q < kWordsPerThread / 4; q++)
        {
            const uint4 b = reinterpret_cast<const uint4*>(bitmap + w0)[q];
            c += __popc(b.x) + __popc(b.y) + __popc(b.z) + __popc(b.w);
        }
    }
    else
        for (u32 w = w0; w < words; w++) c += __popc(bitmap[w]);
    const u32 sum = BlockReduce(temp).Sum(c);
    if (threadIdx.x == 0) tile_count[blockIdx.x] = sum;
}

__global__ void __launch_bounds__(kBlock) coalesced_expand_kernel(u32* __restrict__ bitmap, const u32* __restrict__ tile_count, u32 words,
                                                                  uint2* __restrict__ table, u32* __restrict__ cv_next, u32* __restrict__ n_classes_next)
{
    using BlockReduce = cub::BlockReduce<u32, kBlock>;
    using BlockScan = cub::BlockScan<u32, kBlock>;
    __shared__ union { typename BlockReduce::TempStorage reduce; typename BlockScan::TempStorage scan; } temp;
    __shared__ u32 tile_base;
    u32 before = 0;
    for (u32 t = threadIdx.x; t < blockIdx.x; t += kBlock) before += tile_count[t];
    before = BlockReduce(temp.reduce).Sum(before);
    if (threadIdx.x == 0) tile_base = before;
    __syncthreads();
    static_assert(kWordsPerThread == 4, "one uint4 of bitmap words per thread");
    const u32 w0 = blockIdx.x * kTileWords + threadIdx.x * kWordsPerThread;
    const bool full = w0 + kWordsPerThread <= words;
    u32 bits[kWordsPerThread] = {0, 0, 0, 0};
    if (full)
    {
        const uint4 b = *reinterpret_cast<const uint4*>(bitmap + w0);
        bits[0] = b.x; bits[1] = b.y; bits[2] = b.z; bits[3] = b.w;
    }
    else
        for (u32 q = 0; q < kWordsPerThread; q++) bits[q] = w0 + q < words ? bitmap[w0 + q] : 0u;
    const u32 mine = __popc(bits[0]) + __popc(bits[1]) + __popc(bits[2]) + __popc(bits[3]);
    u32 offset, tile_total;
    BlockScan(temp.scan).ExclusiveSum(mine, offset, tile_total);
    u32 r = tile_base + offset;
    u32 first[kWordsPerThread];
#pragma unroll
    for (u32 q = 0; q < kWordsPerThread; q++)
    {
        first[q] = r;
        u32 b = bits[q];
        while (b)
        {
            cv_next[r++] = ((w0 + q) << 5) + (u32) (__ffs(b) - 1);
            b &= b - 1;
        }
    }
    if (full)
    {
        uint4* t = reinterpret_cast<uint4*>(table + w0);
        t[0] = make_uint4(first[0], bits[0], first[1], bits[1]);
        t[1] = make_uint4(first[2], bits[2], first[3], bits[3]);
        if (mine) *reinterpret_cast<uint4*>(bitmap + w0) = make_uint4(0u, 0u, 0u, 0u);
    }
    else
        for (u32 q = 0; q < kWordsPerThread && w0 + q < words; q++)
        {
            table[w0 + q] = make_uint2(first[q], bits[q]);
            bitmap[w0 + q] = 0;
        }
    if (blockIdx.x == gridDim.x - 1 && threadIdx.x == 0)
    {
        const u32 total = tile_base + tile_total;
        table[words] = make_uint2(total, 0u);
        cv_next[total] = kNoVertex;                    // the class of every walker that stopped
        *n_classes_next = total + 1;
    }
}

// The same from a library scan of the word popcounts (bitmaps of more than kMaxTiles tiles).
__global__ void __launch_bounds__(kBlock) coalesced_expand_scanned_kernel(u32* __restrict__ bitmap, const u32* __restrict__ word_rank, u32 words,
                                                                          uint2* __restrict__ table, u32* __restrict__ cv_next, u32* __restrict__ n_classes_next)
{
    const u32 w = blockIdx.x * kBlock + threadIdx.x;
    if (w == 0)
    {
        const u32 total = word_rank[words];
        table[words] = make_uint2(total, 0u);
        cv_next[total] = kNoVertex;
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
__global__ void __launch_bounds__(kBlock) coalesced_relabel_kernel(u32* __restrict__ rec, const uint2* __restrict__ table, u32 words,
                                                                   const u32* __restrict__ n_classes)
{
    const u32 nc = *n_classes;
    const u32 total = table[words].x;
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
    u32 nl;                          // walkers of the replica on this rank
    unsigned walker_grid, emit_grid, class_grid, word_grid, tiles;
    u32 *cls, *rec, *bitmap, *word_rank;
    uint2* table;
    void* scan_temp;
    size_t scan_temp_bytes;
    unsigned long long* d_steps;
};

// WHARF_RANK_SCAN=1 takes the library-scan ranking at any size (it is otherwise only reached beyond 134 M vertices).
bool force_library_scan()
{
    const char* e = std::getenv("WHARF_RANK_SCAN");
    return e && std::atoi(e) != 0;
}

template <u32 R>
void run_segment(const SegmentCtx& s, const SegDraws& draws, u32 k, const u32* cv, u32* cv_next, const u32* nc, u32* nc_next, u32* first_row, u32 stored,
                 bool merge, const u32* first_step = nullptr)
{
    const LaunchCtx& lc = s.lc;
    coalesced_advance_kernel<R><<<s.class_grid, kBlock, 0, lc.stream>>>(s.wp.desc, s.wp.pool, cv, nc, draws, k, s.rec, merge ? s.bitmap : nullptr, s.wp.shift,
                                                                        first_step);
    if (first_step) stored = 0;                               // the segment's only row is the array just read
    lc.count();
    if (merge)
    {
        if (s.tiles <= kMaxTiles && !force_library_scan())
        {
            coalesced_tile_counts_kernel<<<s.tiles, kBlock, 0, lc.stream>>>(s.bitmap, s.words, s.word_rank);
            coalesced_expand_kernel<<<s.tiles, kBlock, 0, lc.stream>>>(s.bitmap, s.word_rank, s.words, s.table, cv_next, nc_next);
            lc.count(2);
        }
        else
        {
            cub::CountingInputIterator<u32> iota(0);
            cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{s.bitmap, s.words});
            size_t bytes = s.scan_temp_bytes;
            WHARF_CUDA(cub::DeviceScan::ExclusiveSum(s.scan_temp, bytes, pops, s.word_rank, (int) (s.words + 1), lc.stream));
            lc.count_lib();
            coalesced_expand_scanned_kernel<<<s.word_grid, kBlock, 0, lc.stream>>>(s.bitmap, s.word_rank, s.words, s.table, cv_next, nc_next);
            lc.count();
        }
        coalesced_relabel_kernel<R><<<s.class_grid, kBlock, 0, lc.stream>>>(s.rec, s.table, s.words, nc);
        coalesced_emit_kernel<R, true><<<s.emit_grid, kBlock, 0, lc.stream>>>(s.cls, s.rec, first_row, s.wp.w_local, stored, s.nl, s.d_steps);
        lc.count(2);
    }
    else
    {
        coalesced_emit_kernel<R, false><<<s.emit_grid, kBlock, 0, lc.stream>>>(s.cls, s.rec, first_row, s.wp.w_local, stored, s.nl, s.d_steps);
        lc.count();
    }
}
}  // namespace

// Scratch of the schedule plus, when no per-replica callback is wanted, the whole pass as one CUDA graph: ~2700
// kernel launches per generate (scale-25: most of them a few microseconds long) replayed without host work.  The
// replicas are independent, so the graph runs them on up to `branches` parallel branches (each with its own
// scratch): on a shard of a multi-rank run most kernels are too small to fill the GPU on their own.
struct CoalescedBranch
{
    Scratch<u32> cls, cv_a, cv_b, rec, bitmap, word_rank, counts;
    Scratch<uint2> table;
    Scratch<u8> temp;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
};

struct CoalescedPlan
{
    std::vector<CoalescedBranch> branch;
    size_t temp_bytes = 0;
    u32 n_local = 0, words = 0;
    cudaEvent_t fork = nullptr;
    cudaGraphExec_t exec = nullptr;
    u64 graph_launches = 0, graph_lib_calls = 0;
    struct Key
    {
        const void *desc, *pool, *walks, *steps;
        u64 n, w_local, draw_hash;
        u32 n_local, wpv, L, rank, world, whole;
        bool operator==(const Key& o) const
        {
            return desc == o.desc && pool == o.pool && walks == o.walks && steps == o.steps && n == o.n && w_local == o.w_local &&
                   draw_hash == o.draw_hash && n_local == o.n_local && wpv == o.wpv && L == o.L && rank == o.rank && world == o.world &&
                   whole == o.whole;
        }
    } key{};
    void drop_graph()
    {
        if (exec) cudaGraphExecDestroy(exec);
        exec = nullptr;
    }
    ~CoalescedPlan()
    {
        drop_graph();
        for (auto& b : branch)
        {
            if (b.stream) cudaStreamDestroy(b.stream);
            if (b.done) cudaEventDestroy(b.done);
        }
        if (fork) cudaEventDestroy(fork);
    }
};

void coalesced_plan_release(CoalescedPlan*& plan)
{
    delete plan;
    plan = nullptr;
}

namespace
{
// Parallel branches of the graph: two while one replica's kernels fill the GPU (n_local in the tens of millions; the
// small class kernels of one replica then overlap the other's emit), more on smaller shards (measured at scale-25:
// 78.6 -> 73.9 ms on one GPU with 2 branches, 23.8 -> 19.4 ms on a shard of eight with 5+).  WHARF_BRANCHES overrides.
u32 wanted_branches(u32 nl, u32 jobs)
{
    u64 b = std::max<u64>(jobs <= 3 ? 3 : 2, (48ull << 20) / std::max<u32>(nl, 1));
    if (const char* e = std::getenv("WHARF_BRANCHES")) b = (u64) std::max(1, std::atoi(e));
    return (u32) std::max<u64>(1, std::min<u64>(b, jobs));
}

void plan_buffers(CoalescedPlan& pl, const LaunchCtx& lc, u32 nl, u32 words, u32 branches)
{
    if (pl.n_local == nl && pl.words == words && pl.branch.size() >= branches) return;
    pl.drop_graph();
    if (pl.n_local != nl || pl.words != words)
    {
        for (auto& b : pl.branch) { if (b.stream) cudaStreamDestroy(b.stream); if (b.done) cudaEventDestroy(b.done); }
        pl.branch.clear();
    }
    pl.n_local = nl;
    pl.words = words;
    if (!pl.fork) WHARF_CUDA(cudaEventCreateWithFlags(&pl.fork, cudaEventDisableTiming));
    while (pl.branch.size() < branches)
    {
        pl.branch.emplace_back();
        CoalescedBranch& b = pl.branch.back();
        b.cls.allocate(lc.stream, nl);
        b.cv_a.allocate(lc.stream, (u64) nl + 1);
        b.cv_b.allocate(lc.stream, (u64) nl + 1);
        b.rec.allocate(lc.stream, ((u64) nl + 1) * kMaxRec);
        b.bitmap.allocate(lc.stream, words);
        b.word_rank.allocate(lc.stream, (u64) words + 1);
        b.counts.allocate(lc.stream, 2);
        b.table.allocate(lc.stream, (u64) words + 1);
        cub::CountingInputIterator<u32> iota(0);
        cub::TransformInputIterator<u32, PopcountOf, cub::CountingInputIterator<u32>> pops(iota, PopcountOf{b.bitmap.ptr, words});
        pl.temp_bytes = 0;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, pl.temp_bytes, pops, b.word_rank.ptr, (int) (words + 1), lc.stream));
        b.temp.allocate(lc.stream, pl.temp_bytes);
        if (pl.branch.size() > 1)
        {
            WHARF_CUDA(cudaStreamCreateWithFlags(&b.stream, cudaStreamNonBlocking));
            WHARF_CUDA(cudaEventCreateWithFlags(&b.done, cudaEventDisableTiming));
        }
    }
}

// One replica's share on this rank: all n walkers (a replica the rank owns entirely) or the walkers of the owned
// start vertices (a replica sharded over the ranks) - ShardMap, common.cuh.
struct ReplicaJob
{
    u32 replica;      // global replica number (selects the draws)
    u32 nl;           // walkers
    u64 first;        // first local walk = column offset in every corpus row
    u32 rank, world;  // start vertex of walker i: shard_start(i, rank, world, n)
    bool row1_done = false;   // corpus row 1 of this (whole) replica was written by coalesced_first_steps_kernel
};

std::vector<ReplicaJob> replica_jobs(const ShardMap& m)
{
    std::vector<ReplicaJob> jobs;
    for (u32 j = 0; j < m.whole; j++) jobs.push_back({m.whole_first + j, (u32) m.n, (u64) j * m.n, 0u, 1u});
    for (u32 k = 0; k < m.shared; k++) jobs.push_back({m.shared_first + k, m.n_local, m.whole_walks() + (u64) k * m.n_local, m.sub_rank, m.sub_world});
    return jobs;
}

// All segments of one replica job on the stream of `lc`, with the scratch of branch `b`.
void enqueue_replica(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps, const CoalescedPlan& pl,
                     const CoalescedBranch& b, const ReplicaJob& job)
{
    const u32 nl = job.nl, L = wp.L, words = pl.words;
    if (nl == 0) return;
    SegmentCtx s{lc, wp, words, nl, div_up(nl, kBlock), div_up(nl, kBlock * kEmitWalkers), grid_for(nl, kBlock, 8, lc.sm_count), div_up(words, kBlock),
                 div_up(words, kTileWords), b.cls.ptr, b.rec.ptr, b.bitmap.ptr, b.word_rank.ptr, b.table.ptr, b.temp.ptr, pl.temp_bytes, d_steps};
    u32 *cv = b.cv_a.ptr, *cv_next = b.cv_b.ptr, *nc = b.counts.ptr, *nc_next = b.counts.ptr + 1;
    u32* col = wp.walks + job.first;                          // column block of the replica in every corpus row
    coalesced_init_kernel<<<s.walker_grid, kBlock, 0, lc.stream>>>(b.cls.ptr, cv, col, nl, job.rank, job.world, wp.n, nc);
    lc.count();
    for (u32 a = 0; a < L;)
    {
        const u32 R = record_words(a);
        const u32 k = std::min(R - 1, L - a);                 // draws a .. a + k - 1
        SegDraws draws{};
        for (u32 j = 0; j < k; j++) draws.x[j] = host_draw[(u64) job.replica * L + a + j];
        const u32 stored = std::min(k, L - 1 - a);            // cells a + 1 .. a + stored exist; the last draw is only counted
        const bool merge = a + k < L;
        u32* first_row = col + (u64) (a + 1) * wp.w_local;
        const u32* first_step = (a == 0 && job.row1_done && k == 1 && stored == 1) ? first_row : nullptr;
        switch (R)
        {
            case 2: run_segment<2>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge, first_step); break;
            case 4: run_segment<4>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
            default: run_segment<8>(s, draws, k, cv, cv_next, nc, nc_next, first_row, stored, merge); break;
        }
        if (merge) { std::swap(cv, cv_next); std::swap(nc, nc_next); }
        a += k;
    }
    WHARF_CUDA(cudaGetLastError());
}

// Row 1 of every whole replica in one sweep over the graph (two or more of them, walks of at least two positions).
void enqueue_first_steps(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, std::vector<ReplicaJob>& jobs)
{
    const char* env = std::getenv("WHARF_FIRST_STEPS");
    if (wp.L < 2 || (env && std::atoi(env) == 0)) return;
    std::vector<ReplicaJob*> whole;
    for (ReplicaJob& job : jobs)
        if (job.world == 1 && job.nl == wp.n) whole.push_back(&job);
    if (whole.size() < 2) return;
    for (size_t g = 0; g < whole.size(); g += kFirstReplicas)
    {
        const u32 count = (u32) std::min<size_t>(kFirstReplicas, whole.size() - g);
        FirstDraws draws{};
        for (u32 r = 0; r < count; r++)
        {
            draws.x[r] = host_draw[(u64) whole[g + r]->replica * wp.L];
            if (whole[g + r]->first != whole[g]->first + (u64) r * wp.n) return;      // whole replicas lie side by side (ShardMap); anything else: no fusion
        }
        coalesced_first_steps_kernel<<<div_up(wp.n, kBlock), kBlock, 0, lc.stream>>>(wp.desc, wp.pool, (u32) wp.n, wp.shift, draws, count,
                                                                                  wp.walks + wp.w_local + whole[g]->first, wp.n);
        lc.count();
        for (u32 r = 0; r < count; r++) whole[g + r]->row1_done = true;
    }
    WHARF_CUDA(cudaGetLastError());
}

// One replica after the other on the caller's stream.
void enqueue_serial(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps, CoalescedPlan& pl,
                    std::vector<ReplicaJob>& jobs, const std::function<void(u64, u64)>& after_replica)
{
    const CoalescedBranch& b = pl.branch[0];
    WHARF_CUDA(cudaMemsetAsync(b.bitmap.ptr, 0, (u64) pl.words * 4, lc.stream));
    enqueue_first_steps(lc, wp, host_draw, jobs);
    for (const ReplicaJob& job : jobs)
    {
        enqueue_replica(lc, wp, host_draw, d_steps, pl, b, job);
        if (after_replica) after_replica(job.first, job.nl);
    }
}

// Inside a stream capture: job j on branch j % branches; the extra branches fork from and join the caller's stream.
void enqueue_forked(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps, CoalescedPlan& pl,
                    std::vector<ReplicaJob>& jobs, u32 branches)
{
    enqueue_first_steps(lc, wp, host_draw, jobs);
    WHARF_CUDA(cudaEventRecord(pl.fork, lc.stream));
    for (u32 i = 0; i < branches; i++)
    {
        const CoalescedBranch& b = pl.branch[i];
        LaunchCtx blc = lc;
        if (i) { blc.stream = b.stream; WHARF_CUDA(cudaStreamWaitEvent(b.stream, pl.fork, 0)); }
        WHARF_CUDA(cudaMemsetAsync(b.bitmap.ptr, 0, (u64) pl.words * 4, blc.stream));
        for (size_t j = i; j < jobs.size(); j += branches) enqueue_replica(blc, wp, host_draw, d_steps, pl, b, jobs[j]);
    }
    for (u32 i = 1; i < branches; i++)
    {
        WHARF_CUDA(cudaEventRecord(pl.branch[i].done, pl.branch[i].stream));
        WHARF_CUDA(cudaStreamWaitEvent(lc.stream, pl.branch[i].done, 0));
    }
}

bool graphs_enabled()
{
    const char* e = std::getenv("WHARF_GRAPH");
    return !e || std::atoi(e) != 0;
}
}  // namespace

void walks_generate_coalesced(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps,
                              const std::function<void(u64, u64)>& after_replica, CoalescedPlan** plan_slot)
{
    if (wp.w_local == 0) return;
    const u32 words = (u32) ((wp.n + 31) / 32);
    std::vector<ReplicaJob> jobs = replica_jobs(wp.map);
    u32 nl_max = 0;
    for (const ReplicaJob& job : jobs) nl_max = std::max(nl_max, job.nl);
    CoalescedPlan local;
    CoalescedPlan* pl = &local;
    if (plan_slot)
    {
        if (!*plan_slot) *plan_slot = new CoalescedPlan();
        pl = *plan_slot;
    }
    const bool as_graph = !after_replica && plan_slot && graphs_enabled();
    const u32 branches = as_graph ? wanted_branches(nl_max, (u32) jobs.size()) : 1;
    plan_buffers(*pl, lc, nl_max, words, branches);
    if (!as_graph)
    {
        enqueue_serial(lc, wp, host_draw, d_steps, *pl, jobs, after_replica);
        return;
    }
    u64 draw_hash = 1469598103934665603ull;
    for (u64 i = 0; i < (u64) wp.wpv * wp.L; i++) draw_hash = (draw_hash ^ host_draw[i]) * 1099511628211ull;
    const CoalescedPlan::Key key{wp.desc, wp.pool, wp.walks, d_steps, wp.n, wp.w_local, draw_hash, wp.map.n_local, wp.wpv, wp.L, wp.map.rank, wp.map.world,
                                 wp.map.whole * 65536u + wp.map.shared_first * 2u + wp.shift};
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
            enqueue_forked(cap, wp, host_draw, d_steps, *pl, jobs, branches);
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
