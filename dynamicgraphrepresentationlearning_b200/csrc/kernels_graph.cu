// Device adjacency: build, per-batch delta merge, compaction, export, synthetic RMAT input.
//
// Replaces graph/snapshot.h:108-194 (FlatGraph) + graph/wharfmh.h:175-208 (flatten_graph, re-done in full
// for every batch by the reference), the C-tree constructor / uniont / difference
// (libs/compressed_trees/graph/tree_plus/tree_plus.h:373,951,1319 via wharfmh.h:81,511,659),
// sort_edge_batch_by_source (wharfmh.h:1056-1104) and the dedup of wharfmh.h:456-470.
//
// Data layout: desc[v] = {begin, degree} (one aligned 8-byte record per vertex) into an append-only u32
// neighbour pool.  A batch rewrites only the rows of its source vertices: the merged row is appended at the
// pool tail and desc[v] is repointed; the old copy becomes garbage that graph_compact() reclaims when the
// pool fills.  Rows are always ascending and unique, which is all the walk path needs from the C-tree.
#include "engine.hpp"

#include <chrono>
#include <cstdio>
#include <cstdlib>

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace wharf
{
namespace
{
constexpr u32 kChunk = 1024;      // old-row elements copied by one warp in the merge
constexpr u32 kBigRow = 2048;     // rows above this are copied by a whole block in compaction / export

__global__ void build_desc_kernel(const u64* __restrict__ offsets, u64 n, u64 m, RowDesc* __restrict__ desc)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const u64 lo = offsets[v];
    const u64 hi = (v + 1 == n) ? m : offsets[v + 1];      // n offsets only: graph/wharfmh.h:74-75
    desc[v] = RowDesc{(u32) lo, (u32) (hi - lo)};
}

// The constructor's contract (graph/wharfmh.h:58-103: rows sorted and unique, ids < n): offsets monotone and <= m,
// every neighbour < n, every row strictly ascending.  One thread per edge slot finds its row by binary search over the
// n offsets; flag bit 0 = offsets, bit 1 = neighbour id, bit 2 = row order.
__global__ void __launch_bounds__(256) check_csr_kernel(const u64* __restrict__ offsets, u64 n, u64 m, const u32* __restrict__ edges, u32* __restrict__ flag)
{
    const u64 t = (u64) blockIdx.x * 256 + threadIdx.x;
    if (t < n)
    {
        const u64 lo = offsets[t];
        const u64 hi = (t + 1 == n) ? m : offsets[t + 1];
        if (lo > hi || hi > m || (t == 0 && lo != 0)) atomicOr(flag, 1u);
    }
    if (t >= m) return;
    const u32 x = edges[t];
    if (x >= n) atomicOr(flag, 2u);
    if (t == 0) return;
    // same row as the previous slot <=> no offset equals t: largest v with offsets[v] <= t
    u64 lo = 0, hi = n;
    while (hi - lo > 1)
    {
        const u64 mid = (lo + hi) >> 1;
        if (offsets[mid] <= t) lo = mid; else hi = mid;
    }
    if (offsets[lo] < t && edges[t - 1] >= x) atomicOr(flag, 4u);
}

__global__ void check_ids_kernel(const u32* __restrict__ pairs, u64 count2, u64 n, u32* __restrict__ flag)
{
    for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < count2; i += (u64) gridDim.x * blockDim.x)
        if (pairs[i] >= n) *flag = 1;
}

__global__ void pairs_to_keys_kernel(const u32* __restrict__ pairs, u64 count, u64* __restrict__ keys)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const uint2 e = reinterpret_cast<const uint2*>(pairs)[i];
    keys[i] = ((u64) e.x << 32) | e.y;                      // (src << bits) + dst: wharfmh.h:1083-1085
}

__global__ void keys_to_pairs_kernel(const u64* __restrict__ keys, u64 count, u32* __restrict__ pairs)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const u64 k = keys[i];
    reinterpret_cast<uint2*>(pairs)[i] = make_uint2((u32) (k >> 32), (u32) k);
}

// keep[i]: first of a run of equal keys and (unless allowed) not a self-loop: wharfmh.h:459-463, utility.h:121-125
__global__ void keep_flags_kernel(const u64* __restrict__ keys, u64 count, bool drop_self_loops, u8* __restrict__ keep)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const u64 k = keys[i];
    bool f = (i == 0) || keys[i - 1] != k;
    if (drop_self_loops && (u32) (k >> 32) == (u32) k) f = false;
    keep[i] = f;
}

// head[i]: first key of its source (run starts, wharfmh.h:475-481)
__global__ void head_flags_kernel(const u64* __restrict__ keys, u64 count, u8* __restrict__ head)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    head[i] = (i == 0) || (u32) (keys[i - 1] >> 32) != (u32) (keys[i] >> 32);
}

__global__ void sources_kernel(const u64* __restrict__ keys, u64* __restrict__ slice, u64 n_sources, u64 k, u32* __restrict__ sources)
{
    const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (j < n_sources) sources[j] = (u32) (keys[slice[j]] >> 32);
    if (j == 0) slice[n_sources] = k;
}

// Per batch edge: rank of dst in the current row of src and whether it is already there.
// change[i] = 1 when the edge changes the row (insert: absent, delete: present).
__global__ void locate_kernel(const u64* __restrict__ keys, u64 k, const RowDesc* __restrict__ desc, const u32* __restrict__ pool,
                              u32 rs, bool is_delete, u32* __restrict__ rank, u32* __restrict__ change)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i > k) return;
    if (i == k) { change[k] = 0; return; }                   // sentinel so the exclusive scan yields the total
    const u64 key = keys[i];
    const u32 src = (u32) (key >> 32), dst = (u32) key;
    const uint2 t = reinterpret_cast<const uint2*>(desc)[src];
    const u32 r = lower_bound_row(pool + t.x, t.y, dst, rs);
    const bool present = r < t.y && pool[t.x + ((u64) r << rs)] == dst;
    rank[i] = r;
    change[i] = is_delete ? present : !present;
}

// Per source: new degree, merge chunks of the old row.
__global__ void plan_rows_kernel(const u32* __restrict__ sources, const u64* __restrict__ slice, u64 n_sources,
                                 const u64* __restrict__ change_scan, const RowDesc* __restrict__ desc, bool is_delete, u32 row_words,
                                 u64* __restrict__ new_degree, u64* __restrict__ alloc, u64* __restrict__ chunks)
{
    const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (j > n_sources) return;
    if (j == n_sources) { new_degree[j] = 0; alloc[j] = 0; chunks[j] = 0; return; }
    const u32 old_deg = desc[sources[j]].degree;
    const u64 delta = change_scan[slice[j + 1]] - change_scan[slice[j]];
    new_degree[j] = is_delete ? old_deg - delta : old_deg + delta;
    alloc[j] = new_degree[j] * row_words;                       // neighbours (+ the aux tail of the same length)
    chunks[j] = (old_deg + kChunk - 1) / kChunk;
}

// One warp per chunk of an old row: sorted-set union with / difference by the batch slice.
// Position of old element a (row index ia) in the new row:
//   insert: ia + #(new batch dsts < a)     delete: ia - #(removed dsts < a), dropped when a is removed.
__global__ void __launch_bounds__(256) merge_old_rows_kernel(const u32* __restrict__ sources, const u64* __restrict__ slice, u64 n_sources,
                                                            const u64* __restrict__ keys, const u64* __restrict__ change_scan,
                                                            const u32* __restrict__ change, const u64* __restrict__ chunk_scan, u64 n_chunks,
                                                            const u64* __restrict__ alloc_scan, u64 pool_base,
                                                            const RowDesc* __restrict__ desc, u32* __restrict__ pool, u32 rs, bool is_delete)
{
    const u64 chunk = ((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 lane = threadIdx.x & 31;
    if (chunk >= n_chunks) return;
    const u32 j = upper_segment_u64(chunk_scan, (u32) n_sources + 1, chunk);
    const u32 c = (u32) (chunk - chunk_scan[j]);
    const uint2 d = reinterpret_cast<const uint2*>(desc)[sources[j]];
    const u64 b_lo = slice[j];
    const u32 b_n = (u32) (slice[j + 1] - b_lo);
    const u64 scan_lo = change_scan[b_lo];
    u32* out = pool + pool_base + alloc_scan[j];
    const u32 hi = min(d.y, (c + 1) * kChunk);
    for (u32 ia = c * kChunk + lane; ia < hi; ia += 32)
    {
        const u32 a = pool[d.x + ((u64) ia << rs)];
        // lower_bound of a among the slice's dsts
        u32 lo = 0, up = b_n;
        while (lo < up)
        {
            const u32 mid = (lo + up) >> 1;
            if ((u32) keys[b_lo + mid] < a) lo = mid + 1; else up = mid;
        }
        const u32 shift = (u32) (change_scan[b_lo + lo] - scan_lo);
        u64 to;
        if (is_delete)
        {
            if (lo < b_n && (u32) keys[b_lo + lo] == a && change[b_lo + lo]) continue;     // removed
            to = (u64) (ia - shift) << rs;
        }
        else to = (u64) (ia + shift) << rs;
        out[to] = a;
        // aux word of a surviving slot: the filter half moves with its neighbour (triangle blocks are kept current
        // incrementally; per-row blocks are rebuilt after the merge anyway), the sampler cache half starts empty
        if (rs) out[to + 1] = pool[d.x + 2 * (u64) ia + 1] & 0xFFFF0000u;
    }
}

// Insert only: one thread per batch edge that is new to its row.
__global__ void write_new_edges_kernel(const u64* __restrict__ keys, u64 k, const u64* __restrict__ slice, u64 n_sources,
                                       const u32* __restrict__ rank, const u32* __restrict__ change, const u64* __restrict__ change_scan,
                                       const u64* __restrict__ alloc_scan, u64 pool_base, u32* __restrict__ pool, u32 rs)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k || !change[i]) return;
    const u32 j = upper_segment_u64(slice, (u32) n_sources + 1, i);
    const u64 within = change_scan[i] - change_scan[slice[j]];
    const u64 to = pool_base + alloc_scan[j] + ((rank[i] + within) << rs);
    pool[to] = (u32) keys[i];
    if (rs) pool[to + 1] = 0;                                     // aux word of a new slot: empty cache, empty filter block
}

// Repoint the rows, bump the sampler-manager generation (wharfmh.h:504 -> :539) and mark the touched bitmap.
__global__ void commit_rows_kernel(const u32* __restrict__ sources, u64 n_sources, const u64* __restrict__ new_degree,
                                   const u64* __restrict__ alloc_scan, u64 pool_base, RowDesc* __restrict__ desc,
                                   u32* __restrict__ version, u32* __restrict__ bitmap)
{
    const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_sources) return;
    const u32 v = sources[j];
    desc[v] = RowDesc{(u32) (pool_base + alloc_scan[j]), (u32) new_degree[j]};
    version[v] += 1;
    if (bitmap) atomicOr(bitmap + (v >> 5), 1u << (v & 31));
}

// ---- aux words (MH node2vec) ----------------------------------------------------------------------------
// With g.tail a row of degree d owns 2 d pool words, neighbours INTERLEAVED with their aux words:
//     pool[begin + 2 i]      neighbour i (ascending)
//     pool[begin + 2 i + 1]  aux word i:  low 16 bits   sampler cache of the state (cur = neighbour i, prev = this row's
//                                                       vertex)                                      (kernels_walk.cu)
//                                         high 16 bits  block i of the row's neighbour filter           (common.cuh)
// A walker that picks neighbour i of its vertex as the next vertex loads both words with one 8-byte load: the second
// is the sampler cache of the state it is about to enter, so the cache costs no memory access of its own (it used to
// be a separate sector, and on graphs beyond the L2 a separate DRAM fetch, per step).  Rows without aux words
// (g.tail == 0: deterministic engines) keep plain arrays; every row accessor takes the shift.

// One warp per row: zero the aux words, then set the filter bits of every neighbour.
__device__ __forceinline__ void aux_build_row(u32* __restrict__ pool, u32 begin, u32 degree, u32 lane)
{
    u32* row = pool + begin;
    for (u32 i = lane; i < degree; i += 32) row[2 * (u64) i + 1] = 0;
    __syncwarp();
    for (u32 i = lane; i < degree; i += 32)
    {
        const u32 h = filter_hash(row[2 * (u64) i]);
        atomicOr(row + 2 * (u64) filter_block(h, degree) + 1, filter_bits(h) << 16);
    }
}

__global__ void __launch_bounds__(256) aux_build_all_kernel(const RowDesc* __restrict__ desc, u64 n, u32* __restrict__ pool)
{
    const u64 v = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (v >= n) return;
    const uint2 d = reinterpret_cast<const uint2*>(desc)[v];
    aux_build_row(pool, d.x, d.y, threadIdx.x & 31);
}

__global__ void __launch_bounds__(256) aux_build_sources_kernel(const u32* __restrict__ sources, u64 n_sources, const RowDesc* __restrict__ desc,
                                                                u32* __restrict__ pool)
{
    const u64 j = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (j >= n_sources) return;
    const uint2 d = reinterpret_cast<const uint2*>(desc)[sources[j]];
    aux_build_row(pool, d.x, d.y, threadIdx.x & 31);
}

// The cache half of aux word i of row p is a function of the rows of cur = neighbour i and of p and of version[cur].
// Rows re-written by a batch get fresh (empty) aux words; this pass empties every other cache whose cur is a batch
// source (bit set in the touched bitmap).  One warp per row over all rows: no reverse adjacency needed, so it is
// exact for directed graphs.  bitmap == nullptr empties every cache (a fresh corpus, wharfmh.h:41,83,90).
__global__ void __launch_bounds__(256) aux_invalidate_kernel(const RowDesc* __restrict__ desc, u64 n, u32* __restrict__ pool, const u32* __restrict__ bitmap)
{
    const u64 v = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (v >= n) return;
    const uint2 d = reinterpret_cast<const uint2*>(desc)[v];
    unsigned short* halves = reinterpret_cast<unsigned short*>(pool + d.x);           // low half of aux word i = halves[4 i + 2]
    for (u32 i = threadIdx.x & 31; i < d.y; i += 32)
    {
        bool clear = true;
        if (bitmap)
        {
            const u32 c = pool[d.x + 2 * (u64) i];
            clear = (__ldg(bitmap + (c >> 5)) >> (c & 31)) & 1u;
        }
        if (clear) halves[4 * (u64) i + 2] = 0;
    }
}

// ---- triangle blocks (MH node2vec on a symmetric graph) -----------------------------------------------------
// Filter half of aux word i of row v, c = neighbour i:  OR over w in N(v) n N(c) of filter_bits(filter_hash(w)).
// node2vec asks "is the proposal c a neighbour of prev" for a walker standing on v that came from prev (node2vec.h:
// 74-88,112-119).  prev is a neighbour of v, and on a symmetric graph c in N(prev) <=> prev in N(c) <=> prev in N(v) n N(c):
// the block of the slot v -> c, which arrives with the load that picks c, answers it (94 % of the proposals are no
// neighbour of prev and the sets are nearly empty: mean 0.28 members at scale-20).  A block is a Bloom filter: extra bits
// only turn a proof of absence into a "maybe", which the walk kernel resolves with the exact search.
constexpr u32 kTriChunk = 32;        // slots of one row handled by one warp (four 8-lane groups x 8 slots: hub rows spread over many warps)
constexpr u32 kTriHeavy = 2048;      // slots whose smaller row is longer than this go to a block of their own

// Scratch of a build: one 16-bit block per edge slot, row v owning blocks[eoff[v], eoff[v] + degree) - the per-row
// neighbour filter of common.cuh (element x sets filter_bits(h(x)) in block filter_block(h(x), degree)).  A membership
// question is answered by ONE 2-byte load in most cases (a clear bit proves absence); the binary search runs for the rest.
struct TriFilter
{
    const unsigned short* blocks;
    const u32* eoff;
};

__device__ __forceinline__ bool tri_member(const u32* __restrict__ pool, uint2 row, u32 row_off, u32 x, const TriFilter& f)
{
    const u32 h = filter_hash(x);
    const u32 need = filter_bits(h);
    if ((f.blocks[row_off + filter_block(h, row.y)] & need) != need) return false;
    const u32 r = lower_bound_row(pool + row.x, row.y, x, 1);
    return r < row.y && __ldg(pool + row.x + 2 * (u64) r) == x;
}

__device__ __forceinline__ void tri_store(u32* __restrict__ pool, uint2 dv, u32 i, u32 bits)
{
    reinterpret_cast<unsigned short*>(pool + dv.x + 2 * (u64) i + 1)[1] = (unsigned short) bits;      // high half only
}

struct TriChunksOf
{
    const RowDesc* desc;
    u64 count;
    __device__ u64 operator()(u64 v) const { return v < count ? (desc[v].degree + kTriChunk - 1) / kTriChunk : 0; }
};

struct SlotsOfRow
{
    const RowDesc* desc;
    u64 count;
    __device__ u64 operator()(u64 v) const { return v < count ? (u64) desc[v].degree : 0ull; }
};

struct DegreeOfRow
{
    const RowDesc* desc;
    u64 count;
    __device__ u32 operator()(u32 v) const { return v < count ? desc[v].degree : 0u; }
};

// One warp per chunk of kTriChunk elements of a row: the row's per-row filter blocks into the scratch array (zeroed before).
__global__ void __launch_bounds__(256) tri_rowfilter_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, u64 n,
                                                            const u64* __restrict__ chunk_scan, u64 n_chunks, const u32* __restrict__ eoff,
                                                            u32* __restrict__ blocks_as_words)
{
    const u64 chunk = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (chunk >= n_chunks) return;
    const u32 lane = threadIdx.x & 31;
    const u32 v = upper_segment_u64(chunk_scan, (u32) n + 1, chunk);
    const uint2 dv = reinterpret_cast<const uint2*>(desc)[v];
    const u32 first = (u32) (chunk - chunk_scan[v]) * kTriChunk, last = min(dv.y, first + kTriChunk);
    const u32 off = eoff[v];
    for (u32 i = first + lane; i < last; i += 32)
    {
        const u32 h = filter_hash(__ldg(pool + dv.x + 2 * (u64) i));
        const u32 slot = off + filter_block(h, dv.y);                         // 16-bit block `slot`: half of word slot / 2
        atomicOr(blocks_as_words + (slot >> 1), filter_bits(h) << ((slot & 1) * 16));
    }
}

// One warp per chunk of kTriChunk slots of a row u, an 8-lane group per slot u -> w.  A pair {u, w} is handled ONCE, from the
// row of its longer member (ties: the larger id): the lanes scan N(w) - short and contiguous - and look every element up in
// the filter and the row of u, which all slots of the chunk share (L1 / L2-hot; looking the members of the short row up
// from the short row's side sent one random DRAM fetch per lookup: 8.6 ms at scale-20).  The block goes to slot u -> w and
// to its twin w -> u, whose position falls out of the scan.  counters[1] |= 1 when a twin is missing, counters[2] counts the
// slots written: a graph is symmetric exactly when no twin is missing and every slot was written.
__global__ void __launch_bounds__(256) tri_build_kernel(const RowDesc* __restrict__ desc, u32* __restrict__ pool, u64 n,
                                                        const u64* __restrict__ chunk_scan, u64 n_chunks, TriFilter f,
                                                        uint2* __restrict__ heavy, u32 heavy_capacity, u32* __restrict__ counters)
{
    const u64 chunk = ((u64) blockIdx.x * 256 + threadIdx.x) >> 5;
    if (chunk >= n_chunks) return;
    const u32 lane = threadIdx.x & 31, sub = lane & 7, group = lane >> 3;
    const u32 u = upper_segment_u64(chunk_scan, (u32) n + 1, chunk);
    const uint2 du = reinterpret_cast<const uint2*>(desc)[u];
    const u32 off_u = f.eoff[u];
    const u32 first = (u32) (chunk - chunk_scan[u]) * kTriChunk;
    const u32 last = min(du.y, first + kTriChunk);
    u32 written = 0;
    for (u32 i0 = first; i0 < last; i0 += 4)                      // uniform trip count: the shuffles below need the whole warp
    {
        const u32 i = i0 + group;
        const bool has = i < last;
        u32 w = 0;
        uint2 dw = make_uint2(0, 0);
        if (has)
        {
            w = __ldg(pool + du.x + 2 * (u64) i);
            dw = reinterpret_cast<const uint2*>(desc)[w];
        }
        const bool mine = has && (dw.y < du.y || (dw.y == du.y && w <= u));
        const bool is_heavy = mine && dw.y > kTriHeavy;
        u32 bits = 0, twin = 0xFFFFFFFFu;
        if (mine && !is_heavy)
            for (u32 k0 = sub; k0 < dw.y; k0 += 32)               // four elements per lane in flight: the chain x -> filter block is latency
            {
                u32 x[4], blk[4];
#pragma unroll
                for (u32 j = 0; j < 4; j++) x[j] = k0 + 8 * j < dw.y ? __ldg(pool + dw.x + 2 * (u64) (k0 + 8 * j)) : 0xFFFFFFFFu;
#pragma unroll
                for (u32 j = 0; j < 4; j++)
                    blk[j] = x[j] != 0xFFFFFFFFu ? (u32) f.blocks[off_u + filter_block(filter_hash(x[j]), du.y)] : 0u;
#pragma unroll
                for (u32 j = 0; j < 4; j++)
                {
                    if (x[j] == u) twin = k0 + 8 * j;
                    const u32 need = filter_bits(filter_hash(x[j]));
                    if (x[j] != 0xFFFFFFFFu && (blk[j] & need) == need)
                    {
                        const u32 r = lower_bound_row(pool + du.x, du.y, x[j], 1);
                        if (r < du.y && __ldg(pool + du.x + 2 * (u64) r) == x[j]) bits |= need;
                    }
                }
            }
#pragma unroll
        for (u32 d = 1; d < 8; d <<= 1)
        {
            bits |= __shfl_xor_sync(0xffffffffu, bits, d);
            twin = min(twin, __shfl_xor_sync(0xffffffffu, twin, d));
        }
        if (mine && sub == 0)
        {
            if (is_heavy)
            {
                const u32 at = atomicAdd(counters, 1u);
                if (at < heavy_capacity) heavy[at] = make_uint2(u, i);
                else                                                  // list full: "maybe" for every query, the exact search decides
                {
                    const u32 r = lower_bound_row(pool + dw.x, dw.y, u, 1);
                    if (r < dw.y && __ldg(pool + dw.x + 2 * (u64) r) == u)
                    {
                        tri_store(pool, du, i, 0xFFFFu);
                        if (w != u) tri_store(pool, dw, r, 0xFFFFu);
                        written += (w != u) ? 2 : 1;
                    }
                    else counters[1] = 1u;
                }
            }
            else if (twin == 0xFFFFFFFFu) counters[1] = 1u;
            else
            {
                tri_store(pool, du, i, bits);
                if (w != u) tri_store(pool, dw, twin, bits);
                written += (w != u) ? 2 : 1;
            }
        }
    }
#pragma unroll
    for (u32 d = 16; d; d >>= 1) written += __shfl_xor_sync(0xffffffffu, written, d);
    if (lane == 0 && written) atomicAdd(counters + 2, written);
}

// One block per heavy pair (both rows long: hub - hub edges).
__global__ void __launch_bounds__(256) tri_heavy_kernel(const RowDesc* __restrict__ desc, u32* __restrict__ pool, TriFilter f,
                                                        const uint2* __restrict__ heavy, u32 heavy_capacity, u32* __restrict__ counters)
{
    __shared__ u32 s_bits, s_twin;
    const u32 count = min(counters[0], heavy_capacity);
    for (u32 h = blockIdx.x; h < count; h += gridDim.x)
    {
        const uint2 item = heavy[h];
        const u32 u = item.x;
        const uint2 du = reinterpret_cast<const uint2*>(desc)[u];
        const u32 off_u = f.eoff[u];
        const u32 w = pool[du.x + 2 * (u64) item.y];
        const uint2 dw = reinterpret_cast<const uint2*>(desc)[w];
        if (threadIdx.x == 0) { s_bits = 0; s_twin = 0xFFFFFFFFu; }
        __syncthreads();
        u32 bits = 0;
        for (u32 k = threadIdx.x; k < dw.y; k += 256)
        {
            const u32 x = __ldg(pool + dw.x + 2 * (u64) k);
            if (x == u) s_twin = k;
            if (tri_member(pool, du, off_u, x, f)) bits |= filter_bits(filter_hash(x));
        }
        if (bits) atomicOr(&s_bits, bits);
        __syncthreads();
        if (threadIdx.x == 0)
        {
            if (s_twin == 0xFFFFFFFFu) counters[1] = 1u;
            else
            {
                tri_store(pool, du, item.y, s_bits);
                if (w != u) tri_store(pool, dw, s_twin, s_bits);
                atomicAdd(counters + 2, (w != u) ? 2u : 1u);
            }
        }
        __syncthreads();
    }
}

// Insert batch, after the merge: one 8-lane group per NEW undirected edge {a, b} - the key (a, b) with a <= b; the batch is
// symmetric, so its twin key is there too and is skipped: the intersection is the same set.  With I = N(a) n N(b) in the
// NEW graph: the slots a -> b and b -> a get the block of I; b joined N(a) and a joined N(b), so for every w in I the sets
// of a -> w, w -> a gained b and those of b -> w, w -> b gained a.  ORs are idempotent, so edges of one batch that close
// triangles with each other need no ordering.  Hub - hub edges go to the heavy list.
__device__ __forceinline__ void tri_or(u32* __restrict__ pool, uint2 row, u32 at, u32 bits_hi)
{
    atomicOr(pool + row.x + 2 * (u64) at + 1, bits_hi);
}

__device__ __forceinline__ void tri_or_found(u32* __restrict__ pool, uint2 row, u32 x, u32 bits_hi)       // slot row -> x, if there
{
    const u32 r = lower_bound_row(pool + row.x, row.y, x, 1);
    if (r < row.y && pool[row.x + 2 * (u64) r] == x) tri_or(pool, row, r, bits_hi);
}

// The members of I found by the lanes `sub` of `lanes`: ORs into the four slots each touches; returns the block bits.
__device__ __forceinline__ u32 tri_insert_pair(const RowDesc* __restrict__ desc, u32* __restrict__ pool, u32 a, uint2 da, u32 b, uint2 db, u32 sub, u32 lanes)
{
    const bool a_small = da.y <= db.y;
    const uint2 small = a_small ? da : db, large = a_small ? db : da;
    const u32 bits_a = filter_bits(filter_hash(a)) << 16, bits_b = filter_bits(filter_hash(b)) << 16;
    u32 bits = 0;
    for (u32 q = sub; q < small.y; q += lanes)
    {
        const u32 w = __ldg(pool + small.x + 2 * (u64) q);
        const u32 r = lower_bound_row(pool + large.x, large.y, w, 1);
        if (r >= large.y || __ldg(pool + large.x + 2 * (u64) r) != w) continue;
        bits |= filter_bits(filter_hash(w));
        tri_or(pool, da, a_small ? q : r, bits_b);                                                   // a -> w gained b
        if (a != b) tri_or(pool, db, a_small ? r : q, bits_a);                                       // b -> w gained a
        const uint2 dw = reinterpret_cast<const uint2*>(desc)[w];
        tri_or_found(pool, dw, a, bits_b);                                                           // w -> a gained b
        if (a != b) tri_or_found(pool, dw, b, bits_a);                                               // w -> b gained a
    }
    return bits;
}

__global__ void __launch_bounds__(256) tri_insert_kernel(const u64* __restrict__ keys, u64 k, const u32* __restrict__ change,
                                                         const RowDesc* __restrict__ desc, u32* __restrict__ pool,
                                                         uint2* __restrict__ heavy, u32* __restrict__ heavy_count, u32 heavy_capacity)
{
    const u64 item = ((u64) blockIdx.x * 256 + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    bool has = item < k && change[item];
    u32 a = 0, b = 0;
    uint2 da = make_uint2(0, 0), db = make_uint2(0, 0);
    if (has)
    {
        a = (u32) (keys[item] >> 32);
        b = (u32) keys[item];
        has = a <= b;                                                                                  // the twin key does nothing
    }
    if (has)
    {
        da = reinterpret_cast<const uint2*>(desc)[a];
        db = reinterpret_cast<const uint2*>(desc)[b];
    }
    const bool is_heavy = has && min(da.y, db.y) > kTriHeavy;
    u32 bits = (has && !is_heavy) ? tri_insert_pair(desc, pool, a, da, b, db, sub, 8) : 0u;
    bits |= __shfl_xor_sync(0xffffffffu, bits, 1);
    bits |= __shfl_xor_sync(0xffffffffu, bits, 2);
    bits |= __shfl_xor_sync(0xffffffffu, bits, 4);
    if (has && sub == 0)
    {
        if (is_heavy)
        {
            const u32 at = atomicAdd(heavy_count, 1u);
            if (at < heavy_capacity) heavy[at] = make_uint2(a, b);
            else                                                                                       // list full: both slots answer "maybe"
            {
                tri_or_found(pool, da, b, 0xFFFF0000u);
                tri_or_found(pool, db, a, 0xFFFF0000u);
            }
        }
        else if (bits)
        {
            tri_or_found(pool, da, b, bits << 16);                                                     // slot a -> b
            if (a != b) tri_or_found(pool, db, a, bits << 16);                                         // slot b -> a
        }
    }
}

// Heavy inserted edges {a, b}: a block per edge.
__global__ void __launch_bounds__(256) tri_insert_heavy_kernel(const RowDesc* __restrict__ desc, u32* __restrict__ pool, const uint2* __restrict__ heavy,
                                                               const u32* __restrict__ heavy_count, u32 heavy_capacity)
{
    __shared__ u32 s_bits;
    const u32 count = min(*heavy_count, heavy_capacity);
    for (u32 h = blockIdx.x; h < count; h += gridDim.x)
    {
        const u32 a = heavy[h].x, b = heavy[h].y;
        const uint2 da = reinterpret_cast<const uint2*>(desc)[a];
        const uint2 db = reinterpret_cast<const uint2*>(desc)[b];
        if (threadIdx.x == 0) s_bits = 0;
        __syncthreads();
        const u32 bits = tri_insert_pair(desc, pool, a, da, b, db, threadIdx.x, 256);
        if (bits) atomicOr(&s_bits, bits);
        __syncthreads();
        if (threadIdx.x == 0 && s_bits)
        {
            tri_or_found(pool, da, b, s_bits << 16);
            if (a != b) tri_or_found(pool, db, a, s_bits << 16);
        }
        __syncthreads();
    }
}

// Self-check of the membership filter (wmh_filter_check): a filter may say "maybe" too often, never "no" wrongly.  Triangle
// mode: the block of slot v -> c must carry the bits of every w in N(v) n N(c); per-row mode: block filter_block(h(x), deg)
// of row v must carry the bits of every neighbour x.  One 8-lane group per slot; counts the slots that miss a bit.
__global__ void __launch_bounds__(256) filter_check_kernel(const RowDesc* __restrict__ desc, const u32* __restrict__ pool, u64 n,
                                                           const u64* __restrict__ slot_scan, u64 slots, u32 tri, unsigned long long* __restrict__ missing)
{
    const u64 slot = ((u64) blockIdx.x * 256 + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    if (slot >= slots) return;
    const u32 group_mask = __activemask();                       // whole 8-lane groups leave together: a lane's partners below are active
    const u32 v = upper_segment_u64(slot_scan, (u32) n + 1, slot);
    const uint2 dv = reinterpret_cast<const uint2*>(desc)[v];
    const u32 i = (u32) (slot - slot_scan[v]);
    bool bad = false;
    if (tri)
    {
        const u32 c = pool[dv.x + 2 * (u64) i];
        const u32 block = pool[dv.x + 2 * (u64) i + 1] >> 16;
        const uint2 dc = reinterpret_cast<const uint2*>(desc)[c];
        for (u32 q = sub; q < dv.y; q += 8)
        {
            const u32 w = pool[dv.x + 2 * (u64) q];
            const u32 r = lower_bound_row(pool + dc.x, dc.y, w, 1);
            if (r >= dc.y || pool[dc.x + 2 * (u64) r] != w) continue;
            const u32 need = filter_bits(filter_hash(w));
            bad |= (block & need) != need;
        }
    }
    else if (sub == 0)
    {
        const u32 h = filter_hash(pool[dv.x + 2 * (u64) i]);
        const u32 need = filter_bits(h);
        bad = ((pool[dv.x + 2 * (u64) filter_block(h, dv.y) + 1] >> 16) & need) != need;
    }
    u32 any = bad ? 1u : 0u;
    any |= __shfl_xor_sync(group_mask, any, 1);
    any |= __shfl_xor_sync(group_mask, any, 2);
    any |= __shfl_xor_sync(group_mask, any, 4);
    if (any && sub == 0) atomicAdd(missing, 1ull);               // one count per slot
}

// flag |= 1 when some batch key (s, d) has no twin (d, s) in the same (sorted, unique) batch.
__global__ void symmetric_batch_kernel(const u64* __restrict__ keys, u64 k, u32* __restrict__ flag)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k) return;
    const u64 key = keys[i];
    const u64 twin = (key << 32) | (key >> 32);
    u64 lo = 0, hi = k;
    while (lo < hi)
    {
        const u64 mid = (lo + hi) >> 1;
        if (keys[mid] < twin) lo = mid + 1; else hi = mid;
    }
    if (lo >= k || keys[lo] != twin) *flag = 1;
}

// ---- row copy (compaction / export) ---------------------------------------------------------------
__global__ void degrees_kernel(const RowDesc* __restrict__ desc, u64 n, u32 row_words, u64* __restrict__ deg)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n) return;
    deg[v] = v < n ? (u64) desc[v].degree * row_words : 0;
}

// 8 lanes per row; rows above kBigRow are left to copy_big_rows_kernel.
__global__ void __launch_bounds__(256) copy_rows_kernel(const RowDesc* __restrict__ desc, u64 n, const u32* __restrict__ pool,
                                                       const u64* __restrict__ dst_offset, u32* __restrict__ dst, u32 ss, u32 ds)
{
    const u64 g = ((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 3;
    const u32 sub = threadIdx.x & 7;
    if (g >= n) return;
    const uint2 d = reinterpret_cast<const uint2*>(desc)[g];
    if (d.y > kBigRow) return;
    const u64 o = dst_offset[g];
    for (u32 i = sub; i < d.y; i += 8)
    {
        dst[o + ((u64) i << ds)] = pool[d.x + ((u64) i << ss)];
        if (ds) dst[o + 2 * (u64) i + 1] = ss ? pool[d.x + 2 * (u64) i + 1] & 0xFFFF0000u : 0u;    // filter half travels, cache half empties
    }
}

struct IsBigRow
{
    const RowDesc* desc;
    __device__ bool operator()(u32 v) const { return desc[v].degree > kBigRow; }
};

__global__ void __launch_bounds__(256) copy_big_rows_kernel(const u32* __restrict__ rows, const RowDesc* __restrict__ desc,
                                                           const u32* __restrict__ pool, const u64* __restrict__ dst_offset,
                                                           u32* __restrict__ dst, u32 ss, u32 ds)
{
    const u32 v = rows[blockIdx.x];
    const uint2 d = reinterpret_cast<const uint2*>(desc)[v];
    const u64 o = dst_offset[v];
    for (u32 i = threadIdx.x; i < d.y; i += 256)
    {
        dst[o + ((u64) i << ds)] = pool[d.x + ((u64) i << ss)];
        if (ds) dst[o + 2 * (u64) i + 1] = ss ? pool[d.x + 2 * (u64) i + 1] & 0xFFFF0000u : 0u;
    }
}

__global__ void repoint_kernel(RowDesc* __restrict__ desc, u64 n, const u64* __restrict__ new_begin)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) desc[v].begin = (u32) new_begin[v];
}

// Copies the neighbours of every row: element i from pool[begin + (i << ss)] to dst[offset + (i << ds)].
void copy_all_rows(const LaunchCtx& lc, const RowDesc* desc, u64 n, const u32* pool, const u64* d_dst_offset, u32* d_dst, u32 ss, u32 ds)
{
    if (n == 0) return;
    copy_rows_kernel<<<div_up(n * 8, 256), 256, 0, lc.stream>>>(desc, n, pool, d_dst_offset, d_dst, ss, ds);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    Scratch<u32> big(lc.stream, n);
    Scratch<u64> d_count(lc.stream, 1);
    cub::CountingInputIterator<u32> iota(0);
    IsBigRow pred{desc};
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceSelect::If(nullptr, temp_bytes, iota, big.ptr, d_count.ptr, (i64) n, pred, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceSelect::If(temp.ptr, temp_bytes, iota, big.ptr, d_count.ptr, (i64) n, pred, lc.stream));
    lc.count_lib();
    u64 n_big = 0;
    WHARF_CUDA(cudaMemcpyAsync(&n_big, d_count.ptr, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    if (n_big)
    {
        copy_big_rows_kernel<<<(unsigned) n_big, 256, 0, lc.stream>>>(big.ptr, desc, pool, d_dst_offset, d_dst, ss, ds);
        lc.count();
        WHARF_CUDA(cudaGetLastError());
    }
}

template <class T>
void exclusive_sum(const LaunchCtx& lc, const T* in, u64* out, u64 count)
{
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, in, out, (i64) count, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, temp_bytes, in, out, (i64) count, lc.stream));
    lc.count_lib();
}

void sort_keys(const LaunchCtx& lc, u64* keys_in, u64* keys_out, u64 count, int end_bit)
{
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, temp_bytes, keys_in, keys_out, (i64) count, 0, end_bit, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceRadixSort::SortKeys(temp.ptr, temp_bytes, keys_in, keys_out, (i64) count, 0, end_bit, lc.stream));
    lc.count_lib();
}

int bits_for(u64 n)
{
    int b = 1;
    while (b < 32 && (1ull << b) < n) b++;
    return b;
}

// Selects keys[i] with flag[i] into out; returns the number selected (host sync).
u64 select_flagged(const LaunchCtx& lc, const u64* keys, const u8* flags, u64 count, u64* out)
{
    Scratch<u64> d_count(lc.stream, 1);
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceSelect::Flagged(nullptr, temp_bytes, keys, flags, out, d_count.ptr, (i64) count, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceSelect::Flagged(temp.ptr, temp_bytes, keys, flags, out, d_count.ptr, (i64) count, lc.stream));
    lc.count_lib();
    u64 k = 0;
    WHARF_CUDA(cudaMemcpyAsync(&k, d_count.ptr, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    return k;
}

// ---- RMAT (libs/compressed_trees/rmat_util.h:7-43, pbbslib/utilities.h:108-147) ---------------------
__host__ __device__ __forceinline__ u32 hash32(u32 a)
{
    a = (a + 0x7ed55d16) + (a << 12);
    a = (a ^ 0xc761c23c) ^ (a >> 19);
    a = (a + 0x165667b1) + (a << 5);
    a = (a + 0xd3a2646c) ^ (a << 9);
    a = (a + 0xfd7046c5) + (a << 3);
    a = (a ^ 0xb55a4f09) ^ (a >> 16);
    return a;
}

inline u64 hash64(u64 u)
{
    u64 v = u * 3935559000370003845ull + 2691343689449507681ull;
    v ^= v >> 21;
    v ^= v << 37;
    v ^= v >> 4;
    v *= 4768777513237032717ull;
    v ^= v << 20;
    v ^= v >> 41;
    v ^= v << 5;
    return v;
}

__global__ void rmat_kernel(u64 edges_number, u32 n_pow2, u32 h, double a, double ab, double abc, bool directed, u64* __restrict__ keys)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= edges_number) return;
    const u32 ii = (u32) i;
    u32 r = hash32((u32) (2 * ii) * h);                        // rmat_util.h:39-41
    const u32 stride = hash32((u32) (2 * ii + 1) * h);
    u32 x = 0, y = 0;
    for (u32 nn = n_pow2; nn > 1; nn >>= 1)                    // rMatRec, outermost level first (rmat_util.h:26-36)
    {
        const double d = (double) hash32(r) / (double) 0xFFFFFFFFu;
        if (d < a) { }
        else if (d < ab) y += nn >> 1;
        else if (d < abc) x += nn >> 1;
        else { x += nn >> 1; y += nn >> 1; }
        r += stride;
    }
    keys[i] = ((u64) x << 32) | y;
    if (!directed) keys[i + edges_number] = ((u64) y << 32) | x;   // utility.h:91-97
}

__global__ void csr_offsets_kernel(const u32* __restrict__ pairs, u64 count, u64 n, u64* __restrict__ offsets)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n) return;
    u64 lo = 0, hi = count;                                     // first pair with src >= v
    while (lo < hi)
    {
        const u64 mid = (lo + hi) >> 1;
        if (pairs[2 * mid] < v) lo = mid + 1; else hi = mid;
    }
    offsets[v] = lo;
}

__global__ void csr_edges_kernel(const u32* __restrict__ pairs, u64 count, u32* __restrict__ edges)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) edges[i] = pairs[2 * i + 1];
}
}  // namespace

void graph_build_desc(const LaunchCtx& lc, const u64* d_offsets_n, u64 n, u64 m, RowDesc* d_desc)
{
    if (n == 0) return;
    build_desc_kernel<<<div_up(n, 256), 256, 0, lc.stream>>>(d_offsets_n, n, m, d_desc);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void graph_check_csr(const LaunchCtx& lc, const u64* d_offsets_n, u64 n, u64 m, const u32* d_edges, u32* d_flag)
{
    const u64 work = n > m ? n : m;
    if (work == 0) return;
    check_csr_kernel<<<div_up(work, 256), 256, 0, lc.stream>>>(d_offsets_n, n, m, d_edges, d_flag);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void graph_check_ids(const LaunchCtx& lc, const u32* d_pairs, u64 count, u64 n, u32* d_flag)
{
    if (count == 0) return;
    check_ids_kernel<<<grid_for(2 * count, 256, 8, lc.sm_count), 256, 0, lc.stream>>>(d_pairs, 2 * count, n, d_flag);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void graph_prepare_batch(const LaunchCtx& lc, const u32* d_pairs, u64 count, bool drop_self_loops, u64 n, PreparedBatch& out,
                         u32* d_sorted_pairs_out)
{
    out.k = 0;
    out.n_sources = 0;
    if (count == 0) return;
    Trace tr(lc.stream);
    Scratch<u64> raw(lc.stream, count), sorted(lc.stream, count);
    pairs_to_keys_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(d_pairs, count, raw.ptr);
    lc.count();
    sort_keys(lc, raw.ptr, sorted.ptr, count, 32 + bits_for(n));
    if (d_sorted_pairs_out)
    {
        keys_to_pairs_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(sorted.ptr, count, d_sorted_pairs_out);
        lc.count();
    }
    Scratch<u8> flags(lc.stream, count);
    keep_flags_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(sorted.ptr, count, drop_self_loops, flags.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    out.keys.allocate(lc.stream, count);
    out.k = select_flagged(lc, sorted.ptr, flags.ptr, count, out.keys.ptr);
    tr.mark("prepare: sort+dedup");
    if (out.k == 0) return;

    head_flags_kernel<<<div_up(out.k, 256), 256, 0, lc.stream>>>(out.keys.ptr, out.k, flags.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    out.slice.allocate(lc.stream, out.k + 1);
    {
        Scratch<u64> d_count(lc.stream, 1);
        cub::CountingInputIterator<u64> iota(0);
        size_t temp_bytes = 0;
        WHARF_CUDA(cub::DeviceSelect::Flagged(nullptr, temp_bytes, iota, flags.ptr, out.slice.ptr, d_count.ptr, (i64) out.k, lc.stream));
        Scratch<u8> temp(lc.stream, temp_bytes);
        WHARF_CUDA(cub::DeviceSelect::Flagged(temp.ptr, temp_bytes, iota, flags.ptr, out.slice.ptr, d_count.ptr, (i64) out.k, lc.stream));
        lc.count_lib();
        WHARF_CUDA(cudaMemcpyAsync(&out.n_sources, d_count.ptr, 8, cudaMemcpyDeviceToHost, lc.stream));
        WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    }
    tr.mark("prepare: slices");
    out.sources.allocate(lc.stream, out.n_sources);
    sources_kernel<<<div_up(out.n_sources, 256), 256, 0, lc.stream>>>(out.keys.ptr, out.slice.ptr, out.n_sources, out.k, out.sources.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

// Rewrites every live row contiguously in vertex order (in the layout g.tail asks for) into the spare pool and
// swaps the two: garbage left by earlier batches is reclaimed, and a layout change (tail 0 -> 1) is the same copy.
// The spare is kept, so steady-state compaction allocates nothing.  Aux tails come out rebuilt: filter set,
// caches empty (the cache holds pure functions of (graph, key): dropping it is always safe).
void graph_compact(const LaunchCtx& lc, DeviceGraph& g, u64 extra_capacity, int source_tail)
{
    const u32 src_shift = source_tail < 0 ? g.tail : (u32) source_tail;      // layout of the rows being copied
    const u64 row_words = 1 + g.tail;
    const u64 live = g.m * row_words;
    u64 capacity = live + extra_capacity;
    capacity += capacity / 2 + (1u << 20);                       // half as much again of slack
    if (capacity >= (1ull << 32)) capacity = (1ull << 32) - 1;
    if (live + extra_capacity > capacity) throw CudaError(cudaErrorMemoryAllocation, "neighbour pool would exceed 2^32 words");
    Trace tr(lc.stream);
    Scratch<u64> deg(lc.stream, g.n + 1), begin(lc.stream, g.n + 1);
    degrees_kernel<<<div_up(g.n + 1, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, (u32) row_words, deg.ptr);
    lc.count();
    exclusive_sum(lc, deg.ptr, begin.ptr, g.n + 1);
    tr.mark("compact: degrees+scan");
    // the spare pool (the previous pool, or the one sized by graph_enable_tails) is reused as long as it leaves an eighth
    // of slack: growing it means a fresh device allocation of the whole pool (measured: 0.9 s inside an edge batch on
    // the LiveJournal-shape graph), and m only drifts by a fraction of a percent per batch
    const u64 need = live + extra_capacity;
    if (g.spare.count < need + need / 8) g.spare.allocate(capacity);
    tr.mark("compact: spare pool");
    copy_all_rows(lc, g.desc.ptr, g.n, g.pool.ptr, begin.ptr, g.spare.ptr, src_shift, g.tail);
    tr.mark("compact: copy_all_rows");
    repoint_kernel<<<div_up(g.n, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, begin.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    std::swap(g.pool, g.spare);
    g.pool_used = live;
    if (g.tail && src_shift == 0)
    {
        // rows just moved into the interleaved layout (aux words zero): triangle blocks while the graph is symmetric
        // (WHARF_TRI=0 keeps per-row blocks: A/B measurements), per-row blocks otherwise
        const char* env = std::getenv("WHARF_TRI");
        g.tri = (env && std::atoi(env) == 0) ? 0u : 1u;             // the build itself checks the symmetry and falls back
        g.tri_stale = 0;
        graph_aux_rebuild(lc, g);
    }
    else if (g.tail && !g.tri) graph_aux_rebuild(lc, g);        // triangle blocks travelled with their slots (copy_rows_kernel)
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    tr.mark("compact: repoint+aux");
}

namespace
{
constexpr u32 kTriHeavyCapacity = 1u << 22;      // hub - hub slots parked for the block-per-slot kernel (32 MB of scratch)

// Triangle blocks of every slot.  Returns false when the graph is not symmetric (some edge has no twin): the blocks are then
// meaningless and the caller falls back to per-row blocks.
bool tri_build_all(const LaunchCtx& lc, DeviceGraph& g)
{
    const u64 n = g.n;
    if (n == 0 || g.m == 0) return true;
    Scratch<u64> chunk_scan(lc.stream, n + 1);
    Scratch<u32> eoff(lc.stream, n + 1);
    {
        cub::CountingInputIterator<u64> iota(0);
        cub::TransformInputIterator<u64, TriChunksOf, cub::CountingInputIterator<u64>> chunks(iota, TriChunksOf{g.desc.ptr, n});
        cub::CountingInputIterator<u32> iota32(0);
        cub::TransformInputIterator<u32, DegreeOfRow, cub::CountingInputIterator<u32>> degrees(iota32, DegreeOfRow{g.desc.ptr, n});
        size_t bytes_a = 0, bytes_b = 0;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, bytes_a, chunks, chunk_scan.ptr, (i64) (n + 1), lc.stream));
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, bytes_b, degrees, eoff.ptr, (i64) (n + 1), lc.stream));
        Scratch<u8> temp(lc.stream, std::max(bytes_a, bytes_b));
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, bytes_a, chunks, chunk_scan.ptr, (i64) (n + 1), lc.stream));
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, bytes_b, degrees, eoff.ptr, (i64) (n + 1), lc.stream));
        lc.count_lib(2);
    }
    u64 n_chunks = 0;
    u32 slots = 0;
    WHARF_CUDA(cudaMemcpyAsync(&n_chunks, chunk_scan.ptr + n, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemcpyAsync(&slots, eoff.ptr + n, 4, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    if (n_chunks == 0) return true;
    const u64 words = (u64) slots / 2 + 2;
    Scratch<u32> blocks(lc.stream, words);                           // one 16-bit per-row filter block per slot
    Scratch<uint2> heavy(lc.stream, kTriHeavyCapacity);
    Scratch<u32> counters(lc.stream, 4);                             // [0] heavy pairs, [1] a twin is missing, [2] slots written
    WHARF_CUDA(cudaMemsetAsync(blocks.ptr, 0, words * sizeof(u32), lc.stream));
    WHARF_CUDA(cudaMemsetAsync(counters.ptr, 0, 16, lc.stream));
    const unsigned grid = div_up(n_chunks * 32, 256);
    tri_rowfilter_kernel<<<grid, 256, 0, lc.stream>>>(g.desc.ptr, g.pool.ptr, n, chunk_scan.ptr, n_chunks, eoff.ptr, blocks.ptr);
    const TriFilter f{reinterpret_cast<const unsigned short*>(blocks.ptr), eoff.ptr};
    tri_build_kernel<<<grid, 256, 0, lc.stream>>>(g.desc.ptr, g.pool.ptr, n, chunk_scan.ptr, n_chunks, f, heavy.ptr, kTriHeavyCapacity, counters.ptr);
    tri_heavy_kernel<<<lc.sm_count * 4, 256, 0, lc.stream>>>(g.desc.ptr, g.pool.ptr, f, heavy.ptr, kTriHeavyCapacity, counters.ptr);
    lc.count(3);
    WHARF_CUDA(cudaGetLastError());
    u32 host[4] = {0, 0, 0, 0};
    WHARF_CUDA(cudaMemcpyAsync(host, counters.ptr, 16, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    return host[1] == 0 && host[2] == slots;
}
}  // namespace

void graph_aux_rebuild(const LaunchCtx& lc, DeviceGraph& g)
{
    if (!g.tail || g.n == 0) return;
    if (g.tri)
    {
        graph_aux_clear_caches(lc, g);                            // cache halves empty; the build overwrites every filter half
        g.tri_stale = 0;
        if (tri_build_all(lc, g)) return;
        g.tri = 0;                                                // some edge has no twin: per-row blocks
    }
    aux_build_all_kernel<<<div_up(g.n * 32, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, g.pool.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

u64 graph_filter_check(const LaunchCtx& lc, const DeviceGraph& g)
{
    if (!g.tail || g.n == 0 || g.m == 0) return 0;
    Scratch<u64> scan(lc.stream, g.n + 1);
    Scratch<unsigned long long> missing(lc.stream, 1);
    {
        cub::CountingInputIterator<u64> iota(0);
        cub::TransformInputIterator<u64, SlotsOfRow, cub::CountingInputIterator<u64>> degrees(iota, SlotsOfRow{g.desc.ptr, g.n});
        size_t bytes = 0;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, bytes, degrees, scan.ptr, (i64) (g.n + 1), lc.stream));
        Scratch<u8> temp(lc.stream, bytes);
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, bytes, degrees, scan.ptr, (i64) (g.n + 1), lc.stream));
        lc.count_lib();
    }
    u64 slots = 0;
    WHARF_CUDA(cudaMemcpyAsync(&slots, scan.ptr + g.n, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemsetAsync(missing.ptr, 0, 8, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    if (slots == 0) return 0;
    filter_check_kernel<<<div_up(slots * 8, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.pool.ptr, g.n, scan.ptr, slots, g.tri, missing.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    unsigned long long bad = 0;
    WHARF_CUDA(cudaMemcpyAsync(&bad, missing.ptr, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    return bad;
}

void graph_aux_clear_caches(const LaunchCtx& lc, DeviceGraph& g)
{
    if (!g.tail || g.n == 0) return;
    aux_invalidate_kernel<<<div_up(g.n * 32, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, g.pool.ptr, nullptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void graph_enable_tails(const LaunchCtx& lc, DeviceGraph& g)
{
    if (g.tail) return;
    g.tail = 1;
    graph_compact(lc, g, 0, 0);                               // plain rows in, interleaved rows out
    // the pool the rows just left is too small to be the next compaction's target: size the spare now, while the
    // caller is still setting up a corpus, instead of inside an edge batch
    if (g.spare.count < g.pool.count) g.spare.allocate(g.pool.count);
}

void graph_apply_batch(const LaunchCtx& lc, DeviceGraph& g, const PreparedBatch& b, bool is_delete, u32* d_touched_bitmap)
{
    if (b.k == 0) return;
    const u64 k = b.k, ns = b.n_sources;
    Trace tr(lc.stream);
    Scratch<u32> rank(lc.stream, k), change(lc.stream, k + 1);
    Scratch<u64> change_scan(lc.stream, k + 1);
    locate_kernel<<<div_up(k + 1, 256), 256, 0, lc.stream>>>(b.keys.ptr, k, g.desc.ptr, g.pool.ptr, g.tail, is_delete, rank.ptr, change.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    exclusive_sum(lc, change.ptr, change_scan.ptr, k + 1);

    Scratch<u64> new_degree(lc.stream, ns + 1), alloc(lc.stream, ns + 1), chunks(lc.stream, ns + 1), alloc_scan(lc.stream, ns + 1), chunk_scan(lc.stream, ns + 1);
    plan_rows_kernel<<<div_up(ns + 1, 256), 256, 0, lc.stream>>>(b.sources.ptr, b.slice.ptr, ns, change_scan.ptr, g.desc.ptr, is_delete, 1 + g.tail,
                                                                 new_degree.ptr, alloc.ptr, chunks.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    exclusive_sum(lc, alloc.ptr, alloc_scan.ptr, ns + 1);
    exclusive_sum(lc, chunks.ptr, chunk_scan.ptr, ns + 1);

    Scratch<u32> asym(lc.stream, 1);
    u32 batch_asymmetric = 0;
    if (g.tail && g.tri)
    {
        // triangle blocks answer "x in N(prev)" through "prev in N(x)": they stand only while every batch keeps the graph symmetric
        WHARF_CUDA(cudaMemsetAsync(asym.ptr, 0, 4, lc.stream));
        symmetric_batch_kernel<<<div_up(k, 256), 256, 0, lc.stream>>>(b.keys.ptr, k, asym.ptr);
        lc.count();
        WHARF_CUDA(cudaMemcpyAsync(&batch_asymmetric, asym.ptr, 4, cudaMemcpyDeviceToHost, lc.stream));
    }
    u64 totals[3];   // appended pool words, chunks, changed edges
    WHARF_CUDA(cudaMemcpyAsync(&totals[0], alloc_scan.ptr + ns, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemcpyAsync(&totals[1], chunk_scan.ptr + ns, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemcpyAsync(&totals[2], change_scan.ptr + k, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    const u64 total = totals[0], n_chunks = totals[1], changed = totals[2];
    g.last_rewritten = total >> g.tail;
    tr.mark("apply: locate+plan");

    const bool compacted = g.pool_used + total > g.pool.count;
    if (compacted) graph_compact(lc, g, total);                            // rank[] is row-relative: still valid
    const u64 base = g.pool_used;
    tr.mark("apply: compact (if any)");

    if (n_chunks)
    {
        merge_old_rows_kernel<<<div_up(n_chunks * 32, 256), 256, 0, lc.stream>>>(b.sources.ptr, b.slice.ptr, ns, b.keys.ptr, change_scan.ptr,
                                                                                change.ptr, chunk_scan.ptr, n_chunks, alloc_scan.ptr, base,
                                                                                g.desc.ptr, g.pool.ptr, g.tail, is_delete);
        lc.count();
    }
    if (!is_delete)
    {
        write_new_edges_kernel<<<div_up(k, 256), 256, 0, lc.stream>>>(b.keys.ptr, k, b.slice.ptr, ns, rank.ptr, change.ptr, change_scan.ptr,
                                                                      alloc_scan.ptr, base, g.pool.ptr, g.tail);
        lc.count();
    }
    commit_rows_kernel<<<div_up(ns, 256), 256, 0, lc.stream>>>(b.sources.ptr, ns, new_degree.ptr, alloc_scan.ptr, base, g.desc.ptr,
                                                               g.version.ptr, d_touched_bitmap);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    tr.mark("apply: merge+write+commit");
    g.pool_used += total;
    g.m = is_delete ? g.m - changed : g.m + changed;
    if (g.tail && g.tri && batch_asymmetric)
    {
        // a directed batch: back to per-row blocks for good (every aux word rebuilt: filters set, caches empty)
        g.tri = 0;
        graph_aux_rebuild(lc, g);
        tr.mark("apply: aux tails (triangle blocks dropped)");
    }
    else if (g.tail && g.tri)
    {
        // the merge carried the blocks of the surviving slots and left new slots empty.  Inserts: one intersection per new
        // edge ORs the bits it adds.  Deletes leave their bits behind (a superset still proves absence); once a quarter of
        // the edges' worth has been deleted the blocks are rebuilt.  Caches whose cur is a source: emptied.
        if (!is_delete)
        {
            Scratch<uint2> heavy(lc.stream, kTriHeavyCapacity);
            Scratch<u32> heavy_count(lc.stream, 1);
            WHARF_CUDA(cudaMemsetAsync(heavy_count.ptr, 0, 4, lc.stream));
            tri_insert_kernel<<<div_up(k * 8, 256), 256, 0, lc.stream>>>(b.keys.ptr, k, change.ptr, g.desc.ptr, g.pool.ptr, heavy.ptr, heavy_count.ptr,
                                                                         kTriHeavyCapacity);
            tri_insert_heavy_kernel<<<lc.sm_count * 2, 256, 0, lc.stream>>>(g.desc.ptr, g.pool.ptr, heavy.ptr, heavy_count.ptr, kTriHeavyCapacity);
            lc.count(2);
        }
        else g.tri_stale += changed;
        if (g.tri_stale > g.m / 4 + 1024) graph_aux_rebuild(lc, g);
        else
        {
            aux_invalidate_kernel<<<div_up(g.n * 32, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, g.pool.ptr, d_touched_bitmap);
            lc.count();
        }
        WHARF_CUDA(cudaGetLastError());
        tr.mark("apply: aux tails (triangle blocks)");
    }
    else if (g.tail)
    {
        // fresh rows: filter set, caches empty; every other cache whose cur is a source: emptied
        aux_build_sources_kernel<<<div_up(ns * 32, 256), 256, 0, lc.stream>>>(b.sources.ptr, ns, g.desc.ptr, g.pool.ptr);
        aux_invalidate_kernel<<<div_up(g.n * 32, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, g.pool.ptr, d_touched_bitmap);
        lc.count(2);
        WHARF_CUDA(cudaGetLastError());
        tr.mark("apply: aux tails");
    }
}

void graph_flatten(const LaunchCtx& lc, const DeviceGraph& g, u64* d_row_ptr, u32* d_col)
{
    Scratch<u64> deg(lc.stream, g.n + 1);
    degrees_kernel<<<div_up(g.n + 1, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, 1u, deg.ptr);
    lc.count();
    exclusive_sum(lc, deg.ptr, d_row_ptr, g.n + 1);
    if (d_col) copy_all_rows(lc, g.desc.ptr, g.n, g.pool.ptr, d_row_ptr, d_col, g.tail, 0);
    WHARF_CUDA(cudaGetLastError());
}

void rmat_generate_device(const LaunchCtx& lc, u64 edges_number, u64 vertices_number, u64 batch_seed, bool self_loops, bool directed,
                          double a, double b, double c, u32* d_out, u64* count_out)
{
    *count_out = 0;
    if (edges_number == 0) return;
    u64 log2_up = 0;
    for (u64 t = vertices_number - 1; t > 0; t >>= 1) log2_up++;               // pbbslib/utilities.h:286-291
    const u64 n_pow2 = 1ull << (log2_up - 1);                                   // utility.h:78
    const u32 h = hash32((u32) hash64(0 + batch_seed));                         // rand.ith_rand(0) narrowed to intT, rmat_util.h:15
    const u64 total = directed ? edges_number : 2 * edges_number;
    Scratch<u64> raw(lc.stream, total), sorted(lc.stream, total);
    rmat_kernel<<<div_up(edges_number, 256), 256, 0, lc.stream>>>(edges_number, (u32) n_pow2, h, a, a + b, a + b + c, directed, raw.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    sort_keys(lc, raw.ptr, sorted.ptr, total, 32 + bits_for(n_pow2));
    Scratch<u8> flags(lc.stream, total);
    keep_flags_kernel<<<div_up(total, 256), 256, 0, lc.stream>>>(sorted.ptr, total, !self_loops, flags.ptr);
    lc.count();
    const u64 kept = select_flagged(lc, sorted.ptr, flags.ptr, total, raw.ptr);
    if (kept)
    {
        keys_to_pairs_kernel<<<div_up(kept, 256), 256, 0, lc.stream>>>(raw.ptr, kept, d_out);
        lc.count();
    }
    WHARF_CUDA(cudaGetLastError());
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    *count_out = kept;
}

void pairs_to_csr_device(const LaunchCtx& lc, u64 n, const u32* d_pairs, u64 count, u64* d_offsets, u32* d_edges)
{
    csr_offsets_kernel<<<div_up(n + 1, 256), 256, 0, lc.stream>>>(d_pairs, count, n, d_offsets);
    lc.count();
    if (count)
    {
        csr_edges_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(d_pairs, count, d_edges);
        lc.count();
    }
    WHARF_CUDA(cudaGetLastError());
}

}  // namespace wharf
