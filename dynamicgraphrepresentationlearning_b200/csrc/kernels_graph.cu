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
        if (is_delete)
        {
            if (lo < b_n && (u32) keys[b_lo + lo] == a && change[b_lo + lo]) continue;     // removed
            out[(u64) (ia - shift) << rs] = a;
        }
        else
        {
            out[(u64) (ia + shift) << rs] = a;
        }
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
    pool[pool_base + alloc_scan[j] + ((rank[i] + within) << rs)] = (u32) keys[i];
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
    for (u32 i = sub; i < d.y; i += 8) dst[o + ((u64) i << ds)] = pool[d.x + ((u64) i << ss)];
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
    for (u32 i = threadIdx.x; i < d.y; i += 256) dst[o + ((u64) i << ds)] = pool[d.x + ((u64) i << ss)];
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
    if (g.tail) graph_aux_rebuild(lc, g);
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    tr.mark("compact: repoint+aux");
}

void graph_aux_rebuild(const LaunchCtx& lc, DeviceGraph& g)
{
    if (!g.tail || g.n == 0) return;
    aux_build_all_kernel<<<div_up(g.n * 32, 256), 256, 0, lc.stream>>>(g.desc.ptr, g.n, g.pool.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
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

    u64 totals[3];   // appended pool words, chunks, changed edges
    WHARF_CUDA(cudaMemcpyAsync(&totals[0], alloc_scan.ptr + ns, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemcpyAsync(&totals[1], chunk_scan.ptr + ns, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaMemcpyAsync(&totals[2], change_scan.ptr + k, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    const u64 total = totals[0], n_chunks = totals[1], changed = totals[2];
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
    if (g.tail)
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
