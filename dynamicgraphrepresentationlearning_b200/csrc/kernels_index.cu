// Inverted index of the walk corpus and its searches.
//
// Replaces walks/inverted_index.h:12-92 (per-vertex PAM map key=walk_id*L+pos -> next), its construction in
// graph/wharfmh.h:312-354 (cuckoo accumulate + per-vertex sample_sort + PAM build), the map-of-changes scan of
// :519-537 / :669-687 and walks/compressed_walks.h:13-90.
//
// Layout: one radix sort of the walk matrix by vertex.  For vertex v, cells[ptr[v] .. ptr[v+1]) are the flat
// cell ids (position * W_local + local_walk) of the walker steps standing on v, ascending.  The `next` value
// the reference stores beside each key is not materialised: it is the matrix cell one position further.
#include "engine.hpp"
#include "pairings.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cub/device/device_segmented_sort.cuh>

namespace wharf
{
namespace
{
__global__ void iota_kernel(u32* __restrict__ out, u64 count)
{
    for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64) gridDim.x * blockDim.x) out[i] = (u32) i;
}

// ptr[v] = first sorted entry whose (masked) vertex is >= v
__global__ void segment_ptr_kernel(const u32* __restrict__ sorted_vertices, u64 count, u32 mask, u64 n, u64* __restrict__ ptr)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n) return;
    u64 lo = 0, hi = count;
    while (lo < hi)
    {
        const u64 mid = (lo + hi) >> 1;
        if ((sorted_vertices[mid] & mask) < v) lo = mid + 1; else hi = mid;
    }
    ptr[v] = lo;
}

__global__ void segment_len_kernel(const u64* __restrict__ ptr, const u32* __restrict__ sources, u64 n_sources, u64* __restrict__ len)
{
    const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (j > n_sources) return;
    len[j] = j < n_sources ? ptr[sources[j] + 1] - ptr[sources[j]] : 0;
}

// One thread per index entry of a touched vertex: walk = cell % W, position = cell / W, keep the minimum
// position per walk (types::MapOfChanges, graph/wharfmh.h:519-537).
__global__ void __launch_bounds__(256) search_kernel(const u64* __restrict__ ptr, const u32* __restrict__ cells, const u32* __restrict__ sources,
                                                     u64 n_sources, const u64* __restrict__ seg_scan, u64 total, u64 w_local,
                                                     u32* __restrict__ min_pos)
{
    const u64 e = (u64) blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    const u32 j = upper_segment_u64(seg_scan, (u32) n_sources + 1, e);
    const u32 cell = cells[ptr[sources[j]] + (e - seg_scan[j])];
    const u32 pos = (u32) (cell / w_local);
    const u32 w = (u32) (cell - (u64) pos * w_local);
    atomicMin(min_pos + w, pos);
}

// InvertedIndex::find_next (walks/inverted_index.h:74-92) for one key.
__global__ void find_next_kernel(const u64* __restrict__ ptr, const u32* __restrict__ cells, const u32* __restrict__ walks, u64 w_local, u32 L,
                                 u32 vertex, u64 local_walk, u32 position, u32* __restrict__ out2)
{
    const u64 lo0 = ptr[vertex], hi0 = ptr[vertex + 1];
    const u64 cell = (u64) position * w_local + local_walk;
    u64 lo = lo0, hi = hi0;
    while (lo < hi)
    {
        const u64 mid = (lo + hi) >> 1;
        if (cells[mid] < cell) lo = mid + 1; else hi = mid;
    }
    const bool found = lo < hi0 && cells[lo] == cell;
    u32 next = kRefSentinel;
    if (found && position + 1 < L)
    {
        next = walks[(u64) (position + 1) * w_local + local_walk];
        if (next == kNoVertex) next = kRefSentinel;       // walker stopped here (wharfmh.h:277-286)
    }
    out2[0] = next;
    out2[1] = found;
}

// code = pair(walk_id * L + position, next)   (walks/compressed_walks.h:52-86 decodes exactly this)
__global__ void __launch_bounds__(256) encode_codes_kernel(const u32* __restrict__ cells, u64 entries, const u32* __restrict__ walks, u64 w_local,
                                                           u32 L, u64 n, ShardMap map, u64* __restrict__ codes)
{
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= entries) return;
    const u32 cell = cells[i];
    const u32 pos = (u32) (cell / w_local);
    const u64 w = cell - (u64) pos * w_local;
    const u64 walk_id = map.global_id(w);
    u32 next = kRefSentinel;
    if (pos + 1 < L)
    {
        next = walks[(u64) (pos + 1) * w_local + w];
        if (next == kNoVertex) next = kRefSentinel;
    }
    codes[i] = szudzik_pair(walk_id * L + pos, next, nullptr);
}

__global__ void compressed_find_next_kernel(const u64* __restrict__ codes, u64 count, u64 walk_id, u32 position, u32 L, u32* __restrict__ out2)
{
    for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64) gridDim.x * blockDim.x)
    {
        u64 key, next;
        szudzik_unpair(codes[i], &key, &next);
        const u64 w = key / L;
        if (w == walk_id && key - w * L == position)
        {
            out2[0] = (u32) next;
            out2[1] = 1;
        }
    }
}
}  // namespace

void index_build(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, u64 n, DeviceIndex& idx)
{
    const u64 cells = w_local * L;
    if (cells >= (1ull << 32)) throw CudaError(cudaErrorInvalidValue, "inverted index needs local_walks * walk_length < 2^32");
    idx.valid = false;
    idx.ptr.ensure(n + 1);
    idx.cells.ensure(cells ? cells : 1);
    int bits = 1;
    while (bits < 32 && (1ull << bits) < n + 1) bits++;        // kNoVertex's low bits (all ones) sort after every vertex
    const u32 mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1);
    Scratch<u32> iota(lc.stream, cells), sorted_vertices(lc.stream, cells);
    if (cells)
    {
        iota_kernel<<<grid_for(cells, 256, 8, lc.sm_count), 256, 0, lc.stream>>>(iota.ptr, cells);
        lc.count();
        size_t temp_bytes = 0;
        WHARF_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, d_walks, sorted_vertices.ptr, iota.ptr, idx.cells.ptr, (i64) cells, 0, bits, lc.stream));
        Scratch<u8> temp(lc.stream, temp_bytes);
        WHARF_CUDA(cub::DeviceRadixSort::SortPairs(temp.ptr, temp_bytes, d_walks, sorted_vertices.ptr, iota.ptr, idx.cells.ptr, (i64) cells, 0, bits, lc.stream));
        lc.count_lib();
    }
    segment_ptr_kernel<<<div_up(n + 1, 256), 256, 0, lc.stream>>>(sorted_vertices.ptr, cells, mask, n, idx.ptr.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    WHARF_CUDA(cudaMemcpyAsync(&idx.entries, idx.ptr.ptr + n, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    idx.valid = true;
}

static u64 touched_segments(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_sources, u64 n_sources, Scratch<u64>& seg_scan)
{
    Scratch<u64> len(lc.stream, n_sources + 1);
    seg_scan.allocate(lc.stream, n_sources + 1);
    segment_len_kernel<<<div_up(n_sources + 1, 256), 256, 0, lc.stream>>>(idx.ptr.ptr, d_sources, n_sources, len.ptr);
    lc.count();
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, len.ptr, seg_scan.ptr, (i64) (n_sources + 1), lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, temp_bytes, len.ptr, seg_scan.ptr, (i64) (n_sources + 1), lc.stream));
    lc.count_lib();
    u64 total = 0;
    WHARF_CUDA(cudaMemcpyAsync(&total, seg_scan.ptr + n_sources, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    return total;
}

u64 index_count_touched(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_sources, u64 n_sources)
{
    if (n_sources == 0) return 0;
    Scratch<u64> seg_scan;
    return touched_segments(lc, idx, d_sources, n_sources, seg_scan);
}

u64 index_search(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_sources, u64 n_sources, u64 w_local, u32* d_min_pos)
{
    if (n_sources == 0) return 0;
    Scratch<u64> seg_scan;
    const u64 total = touched_segments(lc, idx, d_sources, n_sources, seg_scan);
    if (total)
    {
        search_kernel<<<div_up(total, 256), 256, 0, lc.stream>>>(idx.ptr.ptr, idx.cells.ptr, d_sources, n_sources, seg_scan.ptr, total, w_local, d_min_pos);
        lc.count();
        WHARF_CUDA(cudaGetLastError());
    }
    return total;
}

void index_find_next(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_walks, u64 w_local, u32 L, u32 vertex, u64 local_walk,
                     u32 position, u32* d_out2)
{
    find_next_kernel<<<1, 1, 0, lc.stream>>>(idx.ptr.ptr, idx.cells.ptr, d_walks, w_local, L, vertex, local_walk, position, d_out2);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void compressed_walks_encode(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_walks, u64 w_local, u32 L, u64 n, const ShardMap& map,
                             u64* d_codes)
{
    if (idx.entries == 0) return;
    Scratch<u64> raw(lc.stream, idx.entries);
    encode_codes_kernel<<<div_up(idx.entries, 256), 256, 0, lc.stream>>>(idx.cells.ptr, idx.entries, d_walks, w_local, L, n, map, raw.ptr);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
    // the reference keeps each vertex's codes as an ordered set (C-tree): sort every segment ascending
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceSegmentedSort::SortKeys(nullptr, temp_bytes, raw.ptr, d_codes, (i64) idx.entries, (i64) n, idx.ptr.ptr, idx.ptr.ptr + 1, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceSegmentedSort::SortKeys(temp.ptr, temp_bytes, raw.ptr, d_codes, (i64) idx.entries, (i64) n, idx.ptr.ptr, idx.ptr.ptr + 1, lc.stream));
    lc.count_lib();
}

void compressed_find_next(const LaunchCtx& lc, const u64* d_codes, u64 count, u64 walk_id, u32 position, u32 L, u32* d_out2)
{
    WHARF_CUDA(cudaMemsetAsync(d_out2, 0, 2 * sizeof(u32), lc.stream));
    if (count == 0) return;
    compressed_find_next_kernel<<<grid_for(count, 256, 8, lc.sm_count), 256, 0, lc.stream>>>(d_codes, count, walk_id, position, L, d_out2);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

}  // namespace wharf
