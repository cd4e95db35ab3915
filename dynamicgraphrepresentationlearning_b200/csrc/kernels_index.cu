// Inverted index of the walk corpus and its searches.
//
// Replaces walks/inverted_index.h:12-92 (per-vertex PAM map key=walk_id*L+pos -> next), its construction in
// graph/wharfmh.h:312-354 (cuckoo accumulate + per-vertex sample_sort + PAM build), the map-of-changes scan of
// :519-537 / :669-687 and walks/compressed_walks.h:13-90.
//
// Layout: the walk matrix sorted by vertex, in CHUNKS of whole positions so that a cell id fits 32 bits whatever the size
// of the corpus (the reference's own key walk_id * L + pos is u32 and wraps beyond 5.37 M vertices): chunk c covers the
// positions [c * P, (c + 1) * P), P * W_local <= chunk limit.  For vertex v, cells[ptr[c][v] .. ptr[c][v+1]) are the cell
// ids ((position - c * P) * W_local + local_walk) of the walker steps of chunk c standing on v, ascending; the position-major
// walk matrix makes a chunk's sort keys one contiguous slice of it.  Corpora below the limit are one chunk.  The `next`
// value the reference stores beside each key is not materialised: it is the matrix cell one position further.
#include "engine.hpp"
#include "pairings.cuh"

#include <algorithm>
#include <cstdlib>
#include <vector>

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

// ptr[v] = base + first sorted entry whose (masked) vertex is >= v
__global__ void segment_ptr_kernel(const u32* __restrict__ sorted_vertices, u64 count, u32 mask, u64 n, u64 base, u64* __restrict__ ptr)
{
    const u64 v = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (v > n) return;
    u64 lo = 0, hi = count;
    while (lo < hi)
    {
        const u64 mid = (lo + hi) >> 1;
        if ((sorted_vertices[mid] & mask) < v) lo = mid + 1; else hi = mid;
    }
    ptr[v] = base + lo;
}

__global__ void segment_len_kernel(const u64* __restrict__ ptr, const u32* __restrict__ sources, u64 n_sources, u64* __restrict__ len)
{
    const u64 j = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (j > n_sources) return;
    len[j] = j < n_sources ? ptr[sources[j] + 1] - ptr[sources[j]] : 0;
}

// One thread per index entry of a touched vertex (one chunk): walk = cell % W, position = first_pos + cell / W, keep the
// minimum position per walk (types::MapOfChanges, graph/wharfmh.h:519-537).
__global__ void __launch_bounds__(256) search_kernel(const u64* __restrict__ ptr, const u32* __restrict__ cells, const u32* __restrict__ sources,
                                                     u64 n_sources, const u64* __restrict__ seg_scan, u64 total, u64 w_local, u32 first_pos,
                                                     u32* __restrict__ min_pos)
{
    const u64 e = (u64) blockIdx.x * 256 + threadIdx.x;
    if (e >= total) return;
    const u32 j = upper_segment_u64(seg_scan, (u32) n_sources + 1, e);
    const u32 cell = cells[ptr[sources[j]] + (e - seg_scan[j])];
    const u32 pos = (u32) (cell / w_local);
    const u32 w = (u32) (cell - (u64) pos * w_local);
    atomicMin(min_pos + w, first_pos + pos);
}

// InvertedIndex::find_next (walks/inverted_index.h:74-92) for one key.
__global__ void find_next_kernel(const u64* __restrict__ ptr /* of the position's chunk */, const u32* __restrict__ cells, const u32* __restrict__ walks,
                                 u64 w_local, u32 L, u32 first_pos, u32 vertex, u64 local_walk, u32 position, u32* __restrict__ out2)
{
    const u64 lo0 = ptr[vertex], hi0 = ptr[vertex + 1];
    const u64 cell = (u64) (position - first_pos) * w_local + local_walk;
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

// Cells per chunk: 2^31 keeps the sort's scratch (keys, values and their alternate buffers: 16 bytes per cell) at 32 GB;
// WHARF_INDEX_CHUNK_CELLS lowers it (tests drive several chunks with small corpora).
static u64 index_chunk_limit()
{
    u64 limit = 1ull << 31;
    if (const char* env = std::getenv("WHARF_INDEX_CHUNK_CELLS"))
    {
        const u64 v = std::strtoull(env, nullptr, 10);
        if (v) limit = std::min<u64>(v, (1ull << 32) - 1);
    }
    return limit;
}

void index_build(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, u64 n, DeviceIndex& idx)
{
    if (w_local >= (1ull << 32)) throw CudaError(cudaErrorInvalidValue, "inverted index needs fewer than 2^32 walks per rank");
    const u64 cells = w_local * L;
    idx.valid = false;
    idx.positions_per_chunk = (u32) std::max<u64>(1, std::min<u64>(L ? L : 1, index_chunk_limit() / std::max<u64>(w_local, 1)));
    idx.chunks = L ? (L + idx.positions_per_chunk - 1) / idx.positions_per_chunk : 1;
    idx.ptr.ensure((u64) idx.chunks * (n + 1));
    idx.cells.ensure(cells ? cells : 1);
    int bits = 1;
    while (bits < 32 && (1ull << bits) < n + 1) bits++;        // kNoVertex's low bits (all ones) sort after every vertex
    const u32 mask = bits >= 32 ? 0xFFFFFFFFu : ((1u << bits) - 1);
    const u64 chunk_cells = std::min<u64>(cells, (u64) idx.positions_per_chunk * w_local);
    Scratch<u32> iota(lc.stream, chunk_cells), sorted_vertices(lc.stream, chunk_cells);
    Scratch<u8> temp;
    size_t temp_bytes = 0;
    if (chunk_cells)
    {
        iota_kernel<<<grid_for(chunk_cells, 256, 8, lc.sm_count), 256, 0, lc.stream>>>(iota.ptr, chunk_cells);
        lc.count();
        WHARF_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, d_walks, sorted_vertices.ptr, iota.ptr, idx.cells.ptr, (i64) chunk_cells, 0, bits, lc.stream));
        temp.allocate(lc.stream, temp_bytes);
    }
    for (u32 c = 0; c < idx.chunks; c++)
    {
        const u32 first = c * idx.positions_per_chunk;
        const u64 count = (u64) std::min<u32>(idx.positions_per_chunk, L - first) * w_local, base = (u64) first * w_local;
        if (count)
        {
            size_t bytes = temp_bytes;
            WHARF_CUDA(cub::DeviceRadixSort::SortPairs(temp.ptr, bytes, d_walks + base, sorted_vertices.ptr, iota.ptr, idx.cells.ptr + base, (i64) count, 0, bits, lc.stream));
            lc.count_lib();
        }
        segment_ptr_kernel<<<div_up(n + 1, 256), 256, 0, lc.stream>>>(sorted_vertices.ptr, count, mask, n, base, idx.ptr.ptr + (u64) c * (n + 1));
        lc.count();
    }
    WHARF_CUDA(cudaGetLastError());
    std::vector<u64> ends(idx.chunks);
    for (u32 c = 0; c < idx.chunks; c++)
        WHARF_CUDA(cudaMemcpyAsync(&ends[c], idx.ptr.ptr + (u64) c * (n + 1) + n, 8, cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    idx.entries = 0;
    for (u32 c = 0; c < idx.chunks; c++) idx.entries += ends[c] - (u64) c * idx.positions_per_chunk * w_local;
    idx.valid = true;
}

static u64 touched_segments(const LaunchCtx& lc, const u64* d_ptr, const u32* d_sources, u64 n_sources, Scratch<u64>& seg_scan)
{
    Scratch<u64> len(lc.stream, n_sources + 1);
    seg_scan.allocate(lc.stream, n_sources + 1);
    segment_len_kernel<<<div_up(n_sources + 1, 256), 256, 0, lc.stream>>>(d_ptr, d_sources, n_sources, len.ptr);
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

u64 index_count_touched(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_sources, u64 n_sources)
{
    if (n_sources == 0) return 0;
    u64 total = 0;
    for (u32 c = 0; c < idx.chunks; c++)
    {
        Scratch<u64> seg_scan;
        total += touched_segments(lc, idx.ptr.ptr + (u64) c * (n + 1), d_sources, n_sources, seg_scan);
    }
    return total;
}

u64 index_search(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_sources, u64 n_sources, u64 w_local, u32* d_min_pos)
{
    if (n_sources == 0) return 0;
    u64 visited = 0;
    for (u32 c = 0; c < idx.chunks; c++)
    {
        const u64* d_ptr = idx.ptr.ptr + (u64) c * (n + 1);
        Scratch<u64> seg_scan;
        const u64 total = touched_segments(lc, d_ptr, d_sources, n_sources, seg_scan);
        if (total)
        {
            search_kernel<<<div_up(total, 256), 256, 0, lc.stream>>>(d_ptr, idx.cells.ptr, d_sources, n_sources, seg_scan.ptr, total, w_local,
                                                                     c * idx.positions_per_chunk, d_min_pos);
            lc.count();
            WHARF_CUDA(cudaGetLastError());
        }
        visited += total;
    }
    return visited;
}

void index_find_next(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_walks, u64 w_local, u32 L, u32 vertex, u64 local_walk,
                     u32 position, u32* d_out2)
{
    const u32 c = std::min(position / idx.positions_per_chunk, idx.chunks - 1);
    find_next_kernel<<<1, 1, 0, lc.stream>>>(idx.ptr.ptr + (u64) c * (n + 1), idx.cells.ptr, d_walks, w_local, L, c * idx.positions_per_chunk, vertex,
                                             local_walk, position, d_out2);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void compressed_walks_encode(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_walks, u64 w_local, u32 L, u64 n, const ShardMap& map,
                             u64* d_codes)
{
    if (idx.entries == 0) return;
    if (idx.chunks != 1) throw CudaError(cudaErrorInvalidValue, "the paired corpus is built from a one-chunk index");
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
