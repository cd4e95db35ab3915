// Host-side engine state and the launch wrappers implemented by the kernel translation units.
#pragma once

#include "common.cuh"
#include "../../include/wharfmh_b200.h"

#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <functional>
#include <string>
#include <vector>

namespace wharf
{
// Records the message wmh_last_error() returns on this thread and hands the status code back (engine.cu).
int report_failure(int code, const std::string& msg);

struct LaunchCtx
{
    cudaStream_t stream = nullptr;
    int sm_count = 148;
    u64* launches = nullptr;      // hand-written kernel launches
    u64* library_calls = nullptr; // CUB device-wide algorithm invocations
    void count(u64 k = 1) const { if (launches) *launches += k; }
    void count_lib(u64 k = 1) const { if (library_calls) *library_calls += k; }
};

// WHARF_TRACE=1: host-timed phases with the stream synchronised at every mark; WHARF_TRACE=2: host time between
// marks without synchronising (shows where the submitting thread blocks).  Tuning aid only, stderr.
struct Trace
{
    cudaStream_t stream;
    int mode;
    std::chrono::steady_clock::time_point t0;
    explicit Trace(cudaStream_t s) : stream(s), mode(std::getenv("WHARF_TRACE") ? std::atoi(std::getenv("WHARF_TRACE")) : 0)
    {
        if (mode == 1) cudaStreamSynchronize(s);
        if (mode) t0 = std::chrono::steady_clock::now();
    }
    void mark(const char* what)
    {
        if (!mode) return;
        if (mode == 1) cudaStreamSynchronize(stream);
        const auto t1 = std::chrono::steady_clock::now();
        const char* rank = std::getenv("RANK");
        std::fprintf(stderr, "[wharf trace%s%s] %-28s %9.3f ms\n", rank ? " r" : "", rank ? rank : "", what, std::chrono::duration<double, std::milli>(t1 - t0).count());
        t0 = t1;
    }
};

// Stream-ordered scratch allocation (cudaMallocAsync), released at scope exit.
template <class T>
struct Scratch
{
    T* ptr = nullptr;
    u64 count = 0;
    cudaStream_t stream = nullptr;
    Scratch() = default;
    Scratch(cudaStream_t s, u64 n) { allocate(s, n); }
    Scratch(const Scratch&) = delete;
    Scratch& operator=(const Scratch&) = delete;
    Scratch(Scratch&& o) noexcept : ptr(o.ptr), count(o.count), stream(o.stream) { o.ptr = nullptr; }
    Scratch& operator=(Scratch&& o) noexcept
    {
        if (this != &o) { release(); ptr = o.ptr; count = o.count; stream = o.stream; o.ptr = nullptr; }
        return *this;
    }
    ~Scratch() { release(); }
    void allocate(cudaStream_t s, u64 n)
    {
        release();
        stream = s;
        count = n;
        cudaError_t e = pool_malloc_async((void**) &ptr, (n ? n : 1) * sizeof(T), s);
        if (e != cudaSuccess)
        {
            ptr = nullptr;
            cudaGetLastError();
            throw CudaError(e, "cudaMallocAsync of " + std::to_string(n * sizeof(T)) + " bytes failed: " + cudaGetErrorString(e));
        }
    }
    void release()
    {
        if (ptr) cudaFreeAsync(ptr, stream);
        ptr = nullptr;
    }
};

// ------------------------------------------------------------------------------------------------
// Device adjacency: row descriptors + append-only neighbour pool (graph/snapshot.h FlatGraph on device).
// ------------------------------------------------------------------------------------------------
struct DeviceGraph
{
    u64 n = 0;                       // number_of_vertices()
    u64 m = 0;                       // number_of_edges() (directed entries)
    DeviceBuffer<RowDesc> desc;      // n
    DeviceBuffer<u32> pool;          // neighbour pool, rows ascending & unique (interleaved with aux words, see `tail`)
    u64 pool_used = 0;               // words of pool handed out (live + garbage)
    DeviceBuffer<u32> version;       // n: sampler-manager generation per vertex (graph/vertex.h:12)
    DeviceBuffer<u32> spare;         // second pool: compaction copies into it and swaps (no allocation in steady state)
    u32 tail = 0;                    // 1: neighbour i at begin + 2 i, its aux word (sampler cache | neighbour filter) at begin + 2 i + 1
                                     //    (MH node2vec); 0: plain rows.  Doubles as the index shift of every row accessor.
    u32 tri = 0;                     // with tail: 1 = the filter half of aux word i of row v is the TRIANGLE block of the slot v -> c
                                     //    (Bloom block over N(v) n N(c), c = neighbour i), kept while the graph is symmetric;
                                     //    0 = block i of row v's own neighbour filter (any graph)
    u64 last_rewritten = 0;          // neighbours in the rows the last batch rewrote (= sampler states of those sources' rows)
    u64 tri_stale = 0;               // edges deleted since the triangle blocks were last built: a delete leaves its bits behind
                                     //    (a superset still proves absence, it only says "maybe" more often)
};

// One prepared edge batch (sorted, deduplicated) and what the merge learned about it.
struct PreparedBatch
{
    u64 k = 0;                       // unique (src,dst) keys
    u64 n_sources = 0;               // distinct sources
    Scratch<u64> keys;               // k      : (src << 32) | dst, ascending
    Scratch<u32> sources;            // n_src  : distinct sources ascending
    Scratch<u64> slice;              // n_src+1: batch slice of each source in keys
};

// ------------------------------------------------------------------------------------------------
// Peer memory between the ranks of one box (include/wharfmh_b200.h, wmh_peer_*): every rank owns one exchange
// buffer; buf[g] is rank g's buffer mapped into this process (CUDA IPC over NVLink), buf[rank] the own one.
// Layout: a flag page (arrival epochs written by the peers, an error word), then the payload.
// ------------------------------------------------------------------------------------------------
constexpr u32 kPeerMaxWorld = WMH_PEER_MAX_WORLD;
constexpr u64 kPeerFlagBytes = 4096;
constexpr u64 kPeerErrorOffset = 1024;   // u32 in the own flag page: a barrier gave up waiting
struct PeerPtrs { u8* buf[kPeerMaxWorld]; };
struct PeerLink
{
    u32 world = 1, rank = 0;
    PeerPtrs ptrs{};
    u64 bytes = 0;                   // size of every rank's buffer (flag page included)
    u64 epoch = 0;                   // barriers passed
    u64 exchanges = 0;               // generate passes that used the exchange
    bool created = false, attached = false;
    u8* own() const { return ptrs.buf[rank]; }
    u64 payload_bytes() const { return bytes > kPeerFlagBytes ? bytes - kPeerFlagBytes : 0; }
};
// All ranks meet: every kernel enqueued before it on every rank's stream has completed (and its peer stores are
// visible) before any kernel enqueued after it starts.
void peer_barrier(const LaunchCtx& lc, PeerLink& link);

// ---- graph kernels (kernels_graph.cu) -------------------------------------------------------------
void graph_build_desc(const LaunchCtx& lc, const u64* d_offsets_n, u64 n, u64 m, RowDesc* d_desc);
void graph_prepare_batch(const LaunchCtx& lc, const u32* d_pairs, u64 count, bool drop_self_loops, u64 n,
                         PreparedBatch& out, u32* d_sorted_pairs_out /* nullable: count pairs, sorted (not deduped) */);
// Applies the batch; returns the pool elements appended. Updates g.m, g.pool_used, version[], touched bitmap.
void graph_apply_batch(const LaunchCtx& lc, DeviceGraph& g, const PreparedBatch& b, bool is_delete, u32* d_touched_bitmap);
// source_tail: layout (DeviceGraph::tail) of the rows being copied when it differs from g.tail (a layout switch); -1 = g.tail.
void graph_compact(const LaunchCtx& lc, DeviceGraph& g, u64 extra_capacity, int source_tail = -1);
// Aux tails (MH node2vec): switch the layout on (one compaction), rebuild every tail (filter set, caches empty), or
// only empty the sampler caches.
void graph_enable_tails(const LaunchCtx& lc, DeviceGraph& g);
void graph_aux_rebuild(const LaunchCtx& lc, DeviceGraph& g);
u64  graph_filter_check(const LaunchCtx& lc, const DeviceGraph& g);      // slots whose membership filter lacks a required bit (must be 0)
void graph_aux_clear_caches(const LaunchCtx& lc, DeviceGraph& g);
void graph_flatten(const LaunchCtx& lc, const DeviceGraph& g, u64* d_row_ptr /* n+1 */, u32* d_col /* m */);
// flag |= 1: offsets not monotone / beyond m; 2: a neighbour id >= n; 4: a row not strictly ascending.
void graph_check_csr(const LaunchCtx& lc, const u64* d_offsets_n, u64 n, u64 m, const u32* d_edges, u32* d_flag);
void graph_check_ids(const LaunchCtx& lc, const u32* d_pairs, u64 count, u64 n, u32* d_flag);
void rmat_generate_device(const LaunchCtx& lc, u64 edges_number, u64 vertices_number, u64 batch_seed, bool self_loops,
                          bool directed, double a, double b, double c, u32* d_out, u64* count_out);
void pairs_to_csr_device(const LaunchCtx& lc, u64 n, const u32* d_pairs, u64 count, u64* d_offsets, u32* d_edges);

// ---- walk kernels (kernels_walk.cu) ---------------------------------------------------------------
struct WalkParams
{
    const RowDesc* desc = nullptr;
    const u32* pool = nullptr;
    const u32* version = nullptr;
    u32* pool_rw = nullptr;          // the same pool, writable: aux words (MH node2vec; DeviceGraph::tail must be 1)
    u32 shift = 0;                   // DeviceGraph::tail: neighbour i of a row sits at pool[begin + (i << shift)]
    u32 tri = 0;                     // DeviceGraph::tri: the filter halves of the aux words are triangle blocks
    const u64* draw = nullptr;       // X[wpv][L]
    u32* walks = nullptr;            // [L][W_local] position-major
    u64 n = 0;                       // global vertices
    u64 w_local = 0;                 // local walks
    ShardMap map;                    // which walks this rank owns, local walk -> (replica, start vertex)
    u32 wpv = 0, L = 0;
    u32 model = 0, deterministic = 1, init_strategy = 2;
    float p = 1.f, q = 1.f;
    u64 seed = 0;
};
void walks_generate(const LaunchCtx& lc, const WalkParams& wp, unsigned long long* d_steps);
// MH node2vec with peers attached: the first-touch sampler value of every state (cur, prev) - a pure function of
// the replicated graph - is computed by one rank per state and stored into every peer's exchange buffer, then
// copied into the aux tails, so the walk kernel starts with warm caches on every rank.  Payload: 2 bytes per edge.
void n2v_share_sampler_init(const LaunchCtx& lc, const WalkParams& wp, u64 m, PeerLink& link);
void walks_rewalk(const LaunchCtx& lc, const WalkParams& wp, const u32* d_affected, const u32* d_positions, u64 count,
                  unsigned long long* d_steps);
// Deterministic mode only: step-synchronous generation that advances walkers standing on the same vertex once
// (kernels_coalesced.cu).  host_draw = X[wpv][L] on the host.
// after_replica(first, count), when given, is called once the local walks [first, first + count) of one replica have
// been fully enqueued on lc.stream.
struct CoalescedPlan;                       // scratch + captured CUDA graph of one generate pass (kernels_coalesced.cu)
void coalesced_plan_release(CoalescedPlan*& plan);
void walks_generate_coalesced(const LaunchCtx& lc, const WalkParams& wp, const u64* host_draw, unsigned long long* d_steps,
                              const std::function<void(u64, u64)>& after_replica, CoalescedPlan** plan_slot);
// Bulk generation of the local walks [first, first + count) with one walker per thread.
void walks_generate_range(const LaunchCtx& lc, const WalkParams& wp, u64 first, u64 count, unsigned long long* d_steps);
// min_pos[w] = first position whose vertex is in the touched bitmap (kNoPos when none).
void walks_scan_touched(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, const u32* d_bitmap, u32* d_min_pos);
// The same with the bitmap in static device memory (no allocation of any kind; at most kNoAllocMaxVertices vertices): marks
// `count` device-resident vertex ids and scans.  Throws std::length_error beyond the maximum.
constexpr u64 kNoAllocMaxVertices = 1ull << 25;
void walks_scan_touched_no_alloc(const LaunchCtx& lc, int device, const u32* d_walks, u64 w_local, u32 L, u64 n, const u32* d_vertices, u64 count,
                                 u32* d_min_pos);
// Compacts min_pos != kNoPos into (ids, positions); returns the count.
u64  walks_compact_affected(const LaunchCtx& lc, const u32* d_min_pos, u64 w_local, u32* d_ids, u32* d_positions);
void walks_transpose_rows(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, u64 first, u64 count, u32* d_out);
void walks_gather_rows(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, const u32* d_ids, u64 count, u32* d_out);
void walks_fingerprint(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, unsigned long long* d_sum);
// MetropolisHastingsSampler::sample as a chain on the fixed state (cur, prev): d_counts[j] (degree(cur) entries, zeroed by the
// caller) += samples returning neighbour j; *d_init = the init value.
void sampler_chain(const LaunchCtx& lc, const WalkParams& wp, u32 cur, u32 prev, u64 iters, u32* d_init, unsigned long long* d_counts);
// out[i] = global walk id of local walk local_ids[i] (ascending local ids give ascending global ids).
void walks_global_ids(const LaunchCtx& lc, const ShardMap& map, const u32* d_local_ids, u64 count, u64* d_out);
void fill_u32(const LaunchCtx& lc, u32* d, u64 count, u32 value);

// ---- index kernels (kernels_index.cu) -------------------------------------------------------------
struct DeviceIndex
{
    bool valid = false;
    u64 entries = 0;                 // stored steps (walk matrix cells holding a vertex)
    u32 chunks = 1;                  // the corpus is indexed in chunks of whole positions: cell ids stay 32-bit at any size
    u32 positions_per_chunk = 1;
    DeviceBuffer<u64> ptr;           // chunks x (n+1): segment of each vertex in `cells`, absolute offsets
    DeviceBuffer<u32> cells;         // cell id = (position - chunk's first position) * W_local + local_walk, ascending per vertex and chunk
};
void index_build(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, u64 n, DeviceIndex& idx);
// atomicMin of the position of every index entry of every source into min_pos; returns entries visited.
u64  index_search(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_sources, u64 n_sources, u64 w_local, u32* d_min_pos);
u64  index_count_touched(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_sources, u64 n_sources);
void index_find_next(const LaunchCtx& lc, const DeviceIndex& idx, u64 n, const u32* d_walks, u64 w_local, u32 L, u32 vertex,
                     u64 local_walk, u32 position, u32* d_out2 /* [next, found] */);
void compressed_walks_encode(const LaunchCtx& lc, const DeviceIndex& idx, const u32* d_walks, u64 w_local, u32 L, u64 n,
                             const ShardMap& map, u64* d_codes);
void compressed_find_next(const LaunchCtx& lc, const u64* d_codes, u64 count, u64 walk_id, u32 position, u32 L, u32* d_out2);

// ---- pairing kernels (kernels_pairings.cu) --------------------------------------------------------
void szudzik_pair_u64(const LaunchCtx& lc, const u64* x, const u64* y, u64* out, u64 count, unsigned long long* d_overflow);
void szudzik_unpair_u64(const LaunchCtx& lc, const u64* z, u64* x, u64* y, u64 count);
void szudzik_pair_u32(const LaunchCtx& lc, const u32* x, const u32* y, u32* out, u64 count, unsigned long long* d_overflow);
void szudzik_unpair_u32(const LaunchCtx& lc, const u32* z, u32* x, u32* y, u64 count);
void szudzik_pair_triplet_u64(const LaunchCtx& lc, const u64* a, const u64* b, const u64* c, u64* out, u64 count,
                              unsigned long long* d_overflow);
void szudzik_unpair_triplet_u64(const LaunchCtx& lc, const u64* z, u64* a, u64* b, u64* c, u64 count);

}  // namespace wharf
