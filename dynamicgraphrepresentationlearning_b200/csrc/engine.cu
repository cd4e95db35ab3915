// Host-side engine and the C ABI of include/wharfmh_b200.h.
//
// Mirrors dygrl::WharfMH (graph/wharfmh.h:21): graph construction, generate_initial_random_walks (:250),
// insert/delete_edges_batch (:439/:588) with batch_walk_update (:733), walk retrieval (:365,:404),
// flatten_graph (:175), destroy/destroy_index (:228,:237), memory_footprint (:928) and the four timers of
// config/config.h:10-14.  All compute runs in the CUDA kernels of this directory; there is no host fallback.
#include "engine.hpp"
#include "pairings.cuh"
#include "rng.cuh"

#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>

using namespace wharf;

namespace
{
thread_local std::string g_last_error;

int fail(int code, const std::string& msg)
{
    g_last_error = msg;
    return code;
}

template <class F>
int guarded(F&& f)
{
    try
    {
        return f();
    }
    catch (const CudaError& e)
    {
        cudaGetLastError();
        return fail(e.code == cudaErrorMemoryAllocation ? WMH_ERR_OOM : WMH_ERR_CUDA, e.what());
    }
    catch (const std::bad_alloc&)
    {
        return fail(WMH_ERR_OOM, "host allocation failed");
    }
    catch (const std::exception& e)
    {
        return fail(WMH_ERR_INVALID, e.what());
    }
}
}  // namespace

int wharf::report_failure(int code, const std::string& msg) { return fail(code, msg); }

namespace
{
constexpr int kMaxDevices = 64;
cudaMemPool_t g_pools[kMaxDevices] = {};
std::mutex g_pools_mutex;

// The library's pool on `device` (created on first use; lives as long as the process).
cudaError_t library_pool(int device, cudaMemPool_t* out)
{
    if (device < 0 || device >= kMaxDevices) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_pools_mutex);
    if (!g_pools[device])
    {
        cudaMemPoolProps props{};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        cudaMemPool_t pool = nullptr;
        cudaError_t e = cudaMemPoolCreate(&pool, &props);
        if (e != cudaSuccess) return e;
        unsigned long long threshold = ~0ull;              // keep freed blocks: the next batch / engine takes them again
        e = cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &threshold);
        if (e != cudaSuccess) { cudaMemPoolDestroy(pool); return e; }
        g_pools[device] = pool;
    }
    *out = g_pools[device];
    return cudaSuccess;
}
}  // namespace

cudaError_t wharf::pool_malloc_async(void** ptr, size_t bytes, cudaStream_t stream)
{
    int device = 0;
    cudaError_t e = cudaGetDevice(&device);
    if (e != cudaSuccess) return e;
    cudaMemPool_t pool = nullptr;
    e = library_pool(device, &pool);
    if (e != cudaSuccess) return e;
    return cudaMallocFromPoolAsync(ptr, bytes, pool, stream);
}

namespace
{
void peer_release(PeerLink& link)
{
    for (u32 g = 0; g < kPeerMaxWorld; g++)
    {
        if (!link.ptrs.buf[g]) continue;
        if (g == link.rank) { if (link.created) cudaFree(link.ptrs.buf[g]); }
        else cudaIpcCloseMemHandle(link.ptrs.buf[g]);
        link.ptrs.buf[g] = nullptr;
    }
    cudaGetLastError();
    link.created = link.attached = false;
    link.bytes = 0;
}
}  // namespace

// Declared first in wmh_engine, so destroyed last: every buffer of the engine frees itself on this stream.
struct OwnedStream
{
    cudaStream_t handle = nullptr;
    int device = 0;
    ~OwnedStream()
    {
        if (!handle) return;
        cudaSetDevice(device);
        cudaStreamDestroy(handle);           // pending frees complete first: the driver releases the stream afterwards
    }
};

struct wmh_engine
{
    OwnedStream owned;
    wmh_config cfg{};
    cudaStream_t stream = nullptr;        // = owned.handle
    cudaStream_t copy_stream = nullptr;   // corpus export overlapped with generation (generate_and_export)
    int sm_count = 148;
    u64 launches = 0, library_calls = 0;

    DeviceGraph g;
    DeviceBuffer<u64> draw;            // X[wpv][L]
    std::vector<u64> host_draw;        // the same table on the host (kernel arguments of the coalesced path)
    u32 draw_wpv = 0, draw_len = 0;

    DeviceBuffer<u32> walks;           // [L][w_local]
    bool have_walks = false;
    u64 w_local = 0;
    ShardMap map;                      // which walks this rank owns (fixed when the corpus is generated)
    u32 walks_L = 0, walks_wpv = 0;

    DeviceIndex index;
    DeviceBuffer<u64> codes;           // paired corpus (compressed_walks.h), valid while index.valid && codes_valid
    bool codes_valid = false;

    DeviceBuffer<u32> bitmap;          // touched vertices of the current batch
    DeviceBuffer<u32> min_pos;         // w_local
    DeviceBuffer<u32> aff_ids, aff_pos;
    u64 n_affected = 0;
    DeviceBuffer<u32> touched;         // distinct sources of the last batch
    u64 n_touched = 0;

    DeviceBuffer<unsigned long long> d_steps;
    u64 last_steps = 0;
    double timers[5] = {0, 0, 0, 0, 0};
    double last_walk_kernel_ms = 0;
    double last_index_ms = 0;          // index build (generate) / index maintenance (batch) of the last call, device time
    u32 last_search = 0;               // how the last batch found its affected walks: WMH_SEARCH_INDEX / WMH_SEARCH_SCAN, 0 = none
    cudaEvent_t ev_index[2] = {nullptr, nullptr};
    cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr};
    DeviceBuffer<u32> small;           // misc tiny results
    PeerLink peer;                     // exchange buffers of the box's ranks (wmh_peer_*)
    CoalescedPlan* coalesced_plan = nullptr;   // scratch + CUDA graph of the coalesced deterministic pass

    LaunchCtx lc() { return LaunchCtx{stream, sm_count, &launches, &library_calls}; }
    void bind() const
    {
        WHARF_CUDA(cudaSetDevice(cfg.device));
        g_alloc_stream = stream;          // DeviceBuffer allocations of this thread are ordered on the engine stream
        g_alloc_async = stream != nullptr;
    }

    ~wmh_engine()
    {
        cudaSetDevice(cfg.device);
        if (stream) cudaStreamSynchronize(stream);
        if (copy_stream) { cudaStreamSynchronize(copy_stream); cudaStreamDestroy(copy_stream); }
        peer_release(peer);
        g_alloc_stream = stream;
        coalesced_plan_release(coalesced_plan);
        for (auto& e : ev) if (e) cudaEventDestroy(e);
        for (auto& e : ev_index) if (e) cudaEventDestroy(e);
        // buffers release in their destructors (stream-ordered frees); `owned` goes last and destroys the stream
    }
};

namespace
{
void validate_config(const wmh_config& c)
{
    if (c.walks_per_vertex == 0 || c.walks_per_vertex > 255) throw std::invalid_argument("walks_per_vertex must be in 1..255 (uint8_t in the reference)");
    if (c.walk_length == 0 || c.walk_length > 255) throw std::invalid_argument("walk_length must be in 1..255 (uint8_t in the reference)");
    if (c.model > WMH_NODE2VEC) throw std::invalid_argument("unrecognized random walking model");          // wharfmh.h:269-271
    if (c.init_strategy > WMH_INIT_WEIGHT) throw std::invalid_argument("unrecognized sampler init strategy");
    if (c.search_mode > WMH_SEARCH_SCAN) throw std::invalid_argument("unrecognized search mode");
    if (c.generate_mode > WMH_GENERATE_COALESCED) throw std::invalid_argument("unrecognized generate mode");
    if (c.world == 0 || c.rank >= c.world) throw std::invalid_argument("rank must be < world, world >= 1");
    if (c.model == WMH_NODE2VEC && !c.deterministic && (!(c.param_p > 0) || !(c.param_q > 0))) throw std::invalid_argument("node2vec p and q must be positive");
}

std::unique_ptr<wmh_engine> make_engine(const wmh_config* cfg)
{
    if (!cfg) throw std::invalid_argument("null config");
    validate_config(*cfg);
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
    {
        cudaGetLastError();
        throw CudaError(e == cudaSuccess ? cudaErrorNoDevice : e, "no CUDA device available: libwharfmh_b200 has no CPU path");
    }
    if (cfg->device < 0 || cfg->device >= count) throw std::invalid_argument("device ordinal out of range");
    Trace tr(nullptr);
    auto eng = std::make_unique<wmh_engine>();
    eng->cfg = *cfg;
    eng->bind();
    int major = 0, sms = 0;                 // attribute queries: cudaGetDeviceProperties costs ~3 ms per call
    WHARF_CUDA(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, cfg->device));
    WHARF_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, cfg->device));
    if (major < 10) throw CudaError(cudaErrorInvalidDevice, "device " + std::to_string(cfg->device) + " is not sm_100-class; this library carries sm_100a code only");
    eng->sm_count = sms;
    tr.mark("create: device properties");
    // Process-wide device state is left alone.  Tuning aid only: WHARF_L2_FETCH=32|64|128 sets
    // cudaLimitMaxL2FetchGranularity (measured on B200: the hint changes nothing for this path, profiles/r01c_*).
    if (const char* env = std::getenv("WHARF_L2_FETCH"))
    {
        if (const size_t granularity = (size_t) std::atoi(env)) cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, granularity);
        cudaGetLastError();
    }
    WHARF_CUDA(cudaStreamCreateWithFlags(&eng->owned.handle, cudaStreamNonBlocking));
    eng->owned.device = cfg->device;
    eng->stream = eng->owned.handle;
    eng->bind();
    for (auto& ev : eng->ev) WHARF_CUDA(cudaEventCreate(&ev));
    for (auto& ev : eng->ev_index) WHARF_CUDA(cudaEventCreate(&ev));
    {
        cudaMemPool_t pool = nullptr;                    // the library's own pool (pool_malloc_async), not the default one
        WHARF_CUDA(library_pool(cfg->device, &pool));
    }
    eng->d_steps.allocate(1);
    eng->small.allocate(16);
    tr.mark("create: stream+events+pool");
    return eng;
}

void finish_graph(wmh_engine& e, u64 n, u64 m, const u64* d_offsets, const u32* d_edges)
{
    if (n == 0) throw std::invalid_argument("graph needs at least one vertex");
    if (n >= 0xFFFFFFFEull) throw std::invalid_argument("vertex ids must fit 32 bits");
    if (m >= (1ull << 32)) throw std::invalid_argument("more than 2^32 directed edges are not supported by the u32 row descriptors");
    auto lc = e.lc();
    Trace tr(e.stream);
    e.g.n = n;
    e.g.m = m;
    e.g.desc.allocate(n);
    u64 capacity = m + m / 2 + (1u << 20);
    if (capacity >= (1ull << 32)) capacity = (1ull << 32) - 1;
    e.g.pool.allocate(capacity);
    e.g.version.allocate(n);
    WHARF_CUDA(cudaMemsetAsync(e.g.version.ptr, 0, n * sizeof(u32), e.stream));
    if (d_offsets)
    {
        // the constructor's contract (wharfmh.h:58-103): a bad id or an unsorted row would send every walk kernel out of
        // bounds or silently break the sorted-set merges and searches - refuse it before anything can walk
        WHARF_CUDA(cudaMemsetAsync(e.small.ptr, 0, 4, e.stream));
        graph_check_csr(lc, d_offsets, n, m, d_edges, e.small.ptr);
        u32 bad = 0;
        WHARF_CUDA(cudaMemcpyAsync(&bad, e.small.ptr, 4, cudaMemcpyDeviceToHost, e.stream));
        WHARF_CUDA(cudaStreamSynchronize(e.stream));
        if (bad & 1u) throw std::invalid_argument("CSR offsets must start at 0, be non-decreasing and stay <= m");
        if (bad & 2u) throw std::invalid_argument("CSR names a neighbour id >= number_of_vertices()");
        if (bad & 4u) throw std::invalid_argument("CSR rows must be strictly ascending (sorted, no duplicates)");
    }
    if (m) WHARF_CUDA(cudaMemcpyAsync(e.g.pool.ptr, d_edges, m * sizeof(u32), cudaMemcpyDeviceToDevice, e.stream));
    if (d_offsets) graph_build_desc(lc, d_offsets, n, m, e.g.desc.ptr);
    else WHARF_CUDA(cudaMemsetAsync(e.g.desc.ptr, 0, n * sizeof(RowDesc), e.stream));
    e.g.pool_used = m;
    e.bitmap.allocate((n + 31) / 32);
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
    tr.mark("create: graph buffers+copy");
    // Tuning aid only: WHARF_L2_PERSIST=<MiB> asks the L2 to keep that much of the row descriptors resident (access-policy
    // window on the engine's stream) while the neighbour pool streams past them.  Changes process-wide device state
    // (cudaLimitPersistingL2CacheSize), hence off unless asked for; measured in profiles/r02c_l2_persist_ab.txt.
    if (const char* env = std::getenv("WHARF_L2_PERSIST"))
    {
        const size_t want = (size_t) std::atoi(env) << 20;
        int max_persist = 0, max_window = 0;
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, e.cfg.device);
        cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, e.cfg.device);
        const size_t carve = std::min(want, (size_t) max_persist);
        if (carve)
        {
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, carve);
            cudaStreamAttrValue attr{};
            attr.accessPolicyWindow.base_ptr = e.g.desc.ptr;
            attr.accessPolicyWindow.num_bytes = std::min(n * sizeof(RowDesc), (size_t) max_window);
            attr.accessPolicyWindow.hitRatio = (float) std::min(1.0, (double) carve / (double) attr.accessPolicyWindow.num_bytes);
            attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            cudaStreamSetAttribute(e.stream, cudaStreamAttributeAccessPolicyWindow, &attr);
            if (std::getenv("WHARF_TRACE"))
                std::fprintf(stderr, "[wharf] L2 persist: carve %zu MiB (max %d MiB), window %zu MiB, hit ratio %.2f: %s\n", carve >> 20, max_persist >> 20,
                             attr.accessPolicyWindow.num_bytes >> 20, attr.accessPolicyWindow.hitRatio, cudaGetErrorString(cudaGetLastError()));
            cudaGetLastError();
        }
    }
}

// The partition of the walks the engine would use for a corpus generated now.
ShardMap config_map(const wmh_engine& e) { return make_shard_map(e.g.n, e.cfg.walks_per_vertex, e.cfg.rank, e.cfg.world); }

u64 local_walk_capacity(const wmh_engine* e) { return config_map(*e).local_walks(); }

WalkParams walk_params(wmh_engine& e)
{
    WalkParams wp;
    wp.desc = e.g.desc.ptr;
    wp.pool = e.g.pool.ptr;
    wp.version = e.g.version.ptr;
    wp.pool_rw = e.g.tail ? e.g.pool.ptr : nullptr;
    wp.shift = e.g.tail;
    wp.tri = e.g.tail ? e.g.tri : 0;
    wp.draw = e.draw.ptr;
    wp.walks = e.walks.ptr;
    wp.n = e.g.n;
    wp.w_local = e.w_local;
    wp.map = e.map;
    wp.wpv = e.walks_wpv;
    wp.L = e.walks_L;
    wp.model = e.cfg.model;
    wp.deterministic = e.cfg.deterministic;
    wp.init_strategy = e.cfg.init_strategy;
    wp.p = e.cfg.param_p;
    wp.q = e.cfg.param_q;
    wp.seed = e.cfg.seed;
    return wp;
}

bool needs_memo(const wmh_engine& e) { return !e.cfg.deterministic && e.cfg.model == WMH_NODE2VEC; }

// MH node2vec keeps its sampler caches and neighbour filter in aux tails behind the rows (kernels_graph.cu): switch
// the pool to that layout the first time it is needed.
void ensure_tails(wmh_engine& e)
{
    if (!needs_memo(e) || e.g.tail) return;
    auto lc = e.lc();
    graph_enable_tails(lc, e.g);
}

// Exchange buffer a generate pass needs: MH node2vec shares the sampler init values, 2 bytes per edge.
u64 peer_bytes_needed(const wmh_engine& e)
{
    if (e.cfg.world <= 1 || !needs_memo(e)) return 0;
    return kPeerFlagBytes + 2 * ((e.g.m + 63) / 64 * 64 + 64);
}

bool peer_usable(const wmh_engine& e)
{
    const u64 need = peer_bytes_needed(e);
    return e.peer.attached && need != 0 && e.peer.bytes >= need && e.peer.world == e.cfg.world && e.peer.rank == e.cfg.rank;
}

// A barrier that gave up waiting leaves a mark in the own flag page: the peers are out of step (or gone).
void peer_check(wmh_engine& e)
{
    u32 flag = 0;
    WHARF_CUDA(cudaMemcpyAsync(&flag, e.peer.own() + kPeerErrorOffset, 4, cudaMemcpyDeviceToHost, e.stream));
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
    if (!flag) return;
    WHARF_CUDA(cudaMemsetAsync(e.peer.own() + kPeerErrorOffset, 0, 4, e.stream));
    throw CudaError(cudaErrorLaunchTimeout, "peer barrier timed out after " + std::to_string(WMH_PEER_TIMEOUT_MS) +
                                            " ms: the ranks did not issue the same sequence of generate calls");
}

// Deterministic bulk generation: the coalesced schedule pays ~6 small launches per position, the per-walker kernel
// two random loads per step; the latter wins while descriptors + neighbours stay in the 126 MB L2.
bool use_coalesced(const wmh_engine& e)
{
    if (!e.cfg.deterministic || e.cfg.generate_mode == WMH_GENERATE_PER_WALKER) return false;
    if (e.cfg.generate_mode == WMH_GENERATE_COALESCED) return true;
    return e.g.n * sizeof(RowDesc) + e.g.m * sizeof(u32) > (96ull << 20);
}

// Bulk generation with the corpus streamed to the host while it is produced: the walkers are generated in
// ranges (replica by replica on the coalesced schedule); as soon as a range is complete the copy stream
// transposes it to walk-major rows in a staging buffer and copies it out, overlapping the next range's kernel
// with the PCIe transfer.  Same corpus and step count as generate() + wmh_walks_copy.
void generate_and_export(wmh_engine& e, u32* host_out)
{
    auto lc = e.lc();
    const u32 L = e.walks_L;
    if (!e.copy_stream)
    {
        // highest priority: the transposes must get SM slots while the next range's walk kernel fills the machine
        int least = 0, greatest = 0;
        WHARF_CUDA(cudaDeviceGetStreamPriorityRange(&least, &greatest));
        WHARF_CUDA(cudaStreamCreateWithPriority(&e.copy_stream, cudaStreamNonBlocking, greatest));
    }
    LaunchCtx copy_lc{e.copy_stream, e.sm_count, &e.launches, &e.library_calls};
    const u64 max_stage_walks = std::max<u64>(256, (128ull << 20) / L);                 // 512 MiB of staging at most
    u64 piece = std::min(max_stage_walks, std::max<u64>(256, (e.w_local + 7) / 8));
    piece = (piece + 255) / 256 * 256;
    Scratch<u32> stage(e.copy_stream, piece * L);
    std::vector<cudaEvent_t> events;
    auto export_range = [&](u64 first, u64 count) {                                      // [first, first + count) is complete on e.stream
        cudaEvent_t ev;
        WHARF_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        events.push_back(ev);
        WHARF_CUDA(cudaEventRecord(ev, e.stream));
        WHARF_CUDA(cudaStreamWaitEvent(e.copy_stream, ev, 0));
        for (u64 done = 0; done < count; done += piece)
        {
            const u64 c = std::min(piece, count - done);
            walks_transpose_rows(copy_lc, e.walks.ptr, e.w_local, L, first + done, c, stage.ptr);
            WHARF_CUDA(cudaMemcpyAsync(host_out + (first + done) * L, stage.ptr, c * L * sizeof(u32), cudaMemcpyDeviceToHost, e.copy_stream));
        }
    };
    const WalkParams wp = walk_params(e);
    if (use_coalesced(e))
    {
        walks_generate_coalesced(lc, wp, e.host_draw.data(), e.d_steps.ptr, [&](u64 first, u64 count) { export_range(first, count); }, &e.coalesced_plan);
    }
    else
    {
        // generation ranges: an eighth of the corpus each (a range is several full waves of walker blocks; its copy
        // starts when its kernel ends), copied out in staging-sized pieces
        const u64 range = std::max(piece, ((e.w_local + 7) / 8 + 255) / 256 * 256);
        for (u64 first = 0; first < e.w_local; first += range)
        {
            const u64 c = std::min(range, e.w_local - first);
            walks_generate_range(lc, wp, first, c, e.d_steps.ptr);
            export_range(first, c);
        }
    }
    WHARF_CUDA(cudaStreamSynchronize(e.copy_stream));
    for (cudaEvent_t ev : events) cudaEventDestroy(ev);
}

// index_build between two events on the engine stream: e.last_index_ms.
void timed_index_build(wmh_engine& e)
{
    auto lc = e.lc();
    WHARF_CUDA(cudaEventRecord(e.ev_index[0], e.stream));
    index_build(lc, e.walks.ptr, e.w_local, e.walks_L, e.g.n, e.index);
    WHARF_CUDA(cudaEventRecord(e.ev_index[1], e.stream));
    WHARF_CUDA(cudaEventSynchronize(e.ev_index[1]));
    float ms = 0;
    WHARF_CUDA(cudaEventElapsedTime(&ms, e.ev_index[0], e.ev_index[1]));
    e.last_index_ms = ms;
}

void generate(wmh_engine& e, u32* host_out = nullptr)
{
    e.bind();
    validate_config(e.cfg);
    auto lc = e.lc();
    const u32 wpv = e.cfg.walks_per_vertex, L = e.cfg.walk_length;
    e.map = config_map(e);
    e.w_local = e.map.local_walks();
    if (e.w_local >= (1ull << 32)) throw std::invalid_argument("more than 2^32 walks per rank are not supported");
    e.walks_L = L;
    e.walks_wpv = wpv;
    e.have_walks = false;
    e.index.valid = false;
    e.codes_valid = false;
    e.n_affected = 0;
    Trace tr(e.stream);
    e.walks.ensure(std::max<u64>(1, e.w_local * L));
    e.min_pos.ensure(std::max<u64>(1, e.w_local));
    tr.mark("generate: corpus buffers");
    if (e.draw_wpv != wpv || e.draw_len != L)
    {
        e.host_draw = make_draw_table(wpv, L);
        const std::vector<u64>& table = e.host_draw;
        e.draw.allocate(table.size());
        WHARF_CUDA(cudaMemcpyAsync(e.draw.ptr, table.data(), table.size() * sizeof(u64), cudaMemcpyHostToDevice, e.stream));
        WHARF_CUDA(cudaStreamSynchronize(e.stream));
        e.draw_wpv = wpv;
        e.draw_len = L;
    }
    if (needs_memo(e)) ensure_tails(e);                  // one-time switch of the pool layout
    tr.mark("generate: draws+tails");
    WHARF_CUDA(cudaMemsetAsync(e.d_steps.ptr, 0, 8, e.stream));
    WHARF_CUDA(cudaEventRecord(e.ev[0], e.stream));
    bool exchanged = false;
    if (needs_memo(e))
    {
        // a fresh corpus starts from empty sampler managers (wharfmh.h:41,83,90): every cache is emptied, or overwritten
        // by the value its first touch would compute.  The bulk kernel pays for every state of the graph, so it is used
        // when the corpus is large enough to touch most of them anyway (the quantity compared is the same on every
        // rank: the ranks of a box must take the same branch, they meet at a barrier): with peers each rank computes
        // 1 / world of the states; on one rank it replaces first-touch inits scattered through the walk kernel
        // (scale-20: 30.0 -> 29.3 ms).  WHARF_EAGER_INIT=0 keeps the single-rank pass lazy.
        const u64 steps_per_rank = e.g.n * (u64) wpv * L / e.cfg.world;
        const char* eager_env = std::getenv("WHARF_EAGER_INIT");
        if (peer_usable(e) && steps_per_rank >= 2 * e.g.m)
        {
            n2v_share_sampler_init(lc, walk_params(e), e.g.m, e.peer);
            exchanged = true;
        }
        else if (e.cfg.world == 1 && steps_per_rank >= 8 * e.g.m && !(eager_env && std::atoi(eager_env) == 0))
        {
            Scratch<u8> flags(e.stream, kPeerFlagBytes);                 // a world of one: the barrier only needs its flag page
            WHARF_CUDA(cudaMemsetAsync(flags.ptr, 0, kPeerFlagBytes, e.stream));
            PeerLink solo;
            solo.ptrs.buf[0] = flags.ptr;
            solo.bytes = kPeerFlagBytes;
            n2v_share_sampler_init(lc, walk_params(e), e.g.m, solo);
        }
        else graph_aux_clear_caches(lc, e.g);
    }
    if (host_out) generate_and_export(e, host_out);
    else if (use_coalesced(e)) walks_generate_coalesced(lc, walk_params(e), e.host_draw.data(), e.d_steps.ptr, nullptr, &e.coalesced_plan);
    else walks_generate(lc, walk_params(e), e.d_steps.ptr);
    WHARF_CUDA(cudaEventRecord(e.ev[1], e.stream));
    unsigned long long steps = 0;
    WHARF_CUDA(cudaMemcpyAsync(&steps, e.d_steps.ptr, 8, cudaMemcpyDeviceToHost, e.stream));
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
    if (host_out) WHARF_CUDA(cudaStreamSynchronize(e.copy_stream));
    if (exchanged) peer_check(e);
    float ms = 0;
    WHARF_CUDA(cudaEventElapsedTime(&ms, e.ev[0], e.ev[1]));
    e.last_walk_kernel_ms = ms;
    e.timers[WMH_TIMER_GENERATE] = ms * 1e-3;
    e.last_steps = steps;
    e.have_walks = true;
    tr.mark("generate: walk (+export)");
    e.last_index_ms = 0;
    if (e.cfg.search_mode == WMH_SEARCH_INDEX) timed_index_build(e);     // the reference builds it inside this call (wharfmh.h:312-354)
}

// Shared body of insert_edges_batch / delete_edges_batch (wharfmh.h:439-576 / 588-726).
void update(wmh_engine& e, u64 count, const u32* d_pairs, u32* d_sorted_out, bool is_delete, bool remove_dups, bool apply_walk_updates)
{
    auto lc = e.lc();
    e.n_affected = 0;
    e.n_touched = 0;
    e.last_steps = 0;
    e.last_walk_kernel_ms = 0;
    e.last_index_ms = 0;
    e.last_search = 0;
    if (count == 0) return;

    WHARF_CUDA(cudaMemsetAsync(e.small.ptr, 0, 4, e.stream));
    graph_check_ids(lc, d_pairs, count, e.g.n, e.small.ptr);
    u32 bad = 0;
    WHARF_CUDA(cudaMemcpyAsync(&bad, e.small.ptr, 4, cudaMemcpyDeviceToHost, e.stream));      // read at the first sync below: nothing is changed before

    // ---- graph update (graph_update_time_on_{insert,delete}, wharfmh.h:542-544 / 692-694) ----
    WHARF_CUDA(cudaEventRecord(e.ev[0], e.stream));
    PreparedBatch batch;
    graph_prepare_batch(lc, d_pairs, count, remove_dups, e.g.n, batch, d_sorted_out);                // sorts and dedups a copy; synchronises
    if (bad) throw std::invalid_argument("edge batch names a vertex id >= number_of_vertices()");   // would change walk_id % n (wharfmh.h:254,368)
    WHARF_CUDA(cudaMemsetAsync(e.bitmap.ptr, 0, e.bitmap.bytes(), e.stream));
    if (needs_memo(e) && e.have_walks) ensure_tails(e);
    graph_apply_batch(lc, e.g, batch, is_delete, e.bitmap.ptr);
    e.codes_valid = false;
    e.n_touched = batch.n_sources;
    if (batch.n_sources)
    {
        e.touched.ensure(batch.n_sources);
        WHARF_CUDA(cudaMemcpyAsync(e.touched.ptr, batch.sources.ptr, batch.n_sources * sizeof(u32), cudaMemcpyDeviceToDevice, e.stream));
    }
    WHARF_CUDA(cudaEventRecord(e.ev[1], e.stream));

    // ---- walk update (walk_update_time_on_{insert,delete}, wharfmh.h:546-549 / 696-699) ----
    // Ranks of a box: the batch emptied the sampler caches of every state whose cur is a source - on RMAT batches that is most
    // states, because the hubs are sources - and every rank would re-derive the same init values lazily inside its re-walk.
    // They are a pure function of the replicated graph: compute each once per box, as generate() does.  The rule uses only
    // quantities that are equal on every rank (the ranks meet at a barrier inside).
    bool exchanged = false;
    if (needs_memo(e) && e.have_walks && apply_walk_updates && batch.n_sources && peer_usable(e) && e.g.last_rewritten * 4 >= e.g.m)
    {
        n2v_share_sampler_init(lc, walk_params(e), e.g.m, e.peer);
        exchanged = true;
    }
    float walk_kernel_ms = 0;
    if (e.have_walks && batch.n_sources && e.w_local)
    {
        // 1. map of changes: earliest position of every walk on a touched vertex (wharfmh.h:519-537)
        bool use_index = false;
        if (e.cfg.search_mode == WMH_SEARCH_INDEX)
        {
            if (!e.index.valid) timed_index_build(e);
            use_index = true;
        }
        else if (e.cfg.search_mode == WMH_SEARCH_AUTO && e.index.valid)
        {
            const u64 visited = index_count_touched(lc, e.index, e.g.n, batch.sources.ptr, batch.n_sources);
            use_index = visited * 8 < e.w_local * e.walks_L;    // a random index entry costs about 8 streamed cells
        }
        e.last_search = use_index ? WMH_SEARCH_INDEX : WMH_SEARCH_SCAN;
        if (use_index)
        {
            fill_u32(lc, e.min_pos.ptr, e.w_local, kNoPos);
            index_search(lc, e.index, e.g.n, batch.sources.ptr, batch.n_sources, e.w_local, e.min_pos.ptr);
        }
        else
        {
            walks_scan_touched(lc, e.walks.ptr, e.w_local, e.walks_L, e.bitmap.ptr, e.min_pos.ptr);
        }
        e.aff_ids.ensure(e.w_local);
        e.aff_pos.ensure(e.w_local);
        e.n_affected = walks_compact_affected(lc, e.min_pos.ptr, e.w_local, e.aff_ids.ptr, e.aff_pos.ptr);

        // 2. re-walk every affected walk from its earliest affected position (wharfmh.h:761-859)
        if (apply_walk_updates && e.n_affected)
        {
            WHARF_CUDA(cudaMemsetAsync(e.d_steps.ptr, 0, 8, e.stream));
            WHARF_CUDA(cudaEventRecord(e.ev[3], e.stream));
            walks_rewalk(lc, walk_params(e), e.aff_ids.ptr, e.aff_pos.ptr, e.n_affected, e.d_steps.ptr);
            WHARF_CUDA(cudaEventRecord(e.ev[4], e.stream));
            unsigned long long steps = 0;
            WHARF_CUDA(cudaMemcpyAsync(&steps, e.d_steps.ptr, 8, cudaMemcpyDeviceToHost, e.stream));
            WHARF_CUDA(cudaStreamSynchronize(e.stream));
            e.last_steps = steps;
            WHARF_CUDA(cudaEventElapsedTime(&walk_kernel_ms, e.ev[3], e.ev[4]));
            // 3. index maintenance (wharfmh.h:861-922): the sorted index is re-derived from the patched corpus
            e.index.valid = false;
            if (e.cfg.search_mode == WMH_SEARCH_INDEX) timed_index_build(e);
        }
    }
    WHARF_CUDA(cudaEventRecord(e.ev[2], e.stream));
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
    if (exchanged) peer_check(e);
    float graph_ms = 0, walk_ms = 0;
    WHARF_CUDA(cudaEventElapsedTime(&graph_ms, e.ev[0], e.ev[1]));
    WHARF_CUDA(cudaEventElapsedTime(&walk_ms, e.ev[1], e.ev[2]));
    e.timers[is_delete ? WMH_TIMER_GRAPH_UPDATE_DELETE : WMH_TIMER_GRAPH_UPDATE_INSERT] += graph_ms * 1e-3;
    e.timers[is_delete ? WMH_TIMER_WALK_UPDATE_DELETE : WMH_TIMER_WALK_UPDATE_INSERT] += walk_ms * 1e-3;
    e.last_walk_kernel_ms = walk_kernel_ms;
}

int update_host(wmh_engine* e, u64 m, u32* edges, int sorted, int remove_dups, int apply, u64* n_affected, bool is_delete)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (m && !edges) return fail(WMH_ERR_INVALID, "null edge batch");
    return guarded([&] {
        e->bind();
        Scratch<u32> d_pairs(e->stream, 2 * m), d_sorted;
        if (m) WHARF_CUDA(cudaMemcpyAsync(d_pairs.ptr, edges, 2 * m * sizeof(u32), cudaMemcpyHostToDevice, e->stream));
        if (!sorted && m) d_sorted.allocate(e->stream, 2 * m);
        update(*e, m, d_pairs.ptr, d_sorted.ptr, is_delete, remove_dups != 0, apply != 0);
        if (!sorted && m)
        {
            // the reference sorts the caller's array in place (wharfmh.h:452 -> :1056)
            WHARF_CUDA(cudaMemcpyAsync(edges, d_sorted.ptr, 2 * m * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
            WHARF_CUDA(cudaStreamSynchronize(e->stream));
        }
        if (n_affected) *n_affected = e->n_affected;
        return (int) WMH_OK;
    });
}

int update_device(wmh_engine* e, u64 m, const u32* d_edges, int remove_dups, int apply, u64* n_affected, bool is_delete)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (m && !d_edges) return fail(WMH_ERR_INVALID, "null edge batch");
    return guarded([&] {
        e->bind();
        update(*e, m, d_edges, nullptr, is_delete, remove_dups != 0, apply != 0);
        if (n_affected) *n_affected = e->n_affected;
        return (int) WMH_OK;
    });
}

i64 global_to_local(const wmh_engine& e, u64 walk_id)
{
    if (!e.have_walks || walk_id >= e.g.n * e.walks_wpv) return -1;
    return e.map.local_of(walk_id);
}

u64 local_to_global(const wmh_engine& e, u64 local) { return e.map.global_id(local); }

void fetch_walk(wmh_engine& e, u64 local, std::vector<u32>& row)
{
    auto lc = e.lc();
    const u32 L = e.walks_L;
    Scratch<u32> d_id(e.stream, 1), d_row(e.stream, L);
    const u32 id = (u32) local;
    WHARF_CUDA(cudaMemcpyAsync(d_id.ptr, &id, 4, cudaMemcpyHostToDevice, e.stream));
    walks_gather_rows(lc, e.walks.ptr, e.w_local, L, d_id.ptr, 1, d_row.ptr);
    row.resize(L);
    WHARF_CUDA(cudaMemcpyAsync(row.data(), d_row.ptr, L * sizeof(u32), cudaMemcpyDeviceToHost, e.stream));
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
}

void ensure_index(wmh_engine& e)
{
    if (!e.have_walks) throw std::logic_error("walks have not been generated");
    if (!e.index.valid)
    {
        auto lc = e.lc();
        index_build(lc, e.walks.ptr, e.w_local, e.walks_L, e.g.n, e.index);
        e.codes_valid = false;
    }
}

void ensure_codes(wmh_engine& e)
{
    ensure_index(e);
    if (e.codes_valid) return;
    if (e.g.n * e.walks_wpv * (u64) e.walks_L >= (1ull << 32)) throw std::domain_error("paired corpus needs walks * walk_length < 2^32 (u64 Szudzik domain)");
    if (e.index.chunks != 1) throw std::domain_error("paired corpus needs a one-chunk index (local walks * walk length <= 2^31 cells)");
    auto lc = e.lc();
    e.codes.ensure(std::max<u64>(1, e.index.entries));
    compressed_walks_encode(lc, e.index, e.walks.ptr, e.w_local, e.walks_L, e.g.n, e.map, e.codes.ptr);
    WHARF_CUDA(cudaStreamSynchronize(e.stream));
    e.codes_valid = true;
}

struct DefaultCtx
{
    u64 launches = 0, lib = 0;
    LaunchCtx lc;
    explicit DefaultCtx(int device)
    {
        int count = 0;
        if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count)
        {
            cudaGetLastError();
            throw CudaError(cudaErrorNoDevice, "no usable CUDA device: libwharfmh_b200 has no CPU path");
        }
        WHARF_CUDA(cudaSetDevice(device));
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        lc = LaunchCtx{nullptr, sms, &launches, &lib};
    }
};
}  // namespace

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" {

void wmh_default_config(wmh_config* cfg)
{
    if (!cfg) return;
    std::memset(cfg, 0, sizeof(*cfg));
    cfg->walks_per_vertex = 10;            // config/globals.h:7
    cfg->walk_length = 80;                 // config/globals.h:10
    cfg->model = WMH_NODE2VEC;             // config/globals.h:14
    cfg->param_p = 4.0f;                   // config/globals.h:17
    cfg->param_q = 1.0f;                   // config/globals.h:20
    cfg->init_strategy = WMH_INIT_WEIGHT;  // config/globals.h:23
    cfg->deterministic = 1;                // config/globals.h:29
    cfg->search_mode = WMH_SEARCH_AUTO;
    cfg->seed = 0;
    cfg->device = 0;
    cfg->rank = 0;
    cfg->world = 1;
}

const char* wmh_last_error(void) { return g_last_error.c_str(); }
const char* wmh_version(void) { return "wharfmh_b200 0.1 (sm_100a)"; }

int wmh_create(wmh_engine** out, uint64_t n, uint64_t m, const uint64_t* offsets, const uint32_t* edges, const wmh_config* cfg)
{
    if (!out) return fail(WMH_ERR_INVALID, "null out");
    *out = nullptr;
    if (!offsets || (m && !edges)) return fail(WMH_ERR_INVALID, "null CSR arrays");
    return guarded([&] {
        auto e = make_engine(cfg);
        Scratch<u64> d_off(e->stream, n);
        Scratch<u32> d_edges(e->stream, m);
        WHARF_CUDA(cudaMemcpyAsync(d_off.ptr, offsets, n * sizeof(u64), cudaMemcpyHostToDevice, e->stream));
        if (m) WHARF_CUDA(cudaMemcpyAsync(d_edges.ptr, edges, m * sizeof(u32), cudaMemcpyHostToDevice, e->stream));
        finish_graph(*e, n, m, d_off.ptr, d_edges.ptr);
        *out = e.release();
        return (int) WMH_OK;
    });
}

int wmh_create_device(wmh_engine** out, uint64_t n, uint64_t m, const uint64_t* d_offsets, const uint32_t* d_edges, const wmh_config* cfg)
{
    if (!out) return fail(WMH_ERR_INVALID, "null out");
    *out = nullptr;
    if (!d_offsets || (m && !d_edges)) return fail(WMH_ERR_INVALID, "null CSR arrays");
    return guarded([&] {
        auto e = make_engine(cfg);
        WHARF_CUDA(cudaDeviceSynchronize());     // the caller's producer streams are unknown to us
        finish_graph(*e, n, m, d_offsets, d_edges);
        *out = e.release();
        return (int) WMH_OK;
    });
}

int wmh_create_empty(wmh_engine** out, uint64_t n, const wmh_config* cfg)
{
    if (!out) return fail(WMH_ERR_INVALID, "null out");
    *out = nullptr;
    return guarded([&] {
        auto e = make_engine(cfg);
        finish_graph(*e, n, 0, nullptr, nullptr);
        *out = e.release();
        return (int) WMH_OK;
    });
}

void wmh_destroy(wmh_engine* e)
{
    if (!e) return;
    delete e;                             // the engine owns its stream (OwnedStream)
    g_alloc_stream = nullptr;
    g_alloc_async = false;
}

int wmh_destroy_index(wmh_engine* e)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    return guarded([&] {
        e->bind();
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        e->have_walks = false;
        e->index.valid = false;
        e->index.cells.release();
        e->index.ptr.release();
        e->codes.release();
        e->codes_valid = false;
        e->walks.release();
        coalesced_plan_release(e->coalesced_plan);
        e->min_pos.release();
        e->aff_ids.release();
        e->aff_pos.release();
        e->n_affected = 0;
        return (int) WMH_OK;
    });
}

int wmh_set_config(wmh_engine* e, const wmh_config* cfg)
{
    if (!e || !cfg) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        validate_config(*cfg);
        if (cfg->device != e->cfg.device || cfg->rank != e->cfg.rank || cfg->world != e->cfg.world)
            throw std::invalid_argument("device, rank and world are fixed at creation");
        if (e->have_walks && (cfg->walks_per_vertex != e->walks_wpv || cfg->walk_length != e->walks_L))
            throw std::logic_error("walks_per_vertex / walk_length cannot change while a corpus exists: call wmh_destroy_index first");
        e->cfg = *cfg;
        return (int) WMH_OK;
    });
}

int wmh_get_config(const wmh_engine* e, wmh_config* cfg)
{
    if (!e || !cfg) return fail(WMH_ERR_INVALID, "null argument");
    *cfg = e->cfg;
    return WMH_OK;
}

uint64_t wmh_number_of_vertices(const wmh_engine* e) { return e ? e->g.n : 0; }
uint64_t wmh_number_of_edges(const wmh_engine* e) { return e ? e->g.m : 0; }

int wmh_flatten_graph(wmh_engine* e, uint64_t* row_ptr, uint32_t* col)
{
    if (!e || !row_ptr) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        Scratch<u64> d_ptr(e->stream, e->g.n + 1);
        Scratch<u32> d_col(e->stream, col ? e->g.m : 0);
        graph_flatten(lc, e->g, d_ptr.ptr, col ? d_col.ptr : nullptr);
        WHARF_CUDA(cudaMemcpyAsync(row_ptr, d_ptr.ptr, (e->g.n + 1) * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
        if (col && e->g.m) WHARF_CUDA(cudaMemcpyAsync(col, d_col.ptr, e->g.m * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_degree(wmh_engine* e, uint32_t vertex, uint64_t* degree)
{
    if (!e || !degree) return fail(WMH_ERR_INVALID, "null argument");
    if (vertex >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    return guarded([&] {
        e->bind();
        RowDesc d;
        WHARF_CUDA(cudaMemcpyAsync(&d, e->g.desc.ptr + vertex, sizeof(d), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        *degree = d.degree;
        return (int) WMH_OK;
    });
}

int wmh_neighbors(wmh_engine* e, uint32_t vertex, uint32_t* out, uint64_t capacity, uint64_t* degree)
{
    if (!e) return fail(WMH_ERR_INVALID, "null argument");
    if (vertex >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    return guarded([&] {
        e->bind();
        RowDesc d;
        WHARF_CUDA(cudaMemcpyAsync(&d, e->g.desc.ptr + vertex, sizeof(d), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        if (degree) *degree = d.degree;
        if (out)
        {
            if (capacity < d.degree) return fail(WMH_ERR_RANGE, "neighbour buffer too small");
            // neighbour i sits at begin + (i << tail): a strided copy when the row is interleaved with aux words
            if (d.degree) WHARF_CUDA(cudaMemcpy2DAsync(out, sizeof(u32), e->g.pool.ptr + d.begin, sizeof(u32) << e->g.tail, sizeof(u32), d.degree,
                                                       cudaMemcpyDeviceToHost, e->stream));
            WHARF_CUDA(cudaStreamSynchronize(e->stream));
        }
        return (int) WMH_OK;
    });
}

int wmh_generate_initial_random_walks(wmh_engine* e)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    return guarded([&] {
        generate(*e);
        return (int) WMH_OK;
    });
}

int wmh_generate_initial_random_walks_to_host(wmh_engine* e, uint32_t* out, uint64_t capacity_walks)
{
    if (!e || !out) return fail(WMH_ERR_INVALID, "null argument");
    if (capacity_walks < local_walk_capacity(e)) return fail(WMH_ERR_RANGE, "corpus buffer too small");
    return guarded([&] {
        generate(*e, out);
        return (int) WMH_OK;
    });
}

uint64_t wmh_last_steps(const wmh_engine* e) { return e ? e->last_steps : 0; }
uint64_t wmh_number_of_walks(const wmh_engine* e) { return e ? e->g.n * (e->have_walks ? e->walks_wpv : e->cfg.walks_per_vertex) : 0; }
uint64_t wmh_number_of_local_walks(const wmh_engine* e)
{
    if (!e) return 0;
    return e->have_walks ? e->w_local : local_walk_capacity(e);
}

int wmh_insert_edges_batch(wmh_engine* e, uint64_t m, uint32_t* edges, int sorted, int remove_dups, uint64_t, int apply, uint64_t* n_affected)
{ return update_host(e, m, edges, sorted, remove_dups, apply, n_affected, false); }
int wmh_delete_edges_batch(wmh_engine* e, uint64_t m, uint32_t* edges, int sorted, int remove_dups, uint64_t, int apply, uint64_t* n_affected)
{ return update_host(e, m, edges, sorted, remove_dups, apply, n_affected, true); }
int wmh_insert_edges_batch_device(wmh_engine* e, uint64_t m, const uint32_t* d_edges, int remove_dups, int apply, uint64_t* n_affected)
{ return update_device(e, m, d_edges, remove_dups, apply, n_affected, false); }
int wmh_delete_edges_batch_device(wmh_engine* e, uint64_t m, const uint32_t* d_edges, int remove_dups, int apply, uint64_t* n_affected)
{ return update_device(e, m, d_edges, remove_dups, apply, n_affected, true); }

int wmh_affected_walks(wmh_engine* e, uint64_t* ids, uint32_t* positions, uint64_t capacity)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (capacity < e->n_affected) return fail(WMH_ERR_RANGE, "affected-walk buffer too small");
    return guarded([&] {
        e->bind();
        const u64 k = e->n_affected;
        if (k == 0) return (int) WMH_OK;
        Scratch<u64> d_global;
        if (ids)
        {
            d_global.allocate(e->stream, k);
            walks_global_ids(e->lc(), e->map, e->aff_ids.ptr, k, d_global.ptr);
            WHARF_CUDA(cudaMemcpyAsync(ids, d_global.ptr, k * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
        }
        if (positions) WHARF_CUDA(cudaMemcpyAsync(positions, e->aff_pos.ptr, k * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_affected_walks_global_device(wmh_engine* e, uint64_t* d_ids, uint64_t capacity)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (capacity < e->n_affected) return fail(WMH_ERR_RANGE, "affected-walk buffer too small");
    if (e->n_affected && !d_ids) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        e->bind();
        walks_global_ids(e->lc(), e->map, e->aff_ids.ptr, e->n_affected, d_ids);
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_affected_walks_device(wmh_engine* e, const uint32_t** d_local_ids, const uint32_t** d_positions, uint64_t* count)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (d_local_ids) *d_local_ids = e->aff_ids.ptr;
    if (d_positions) *d_positions = e->aff_pos.ptr;
    if (count) *count = e->n_affected;
    return WMH_OK;
}

int wmh_touched_vertices(wmh_engine* e, uint32_t* out, uint64_t capacity, uint64_t* count)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (count) *count = e->n_touched;
    if (!out) return WMH_OK;
    if (capacity < e->n_touched) return fail(WMH_ERR_RANGE, "touched-vertex buffer too small");
    return guarded([&] {
        e->bind();
        if (e->n_touched) WHARF_CUDA(cudaMemcpyAsync(out, e->touched.ptr, e->n_touched * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_scan_touched_no_alloc(wmh_engine* e, const uint32_t* d_vertices, uint64_t count, uint32_t* d_min_pos)
{
    if (!e || !d_min_pos || (count && !d_vertices)) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (e->g.n > kNoAllocMaxVertices)
        return fail(WMH_ERR_RANGE, "wmh_scan_touched_no_alloc: the graph has more vertices than its static scratch holds (2^25 = 33554432)");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        walks_scan_touched_no_alloc(lc, e->cfg.device, e->walks.ptr, e->w_local, e->walks_L, e->g.n, d_vertices, count, d_min_pos);
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_walk(wmh_engine* e, uint64_t walk_id, uint32_t* vertices, uint32_t* length)
{
    if (!e || !vertices) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    const i64 local = global_to_local(*e, walk_id);
    if (local < 0) return fail(WMH_ERR_RANGE, "walk id not owned by this rank");
    return guarded([&] {
        e->bind();
        std::vector<u32> row;
        fetch_walk(*e, (u64) local, row);
        u32 len = 0;
        while (len < e->walks_L && row[len] != kNoVertex) len++;
        std::memcpy(vertices, row.data(), e->walks_L * sizeof(u32));
        if (length) *length = len;
        return (int) WMH_OK;
    });
}

int wmh_walk_string(wmh_engine* e, uint64_t walk_id, char* buf, size_t capacity, size_t* needed)
{
    if (!e) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    const i64 local = global_to_local(*e, walk_id);
    if (local < 0) return fail(WMH_ERR_RANGE, "walk id not owned by this rank");
    return guarded([&] {
        e->bind();
        std::vector<u32> row;
        fetch_walk(*e, (u64) local, row);
        std::string s;
        for (u32 p = 0; p < e->walks_L && row[p] != kNoVertex; p++) { s += std::to_string(row[p]); s += ' '; }   // wharfmh.h:376
        if (needed) *needed = s.size() + 1;
        if (buf)
        {
            if (capacity < s.size() + 1) return fail(WMH_ERR_RANGE, "string buffer too small");
            std::memcpy(buf, s.c_str(), s.size() + 1);
        }
        return (int) WMH_OK;
    });
}

int wmh_vertex_at_walk(wmh_engine* e, uint64_t walk_id, uint32_t position, uint32_t* vertex)
{
    if (!e || !vertex) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (position >= e->walks_L) return fail(WMH_ERR_RANGE, "position out of range");
    const i64 local = global_to_local(*e, walk_id);
    if (local < 0) return fail(WMH_ERR_RANGE, "walk id not owned by this rank");
    return guarded([&] {
        e->bind();
        WHARF_CUDA(cudaMemcpyAsync(vertex, e->walks.ptr + (u64) position * e->w_local + (u64) local, 4, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_walks_copy(wmh_engine* e, uint64_t first, uint64_t count, uint32_t* out)
{
    if (!e || (count && !out)) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (first + count > e->w_local) return fail(WMH_ERR_RANGE, "walk range out of bounds");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        const u32 L = e->walks_L;
        const u64 chunk = std::max<u64>(1, (64ull << 20) / L);              // 256 MiB staging
        Scratch<u32> stage(e->stream, std::min(chunk, std::max<u64>(count, 1)) * L);
        for (u64 done = 0; done < count; done += chunk)
        {
            const u64 c = std::min(chunk, count - done);
            walks_transpose_rows(lc, e->walks.ptr, e->w_local, L, first + done, c, stage.ptr);
            WHARF_CUDA(cudaMemcpyAsync(out + done * L, stage.ptr, c * L * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
            WHARF_CUDA(cudaStreamSynchronize(e->stream));
        }
        return (int) WMH_OK;
    });
}

int wmh_walks_gather(wmh_engine* e, const uint32_t* local_ids, uint64_t count, uint32_t* out)
{
    if (!e || (count && (!out || !local_ids))) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    for (u64 i = 0; i < count; i++) if (local_ids[i] >= e->w_local) return fail(WMH_ERR_RANGE, "local walk id out of range");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        const u32 L = e->walks_L;
        Scratch<u32> d_ids(e->stream, count), d_rows(e->stream, count * L);
        if (count == 0) return (int) WMH_OK;
        WHARF_CUDA(cudaMemcpyAsync(d_ids.ptr, local_ids, count * sizeof(u32), cudaMemcpyHostToDevice, e->stream));
        walks_gather_rows(lc, e->walks.ptr, e->w_local, L, d_ids.ptr, count, d_rows.ptr);
        WHARF_CUDA(cudaMemcpyAsync(out, d_rows.ptr, count * L * sizeof(u32), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

uint64_t wmh_local_to_global_walk(const wmh_engine* e, uint64_t local_walk)
{
    if (!e || !e->have_walks || local_walk >= e->w_local) return ~0ull;
    return local_to_global(*e, local_walk);
}

int64_t wmh_global_to_local_walk(const wmh_engine* e, uint64_t walk_id) { return e ? global_to_local(*e, walk_id) : -1; }

uint32_t wmh_walk_owner(const wmh_engine* e, uint64_t walk_id)
{
    if (!e || e->g.n == 0) return 0;
    return (e->have_walks ? e->map : config_map(*e)).owner(walk_id);
}

// ---- the generators, host forms (no device needed): the very functions the draw table and the kernels are built from ----
void wmh_random_stream(uint64_t seed, uint64_t count, uint64_t* out)
{
    if (!out) return;
    u64 s[2];
    random_seed(seed, s);
    for (u64 k = 0; k < count; k++) out[k] = random_next(s);
}

double wmh_unit_double(uint64_t draw) { return unit_double(draw); }

uint64_t wmh_keyed_draw(uint64_t seed, uint64_t domain, uint64_t a, uint64_t b, uint64_t c) { return keyed_draw(seed, domain, a, b, c); }

uint32_t wmh_partition_owner(uint64_t n, uint32_t walks_per_vertex, uint32_t world, uint64_t walk_id)
{
    if (n == 0 || world == 0 || walks_per_vertex == 0) return 0;
    return make_shard_map(n, walks_per_vertex, 0, world).owner(walk_id);
}

uint64_t wmh_partition_local_walk_ids(uint64_t n, uint32_t walks_per_vertex, uint32_t rank, uint32_t world, uint64_t* out, uint64_t capacity)
{
    if (n == 0 || world == 0 || rank >= world || walks_per_vertex == 0) return 0;
    const ShardMap m = make_shard_map(n, walks_per_vertex, rank, world);
    const u64 count = m.local_walks();
    if (out)
        for (u64 w = 0; w < count && w < capacity; w++) out[w] = m.global_id(w);
    return count;
}

uint64_t wmh_local_walk_ids(const wmh_engine* e, uint64_t* out, uint64_t capacity)
{
    if (!e) return 0;
    const ShardMap m = e->have_walks ? e->map : config_map(*e);
    const u64 count = m.local_walks();
    if (out)
        for (u64 w = 0; w < count && w < capacity; w++) out[w] = m.global_id(w);
    return count;
}

int wmh_corpus_fingerprint(wmh_engine* e, uint64_t* sum)
{
    if (!e || !sum) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        Scratch<unsigned long long> d_sum(e->stream, 1);
        WHARF_CUDA(cudaMemsetAsync(d_sum.ptr, 0, 8, e->stream));
        walks_fingerprint(lc, e->walks.ptr, e->w_local, e->walks_L, d_sum.ptr);
        unsigned long long h = 0;
        WHARF_CUDA(cudaMemcpyAsync(&h, d_sum.ptr, 8, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        *sum = h;
        return (int) WMH_OK;
    });
}

int wmh_sampler_chain(wmh_engine* e, uint32_t cur, uint32_t prev, uint64_t iters, uint32_t* init_value, uint64_t* counts, uint64_t capacity)
{
    if (!e || !counts) return fail(WMH_ERR_INVALID, "null argument");
    if (cur >= e->g.n || prev >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    return guarded([&] {
        e->bind();
        validate_config(e->cfg);
        RowDesc d;
        WHARF_CUDA(cudaMemcpyAsync(&d, e->g.desc.ptr + cur, sizeof(d), cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        if (capacity < d.degree) return fail(WMH_ERR_RANGE, "counts must hold degree(cur) entries");
        Scratch<unsigned long long> d_counts(e->stream, std::max<u64>(1, d.degree));
        Scratch<u32> d_init(e->stream, 1);
        WHARF_CUDA(cudaMemsetAsync(d_counts.ptr, 0, std::max<u64>(1, d.degree) * 8, e->stream));
        WalkParams wp = walk_params(*e);
        sampler_chain(e->lc(), wp, cur, prev, iters, d_init.ptr, d_counts.ptr);
        u32 init = kNoVertex;
        WHARF_CUDA(cudaMemcpyAsync(&init, d_init.ptr, 4, cudaMemcpyDeviceToHost, e->stream));
        if (d.degree) WHARF_CUDA(cudaMemcpyAsync(counts, d_counts.ptr, d.degree * 8, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        if (init_value) *init_value = init;
        return (int) WMH_OK;
    });
}

int wmh_build_index(wmh_engine* e)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    return guarded([&] {
        e->bind();
        e->index.valid = false;
        ensure_index(*e);
        return (int) WMH_OK;
    });
}

int wmh_index_find_next(wmh_engine* e, uint32_t vertex, uint64_t walk_id, uint32_t position, uint32_t* next, int* found)
{
    if (!e || !next || !found) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (vertex >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    const i64 local = global_to_local(*e, walk_id);
    if (local < 0 || position >= e->walks_L)
    {
        *found = 0;
        *next = kRefSentinel;
        return WMH_OK;
    }
    return guarded([&] {
        e->bind();
        ensure_index(*e);
        auto lc = e->lc();
        index_find_next(lc, e->index, e->g.n, e->walks.ptr, e->w_local, e->walks_L, vertex, (u64) local, position, e->small.ptr);
        u32 r[2];
        WHARF_CUDA(cudaMemcpyAsync(r, e->small.ptr, 8, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        *next = r[0];
        *found = (int) r[1];
        return (int) WMH_OK;
    });
}

int wmh_index_size(wmh_engine* e, uint32_t vertex, uint64_t* entries)
{
    if (!e || !entries) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (vertex >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    return guarded([&] {
        e->bind();
        ensure_index(*e);
        std::vector<u64> p(2 * (size_t) e->index.chunks);
        for (u32 c = 0; c < e->index.chunks; c++)
            WHARF_CUDA(cudaMemcpyAsync(&p[2 * c], e->index.ptr.ptr + (u64) c * (e->g.n + 1) + vertex, 16, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        *entries = 0;
        for (u32 c = 0; c < e->index.chunks; c++) *entries += p[2 * c + 1] - p[2 * c];
        return (int) WMH_OK;
    });
}

int wmh_index_sizes(wmh_engine* e, uint64_t* entries)
{
    if (!e || !entries) return fail(WMH_ERR_INVALID, "null argument");
    const u64 n = e->g.n;
    if (!e->have_walks)                                          // no corpus: every vertex's index is empty (wharfmh.h:237-245)
    {
        std::fill(entries, entries + n, 0ull);
        return WMH_OK;
    }
    return guarded([&] {
        e->bind();
        ensure_index(*e);
        std::vector<u64> ptr(n + 1);
        std::fill(entries, entries + n, 0ull);
        for (u32 c = 0; c < e->index.chunks; c++)
        {
            WHARF_CUDA(cudaMemcpyAsync(ptr.data(), e->index.ptr.ptr + (u64) c * (n + 1), (n + 1) * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
            WHARF_CUDA(cudaStreamSynchronize(e->stream));
            for (u64 v = 0; v < n; v++) entries[v] += ptr[v + 1] - ptr[v];
        }
        return (int) WMH_OK;
    });
}

int wmh_compressed_walks(wmh_engine* e, uint64_t* ptr, uint64_t* codes, uint64_t capacity, uint64_t* total)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    int rc = guarded([&] {
        e->bind();
        try { ensure_codes(*e); }
        catch (const std::domain_error& d) { return fail(WMH_ERR_DOMAIN, d.what()); }
        if (total) *total = e->index.entries;
        if (codes && capacity < e->index.entries) return fail(WMH_ERR_RANGE, "code buffer too small");   // before any output is written
        if (ptr) WHARF_CUDA(cudaMemcpyAsync(ptr, e->index.ptr.ptr, (e->g.n + 1) * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
        if (codes)
        {
            if (e->index.entries) WHARF_CUDA(cudaMemcpyAsync(codes, e->codes.ptr, e->index.entries * sizeof(u64), cudaMemcpyDeviceToHost, e->stream));
        }
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
    return rc;
}

int wmh_compressed_find_next(wmh_engine* e, uint32_t vertex, uint64_t walk_id, uint32_t position, uint32_t* next, int* found)
{
    if (!e || !next || !found) return fail(WMH_ERR_INVALID, "null argument");
    if (!e->have_walks) return fail(WMH_ERR_STATE, "walks have not been generated");
    if (vertex >= e->g.n) return fail(WMH_ERR_RANGE, "vertex out of range");
    return guarded([&] {
        e->bind();
        try { ensure_codes(*e); }
        catch (const std::domain_error& d) { return fail(WMH_ERR_DOMAIN, d.what()); }
        u64 p[2];
        WHARF_CUDA(cudaMemcpyAsync(p, e->index.ptr.ptr + vertex, 16, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        auto lc = e->lc();
        compressed_find_next(lc, e->codes.ptr + p[0], p[1] - p[0], walk_id, position, e->walks_L, e->small.ptr);
        u32 r[2];
        WHARF_CUDA(cudaMemcpyAsync(r, e->small.ptr, 8, cudaMemcpyDeviceToHost, e->stream));
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        *next = r[1] ? r[0] : 0xFFFFFFFFu;                // compressed_walks.h:54: next_vertex = -1 when absent
        *found = (int) r[1];
        return (int) WMH_OK;
    });
}

// ---- Szudzik over device arrays ------------------------------------------------------------------------
#define WMH_PAIR_OP(body, with_overflow)                                                        \
    return guarded([&] {                                                                        \
        DefaultCtx ctx(device);                                                                 \
        Scratch<unsigned long long> d_over(nullptr, 1);                                         \
        WHARF_CUDA(cudaMemsetAsync(d_over.ptr, 0, 8, nullptr));                                 \
        body;                                                                                   \
        unsigned long long over = 0;                                                            \
        WHARF_CUDA(cudaMemcpy(&over, d_over.ptr, 8, cudaMemcpyDeviceToHost));                   \
        if (with_overflow) *(with_overflow) = over;                                             \
        return (int) WMH_OK;                                                                    \
    })

int wmh_szudzik_pair_u64(int device, const uint64_t* x, const uint64_t* y, uint64_t* out, uint64_t count, uint64_t* overflow)
{ WMH_PAIR_OP(szudzik_pair_u64(ctx.lc, x, y, out, count, d_over.ptr), overflow); }
int wmh_szudzik_unpair_u64(int device, const uint64_t* z, uint64_t* x, uint64_t* y, uint64_t count)
{ WMH_PAIR_OP(szudzik_unpair_u64(ctx.lc, z, x, y, count), (uint64_t*) nullptr); }
int wmh_szudzik_pair_u32(int device, const uint32_t* x, const uint32_t* y, uint32_t* out, uint64_t count, uint64_t* overflow)
{ WMH_PAIR_OP(szudzik_pair_u32(ctx.lc, x, y, out, count, d_over.ptr), overflow); }
int wmh_szudzik_unpair_u32(int device, const uint32_t* z, uint32_t* x, uint32_t* y, uint64_t count)
{ WMH_PAIR_OP(szudzik_unpair_u32(ctx.lc, z, x, y, count), (uint64_t*) nullptr); }
int wmh_szudzik_pair_triplet_u64(int device, const uint64_t* a, const uint64_t* b, const uint64_t* c, uint64_t* out, uint64_t count, uint64_t* overflow)
{ WMH_PAIR_OP(szudzik_pair_triplet_u64(ctx.lc, a, b, c, out, count, d_over.ptr), overflow); }
int wmh_szudzik_unpair_triplet_u64(int device, const uint64_t* z, uint64_t* a, uint64_t* b, uint64_t* c, uint64_t count)
{ WMH_PAIR_OP(szudzik_unpair_triplet_u64(ctx.lc, z, a, b, c, count), (uint64_t*) nullptr); }

uint64_t wmh_szudzik_pair(uint64_t x, uint64_t y) { return szudzik_pair(x, y, nullptr); }
void wmh_szudzik_unpair(uint64_t z, uint64_t* x, uint64_t* y)
{
    u64 a, b;
    szudzik_unpair(z, &a, &b);
    if (x) *x = a;
    if (y) *y = b;
}

// ---- reporting -----------------------------------------------------------------------------------------
double wmh_timer_get(const wmh_engine* e, int which) { return (e && which >= 0 && which < 5) ? e->timers[which] : 0.0; }
void wmh_timer_reset(wmh_engine* e, int which) { if (e && which >= 0 && which < 5) e->timers[which] = 0.0; }

int wmh_memory_footprint(wmh_engine* e, wmh_memory* out)
{
    if (!e || !out) return fail(WMH_ERR_INVALID, "null argument");
    std::memset(out, 0, sizeof(*out));
    out->graph_bytes = e->g.desc.bytes() + e->g.m * sizeof(u32) + e->g.version.bytes();
    out->graph_pool_bytes = e->g.pool.bytes() + e->g.spare.bytes();
    out->walks_bytes = e->have_walks ? e->w_local * e->walks_L * sizeof(u32) : 0;
    out->index_bytes = e->index.valid ? e->index.ptr.bytes() + e->index.entries * sizeof(u32) + (e->codes_valid ? e->codes.bytes() : 0) : 0;
    out->scratch_bytes = e->min_pos.bytes() + e->aff_ids.bytes() + e->aff_pos.bytes() + e->bitmap.bytes() + e->touched.bytes() + e->draw.bytes();
    out->total_bytes = e->g.desc.bytes() + e->g.version.bytes() + out->graph_pool_bytes + e->walks.bytes() + e->index.ptr.bytes() +
                       e->index.cells.bytes() + e->codes.bytes() + out->scratch_bytes;
    return WMH_OK;
}

uint64_t wmh_kernel_launches(const wmh_engine* e) { return e ? e->launches : 0; }
double wmh_last_walk_kernel_ms(const wmh_engine* e) { return e ? e->last_walk_kernel_ms : 0.0; }
double wmh_last_index_ms(const wmh_engine* e) { return e ? e->last_index_ms : 0.0; }
uint32_t wmh_last_search_mode(const wmh_engine* e) { return e ? e->last_search : 0u; }
uint32_t wmh_neighbour_filter(const wmh_engine* e) { return (e && e->g.tail) ? (e->g.tri ? 2u : 1u) : 0u; }
int wmh_filter_check(wmh_engine* e, uint64_t* missing)
{
    if (!e || !missing) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        e->bind();
        auto lc = e->lc();
        *missing = graph_filter_check(lc, e->g);
        return (int) WMH_OK;
    });
}

const uint32_t* wmh_device_walks(const wmh_engine* e) { return (e && e->have_walks) ? e->walks.ptr : nullptr; }
void* wmh_stream(const wmh_engine* e) { return e ? (void*) e->stream : nullptr; }

int wmh_synchronize(wmh_engine* e)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    return guarded([&] {
        e->bind();
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        return (int) WMH_OK;
    });
}

int wmh_peer_create(wmh_engine* e, uint64_t bytes, unsigned char handle_out[WMH_PEER_HANDLE_BYTES])
{
    if (!e || !handle_out) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        e->bind();
        if (e->cfg.world > kPeerMaxWorld) throw std::invalid_argument("peer memory supports up to 8 ranks (one box)");
        if (bytes < kPeerFlagBytes + 64) throw std::invalid_argument("exchange buffer too small");
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        peer_release(e->peer);
        e->peer = PeerLink{};
        e->peer.world = e->cfg.world;
        e->peer.rank = e->cfg.rank;
        void* p = nullptr;
        WHARF_CUDA(cudaMalloc(&p, bytes));                    // IPC handles need a plain allocation, not the stream-ordered pool
        e->peer.ptrs.buf[e->peer.rank] = static_cast<u8*>(p);
        e->peer.bytes = bytes;
        e->peer.created = true;
        WHARF_CUDA(cudaMemset(p, 0, bytes));
        WHARF_CUDA(cudaDeviceSynchronize());
        cudaIpcMemHandle_t h;
        WHARF_CUDA(cudaIpcGetMemHandle(&h, p));
        static_assert(sizeof(h) == WMH_PEER_HANDLE_BYTES, "CUDA IPC handle size");
        std::memcpy(handle_out, &h, sizeof(h));
        return (int) WMH_OK;
    });
}

int wmh_peer_attach(wmh_engine* e, const unsigned char* handles)
{
    if (!e || !handles) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        e->bind();
        PeerLink& link = e->peer;
        if (!link.created) throw std::invalid_argument("wmh_peer_create must be called first");
        if (link.world != e->cfg.world || link.rank != e->cfg.rank) throw std::invalid_argument("rank/world changed since wmh_peer_create");
        for (u32 g = 0; g < link.world; g++)
        {
            if (g == link.rank || link.ptrs.buf[g]) continue;
            cudaIpcMemHandle_t h;
            std::memcpy(&h, handles + (size_t) g * WMH_PEER_HANDLE_BYTES, sizeof(h));
            void* p = nullptr;
            WHARF_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
            link.ptrs.buf[g] = static_cast<u8*>(p);
        }
        link.attached = true;
        return (int) WMH_OK;
    });
}

int wmh_peer_detach(wmh_engine* e)
{
    if (!e) return fail(WMH_ERR_INVALID, "null engine");
    return guarded([&] {
        e->bind();
        WHARF_CUDA(cudaStreamSynchronize(e->stream));
        peer_release(e->peer);
        return (int) WMH_OK;
    });
}

uint64_t wmh_peer_bytes_needed(const wmh_engine* e) { return e ? peer_bytes_needed(*e) : 0; }
uint64_t wmh_peer_exchanges(const wmh_engine* e) { return e ? e->peer.exchanges : 0; }

int wmh_trim_memory(int device)
{
    return guarded([&] {
        DefaultCtx ctx(device);
        WHARF_CUDA(cudaDeviceSynchronize());
        cudaMemPool_t pool = nullptr;
        WHARF_CUDA(library_pool(device, &pool));
        WHARF_CUDA(cudaMemPoolTrimTo(pool, 0));
        return (int) WMH_OK;
    });
}

int wmh_generate_batch_of_edges_device(int device, uint64_t edges_number, uint64_t vertices_number, uint64_t batch_seed, int self_loops,
                                       int directed, double a, double b, double c, uint32_t* d_out, uint64_t* count)
{
    if (!count || (edges_number && !d_out)) return fail(WMH_ERR_INVALID, "null argument");
    if (vertices_number < 2) return fail(WMH_ERR_INVALID, "vertices_number must be >= 2");
    if (a + b + c > 1) return fail(WMH_ERR_INVALID, "a + b + c add to more than 1");      // rmat_util.h:16
    return guarded([&] {
        DefaultCtx ctx(device);
        u64 kept = 0;
        rmat_generate_device(ctx.lc, edges_number, vertices_number, batch_seed, self_loops != 0, directed != 0, a, b, c, d_out, &kept);
        *count = kept;
        return (int) WMH_OK;
    });
}

int wmh_generate_batch_of_edges(int device, uint64_t edges_number, uint64_t vertices_number, uint64_t batch_seed, int self_loops, int directed,
                                double a, double b, double c, uint32_t* out, uint64_t capacity, uint64_t* count)
{
    if (!count || (edges_number && !out)) return fail(WMH_ERR_INVALID, "null argument");
    if (vertices_number < 2) return fail(WMH_ERR_INVALID, "vertices_number must be >= 2");
    if (a + b + c > 1) return fail(WMH_ERR_INVALID, "a + b + c add to more than 1");      // rmat_util.h:16
    const u64 slots = edges_number * (directed ? 1 : 2);
    if (capacity < slots) return fail(WMH_ERR_RANGE, "out must hold edges_number pairs (twice that when not directed)");
    return guarded([&] {
        DefaultCtx ctx(device);
        Scratch<u32> d_pairs(nullptr, 2 * slots);
        u64 kept = 0;
        rmat_generate_device(ctx.lc, edges_number, vertices_number, batch_seed, self_loops != 0, directed != 0, a, b, c, d_pairs.ptr, &kept);
        if (kept) WHARF_CUDA(cudaMemcpyAsync(out, d_pairs.ptr, kept * 2 * sizeof(u32), cudaMemcpyDeviceToHost, nullptr));
        WHARF_CUDA(cudaStreamSynchronize(nullptr));
        *count = kept;
        return (int) WMH_OK;
    });
}

int wmh_pairs_to_csr_device(int device, uint64_t n, const uint32_t* d_pairs, uint64_t count, uint64_t* d_offsets, uint32_t* d_edges)
{
    if (!d_offsets || (count && (!d_pairs || !d_edges))) return fail(WMH_ERR_INVALID, "null argument");
    return guarded([&] {
        DefaultCtx ctx(device);
        pairs_to_csr_device(ctx.lc, n, d_pairs, count, d_offsets, d_edges);
        WHARF_CUDA(cudaStreamSynchronize(nullptr));
        return (int) WMH_OK;
    });
}

}  // extern "C"
