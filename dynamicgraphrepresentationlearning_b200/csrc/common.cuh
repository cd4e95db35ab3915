// Shared device/host helpers of libwharfmh_b200 (sm_100a only).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <stdexcept>
#include <string>

namespace wharf
{
using u8  = uint8_t;
using u32 = uint32_t;
using u64 = uint64_t;
using i64 = int64_t;

constexpr u32 kNoVertex    = 0xFFFFFFFFu;   // "no vertex" in the walk matrix
constexpr u32 kRefSentinel = 0xFFFFFFFEu;   // the reference's end-of-walk marker (graph/wharfmh.h:283,317)
constexpr u32 kNoPos       = 0xFFFFFFFFu;   // "walk not affected" in the min-position array

// Row descriptor of the device adjacency: neighbours of v live at pool[begin, begin + degree), ascending.
// One aligned 8-byte load per walker step (never straddles a 32-byte sector).
struct __align__(8) RowDesc
{
    u32 begin;
    u32 degree;
};

struct CudaError : std::runtime_error
{
    cudaError_t code;
    CudaError(cudaError_t c, const std::string& what) : std::runtime_error(what), code(c) {}
};

#define WHARF_CUDA(expr)                                                                              \
    do {                                                                                              \
        cudaError_t _e = (expr);                                                                      \
        if (_e != cudaSuccess)                                                                        \
            throw ::wharf::CudaError(_e, std::string(#expr) + " failed: " + cudaGetErrorString(_e) +  \
                                             " (" + __FILE__ + ":" + std::to_string(__LINE__) + ")"); \
    } while (0)

inline unsigned div_up(u64 a, u64 b) { return (unsigned) ((a + b - 1) / b); }

// Grid for a grid-stride kernel: a whole number of waves on the 148-SM part, capped by the work.
inline unsigned grid_for(u64 work_items, unsigned block, unsigned blocks_per_sm, int sm_count)
{
    u64 need = (work_items + block - 1) / block;
    u64 full = (u64) sm_count * blocks_per_sm;
    if (need == 0) need = 1;
    return (unsigned) (need < full ? need : full);
}

// Stream on which DeviceBuffer allocates and frees (cudaMallocAsync / cudaFreeAsync).  An engine sets it to its
// own stream while it runs (wmh_engine::bind); with no stream set the buffers fall back to cudaMalloc/cudaFree.
// Stream-ordered allocation keeps engine create/destroy and pool growth out of the driver's synchronous
// allocator: freed blocks stay in the device memory pool (release threshold = max) and are handed out again.
inline thread_local cudaStream_t g_alloc_stream = nullptr;
inline thread_local bool g_alloc_async = false;

// The library's own stream-ordered memory pool on the current device (engine.cu): created on first use with release
// threshold = max, so freed blocks are handed out again by the next engine or batch, and the host application's
// default pool (and its cudaMalloc) is left alone; wmh_trim_memory returns the cached blocks to the driver.
cudaError_t pool_malloc_async(void** ptr, size_t bytes, cudaStream_t stream);

// Owning device allocation.  Not copyable.
template <class T>
struct DeviceBuffer
{
    T*  ptr = nullptr;
    u64 count = 0;
    cudaStream_t stream = nullptr;   // stream the block was allocated on (async blocks only)
    bool async = false;

    DeviceBuffer() = default;
    DeviceBuffer(const DeviceBuffer&) = delete;
    DeviceBuffer& operator=(const DeviceBuffer&) = delete;
    DeviceBuffer(DeviceBuffer&& o) noexcept : ptr(o.ptr), count(o.count), stream(o.stream), async(o.async) { o.ptr = nullptr; o.count = 0; }
    DeviceBuffer& operator=(DeviceBuffer&& o) noexcept
    {
        if (this != &o) { release(); ptr = o.ptr; count = o.count; stream = o.stream; async = o.async; o.ptr = nullptr; o.count = 0; }
        return *this;
    }
    ~DeviceBuffer() { release(); }

    void allocate(u64 n)
    {
        release();
        if (n == 0) return;
        async = g_alloc_async;
        stream = g_alloc_stream;
        cudaError_t e = async ? pool_malloc_async((void**) &ptr, n * sizeof(T), stream) : cudaMalloc((void**) &ptr, n * sizeof(T));
        if (e != cudaSuccess)
        {
            ptr = nullptr;
            cudaGetLastError();
            throw CudaError(e, "device allocation of " + std::to_string(n * sizeof(T)) + " bytes failed: " + cudaGetErrorString(e));
        }
        count = n;
    }
    void ensure(u64 n) { if (n > count) allocate(n); }
    void release()
    {
        if (ptr) { if (async) cudaFreeAsync(ptr, stream); else cudaFree(ptr); }
        ptr = nullptr;
        count = 0;
    }
    u64 bytes() const { return count * sizeof(T); }
};

#ifdef __CUDACC__
// ---- integer helpers shared by kernels -------------------------------------------------------------

// x % d for a 64-bit draw and a 32-bit degree (utility::Random::irand, utils/utility.h:220, whose int
// argument is the degree).  Two chained 32-bit-divisor reductions: x = hi*2^32 + lo,
// r1 = hi % d, result = (r1*2^32 + lo) % d; the second is done with one double-precision quotient
// estimate (error <= 1) and an exact integer correction.
__device__ __forceinline__ u32 mod_u64_u32(u64 x, u32 d)
{
    const u32 hi = (u32) (x >> 32);
    const u32 r1 = hi % d;
    const u64 t  = ((u64) r1 << 32) | (u32) x;                 // < d * 2^32
    const double qd = __ddiv_rz(__ull2double_rz(t), (double) d);
    u64 q = __double2ull_rz(qd);
    i64 rem = (i64) (t - q * (u64) d);
    if (rem < 0) rem += d;                                     // estimate one too high
    if (rem >= (i64) d) rem -= d;                              // estimate one too low
    return (u32) rem;
}

// The same remainder for several x and ONE divisor: inv = 1.0 / d computed once (reciprocal_of), each reduction is then two
// double multiplications instead of an integer modulo and a double division.  Exact: both quotient estimates are within
// 2^-19 of the true quotient (hi < 2^32 and t < d * 2^32 carry at most 2^-52 of relative error through the product),
// so truncation lands on the true quotient or one beside it, and the two corrections below restore the remainder.
__device__ __forceinline__ double reciprocal_of(u32 d) { return __drcp_rn((double) d); }

__device__ __forceinline__ u32 mod_u64_u32_rcp(u64 x, u32 d, double inv)
{
    const u32 hi = (u32) (x >> 32);
    const u32 q1 = __double2uint_rz(__dmul_rz((double) hi, inv));
    i64 r1 = (i64) hi - (i64) ((u64) q1 * d);
    if (r1 < 0) r1 += d;
    if (r1 >= (i64) d) r1 -= d;
    const u64 t = ((u64) r1 << 32) | (u32) x;                  // < d * 2^32
    const u64 q = __double2ull_rz(__dmul_rz(__ull2double_rz(t), inv));
    i64 rem = (i64) (t - q * (u64) d);
    if (rem < 0) rem += d;
    if (rem >= (i64) d) rem -= d;
    return (u32) rem;
}

__device__ __forceinline__ u32 warp_sum(u32 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum folded into one 64-bit atomic per block.
__device__ __forceinline__ void block_add_u64(unsigned long long* target, u32 v)
{
    __shared__ u32 s_part[32];
    v = warp_sum(v);
    const u32 lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) s_part[wid] = v;
    __syncthreads();
    if (wid == 0)
    {
        u32 t = lane < ((blockDim.x + 31) >> 5) ? s_part[lane] : 0;
        t = warp_sum(t);
        if (lane == 0 && t) atomicAdd(target, (unsigned long long) t);
    }
    __syncthreads();
}

// ---- neighbour filter ---------------------------------------------------------------------------------
// A blocked Bloom filter aligned with the neighbour pool: row v owns the 16-bit blocks filt[begin, begin + degree)
// (16 bits per neighbour); element x sets two bits of block begin + umulhi(h(x), degree).  A clear bit proves
// "x is not a neighbour of v" with one 2-byte load; node2vec's weight test (node2vec.h:112-119) is a miss for
// ~94 % of the proposals on the RMAT workloads, so most binary searches never start.  A hit decides nothing.
// Which start vertices a rank owns (walks shard by start vertex).  Vertices are taken in blocks of `world`
// consecutive ids; rank r owns one vertex of every block: slot (r + rot(block)) % world.  The rotation spreads RMAT's
// dense residues (ids whose low bits are zero have the highest degrees and the fewest isolated vertices) evenly:
// live start vertices per rank differ by 0.1 % instead of 4-5 % at world = 8.  The last, partial block is not rotated,
// so rank r owns ceil((n - r) / world) starts, in ascending id order.
__host__ __device__ __forceinline__ u32 shard_rotation(u32 block, u32 world) { return world > 1 ? ((block * 0x9E3779B1u) >> 16) % world : 0u; }
__host__ __device__ __forceinline__ u32 shard_start(u32 local_index, u32 rank, u32 world, u64 n)
{
    const u32 full_blocks = (u32) (n / world);
    const u32 slot = local_index < full_blocks ? (rank + shard_rotation(local_index, world)) % world : rank;
    return local_index * world + slot;
}
__host__ __device__ __forceinline__ u32 shard_owner(u32 vertex, u32 world, u64 n)
{
    const u32 block = vertex / world, slot = vertex % world;
    return block < (u32) (n / world) ? (slot + world - shard_rotation(block, world)) % world : slot;
}

// Which walks a rank owns.  Replicas (walk_id / n) are independent, so whole replicas go first: with q = wpv / world
// every rank owns the q replicas [rank * q, rank * q + q) entirely (all n start vertices).  The remaining
// R = wpv - world * q replicas are shared out in g = gcd(world, R) groups of world / g ranks: group k shards the R / g
// replicas [world * q + k * R / g, ...) by start vertex among its ranks (shard_start above, with the rank's position
// in its group).  So at wpv = 10: 2 ranks own 5 replicas each; 4 ranks 2 replicas + half of one more; 8 ranks 1
// replica + a quarter of one more.  Fewer, larger shares matter to the deterministic coalesced schedule: the walkers
// of a replica merge, so a quarter of one replica costs less than two eighths of two (measured);
// wpv < world with gcd 1 degenerates to pure start-vertex sharding, wpv % world == 0 to pure replica sharding.
// Local walk order: the whole replicas (start ascending), then the shared ones (owned starts ascending), i.e.
// ascending global walk id.
struct ShardMap
{
    u64 n = 0;                // vertices
    u32 rank = 0, world = 1;
    u32 whole = 0;            // replicas owned entirely
    u32 whole_first = 0;      // first of them (global replica number)
    u32 shared_base = 0;      // first replica that is not owned whole by anyone (= world * q)
    u32 shared = 0;           // shared replicas this rank's group works on
    u32 shared_first = 0;     // first of them
    u32 sub_rank = 0, sub_world = 1;   // position in / size of the group
    u32 n_local = 0;          // start vertices this rank owns in a shared replica

    __host__ __device__ u64 whole_walks() const { return (u64) whole * n; }
    __host__ __device__ u64 local_walks() const { return whole_walks() + (u64) shared * n_local; }
    // local walk -> (global replica, start vertex)
    __host__ __device__ void locate(u64 w, u32& replica, u32& start) const
    {
        const u64 ww = whole_walks();
        if (w < ww)
        {
            const u32 r = (u32) (w / n);
            replica = whole_first + r;
            start = (u32) (w - (u64) r * n);
        }
        else
        {
            const u64 x = w - ww;
            const u32 r = (u32) (x / n_local);
            replica = shared_first + r;
            start = shard_start((u32) (x - (u64) r * n_local), sub_rank, sub_world, n);
        }
    }
    __host__ __device__ u64 global_id(u64 w) const
    {
        u32 replica, start;
        locate(w, replica, start);
        return (u64) replica * n + start;
    }
    // rank that owns a global walk id
    __host__ __device__ u32 owner(u64 walk_id) const
    {
        const u32 replica = (u32) (walk_id / n);
        if (replica < shared_base) return whole ? replica / whole : 0u;       // every rank owns the same number of whole replicas
        const u32 group = shared ? (replica - shared_base) / shared : 0u;     // every group works on the same number
        return group * sub_world + shard_owner((u32) (walk_id % n), sub_world, n);
    }
    // local index of an owned global walk id, -1 otherwise
    __host__ __device__ i64 local_of(u64 walk_id) const
    {
        if (owner(walk_id) != rank) return -1;
        const u32 replica = (u32) (walk_id / n);
        const u32 start = (u32) (walk_id % n);
        if (replica < shared_base) return (i64) ((u64) (replica - whole_first) * n + start);
        return (i64) (whole_walks() + (u64) (replica - shared_first) * n_local + start / sub_world);
    }
};

inline ShardMap make_shard_map(u64 n, u32 wpv, u32 rank, u32 world)
{
    ShardMap m;
    m.n = n;
    m.rank = rank;
    m.world = world;
    const u32 q = world > 1 ? wpv / world : 0;       // one rank: everything is "shared" by a group of one
    m.whole = q;
    m.whole_first = rank * q;
    m.shared_base = world > 1 ? world * q : 0;
    const u32 rest = wpv - m.shared_base;
    u32 g = world, b = rest;                         // gcd(world, rest); gcd(world, 0) = world
    while (b) { const u32 t = g % b; g = b; b = t; }
    if (rest == 0) g = world;
    m.sub_world = world / g;
    m.sub_rank = rank % m.sub_world;
    m.shared = rest / g;
    m.shared_first = m.shared_base + (rank / m.sub_world) * m.shared;
    m.n_local = m.sub_rank < n ? (u32) ((n - m.sub_rank + m.sub_world - 1) / m.sub_world) : 0;
    return m;
}

__device__ __forceinline__ u32 filter_hash(u32 x)
{
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ u32 filter_block(u32 h, u32 degree) { return __umulhi(h, degree); }
__device__ __forceinline__ u32 filter_bits(u32 h) { return (1u << (h & 15)) | (1u << ((h >> 4) & 15)); }

// lower_bound over an ascending row whose element i sits at a[i << shift] (shift 1: neighbours interleaved with their
// aux words, kernels_graph.cu).
__device__ __forceinline__ u32 lower_bound_row(const u32* __restrict__ a, u32 n, u32 key, u32 shift)
{
    u32 lo = 0, hi = n;
    while (lo < hi)
    {
        u32 mid = (lo + hi) >> 1;
        if (__ldg(a + ((u64) mid << shift)) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// lower_bound over an ascending u32 array slice.
__device__ __forceinline__ u32 lower_bound_u32(const u32* __restrict__ a, u32 n, u32 key)
{
    u32 lo = 0, hi = n;
    while (lo < hi)
    {
        u32 mid = (lo + hi) >> 1;
        if (__ldg(a + mid) < key) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Largest i with a[i] <= key in an ascending u64 array of n >= 1 entries whose a[0] <= key.
__device__ __forceinline__ u32 upper_segment_u64(const u64* __restrict__ a, u32 n, u64 key)
{
    u32 lo = 0, hi = n;
    while (hi - lo > 1)
    {
        u32 mid = (lo + hi) >> 1;
        if (__ldg(a + mid) <= key) lo = mid; else hi = mid;
    }
    return lo;
}
#endif  // __CUDACC__

}  // namespace wharf
