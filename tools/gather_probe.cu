// Random-gather probe for B200: how many independent/dependent random sector loads per second does the
// memory system deliver, as a function of table size, load width, chains per thread, cache policy and
// cudaLimitMaxL2FetchGranularity?  The answers set the denominators of DESIGN.md section 3 (a walk step is a
// chain of dependent random loads).
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/gather_probe tools/gather_probe.cu
//   tools/_build/gather_probe [table_MiB ...]
//
// Every kernel runs `threads` walkers x K chains for `steps` dependent hops:  idx = mix(tbl[idx], salt) % N.
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>
#include <vector>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); std::exit(1); } } while (0)

typedef uint32_t u32;
typedef uint64_t u64;

__device__ __forceinline__ u32 mix(u32 x) { x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x; }

enum Policy { kNc = 0, kCa = 1, kCg = 2, kCs = 3, kNoAlloc = 4, kEvictFirst = 5 };

template <int kPolicy>
__device__ __forceinline__ u32 load32(const u32* p)
{
    u32 v;
    if (kPolicy == kNc) v = __ldg(p);
    else if (kPolicy == kCa) v = *((const volatile u32*) p);                 // plain ld.global (volatile keeps it a real load)
    else if (kPolicy == kCg) v = __ldcg(p);
    else if (kPolicy == kCs) v = __ldcs(p);
    else if (kPolicy == kNoAlloc) asm volatile("ld.global.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p));
    else
    {
        u64 pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("ld.global.L1::no_allocate.L2::cache_hint.u32 %0, [%1], %2;" : "=r"(v) : "l"(p), "l"(pol));
    }
    return v;
}

template <int kPolicy>
__device__ __forceinline__ uint4 load128(const uint4* p)
{
    uint4 v;
    if (kPolicy == kNc) v = __ldg(p);
    else if (kPolicy == kCg) v = __ldcg(p);
    else if (kPolicy == kCs) v = __ldcs(p);
    else if (kPolicy == kNoAlloc) asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    else if (kPolicy == kEvictFirst)
    {
        u64 pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("ld.global.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p), "l"(pol));
    }
    else v = *p;
    return v;
}

// one-level chase, 4-byte loads
template <int K, int kPolicy>
__global__ void __launch_bounds__(256) chase32(const u32* __restrict__ tbl, u32 mask, int steps, u32* __restrict__ sink)
{
    const u32 t = blockIdx.x * 256 + threadIdx.x;
    u32 idx[K];
#pragma unroll
    for (int k = 0; k < K; k++) idx[k] = mix(t * K + k + 1) & mask;
    for (int s = 0; s < steps; s++)
    {
        u32 v[K];
#pragma unroll
        for (int k = 0; k < K; k++) v[k] = load32<kPolicy>(tbl + idx[k]);
#pragma unroll
        for (int k = 0; k < K; k++) idx[k] = mix(v[k] + t + s) & mask;
    }
    u32 acc = 0;
#pragma unroll
    for (int k = 0; k < K; k++) acc ^= idx[k];
    if (acc == 0x12345678u) sink[0] = acc;
}

// one-level chase, 16-byte loads (index counts uint4 entries)
template <int K, int kPolicy>
__global__ void __launch_bounds__(256) chase128(const uint4* __restrict__ tbl, u32 mask, int steps, u32* __restrict__ sink)
{
    const u32 t = blockIdx.x * 256 + threadIdx.x;
    u32 idx[K];
#pragma unroll
    for (int k = 0; k < K; k++) idx[k] = mix(t * K + k + 1) & mask;
    for (int s = 0; s < steps; s++)
    {
        uint4 v[K];
#pragma unroll
        for (int k = 0; k < K; k++) v[k] = load128<kPolicy>(tbl + idx[k]);
#pragma unroll
        for (int k = 0; k < K; k++) idx[k] = mix(v[k].x + v[k].w + t + s) & mask;
    }
    u32 acc = 0;
#pragma unroll
    for (int k = 0; k < K; k++) acc ^= idx[k];
    if (acc == 0x12345678u) sink[0] = acc;
}

// two-level chase like a CSR walk step: 8-byte descriptor (random) then 4-byte neighbour (random), plus a coalesced store
template <int K, int kPolicy>
__global__ void __launch_bounds__(256) chase_csr(const uint2* __restrict__ desc, u32 dmask, const u32* __restrict__ pool, u32 pmask, int steps,
                                                 u32* __restrict__ out, u64 stride)
{
    const u32 t = blockIdx.x * 256 + threadIdx.x;
    u32 cur[K];
#pragma unroll
    for (int k = 0; k < K; k++) cur[k] = mix(t * K + k + 1) & dmask;
    u32* o = out + t;
    for (int s = 0; s < steps; s++)
    {
        uint2 d[K];
#pragma unroll
        for (int k = 0; k < K; k++) d[k] = kPolicy == kNc ? __ldg(desc + cur[k]) : __ldcg(desc + cur[k]);
        u32 v[K];
#pragma unroll
        for (int k = 0; k < K; k++) v[k] = load32<kPolicy>(pool + ((d[k].x + (mix(t + s) & 7)) & pmask));
#pragma unroll
        for (int k = 0; k < K; k++) { cur[k] = v[k] & dmask; o[(u64) k * stride] = cur[k]; }
        o += (u64) K * stride;
    }
}

__global__ void fill_random(u32* p, u64 n, u32 salt)
{
    for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (u64) gridDim.x * blockDim.x) p[i] = mix((u32) i ^ salt) ^ mix((u32) (i >> 32) + salt);
}

struct Timer
{
    cudaEvent_t a, b;
    Timer() { CK(cudaEventCreate(&a)); CK(cudaEventCreate(&b)); }
    void start() { CK(cudaEventRecord(a)); }
    float stop() { CK(cudaEventRecord(b)); CK(cudaEventSynchronize(b)); float ms; CK(cudaEventElapsedTime(&ms, a, b)); return ms; }
};

static u32* g_sink;
static Timer* g_t;
static int g_sms = 148;

template <int K, int kPolicy>
void run32(const char* name, const u32* tbl, u64 entries, int blocks_per_sm, int steps)
{
    const int grid = g_sms * blocks_per_sm;
    chase32<K, kPolicy><<<grid, 256>>>(tbl, (u32) (entries - 1), 4, g_sink);
    g_t->start();
    chase32<K, kPolicy><<<grid, 256>>>(tbl, (u32) (entries - 1), steps, g_sink);
    const float ms = g_t->stop();
    const double loads = (double) grid * 256 * K * steps;
    std::printf("  %-28s K=%d bps=%d  %8.3f ms  %7.2f Gload/s  (x32B = %7.1f GB/s)\n", name, K, blocks_per_sm, ms, loads / ms / 1e6, loads * 32 / ms / 1e6);
}

template <int K, int kPolicy>
void run128(const char* name, const uint4* tbl, u64 entries, int blocks_per_sm, int steps)
{
    const int grid = g_sms * blocks_per_sm;
    chase128<K, kPolicy><<<grid, 256>>>(tbl, (u32) (entries - 1), 4, g_sink);
    g_t->start();
    chase128<K, kPolicy><<<grid, 256>>>(tbl, (u32) (entries - 1), steps, g_sink);
    const float ms = g_t->stop();
    const double loads = (double) grid * 256 * K * steps;
    std::printf("  %-28s K=%d bps=%d  %8.3f ms  %7.2f Gload/s  (x32B = %7.1f GB/s)\n", name, K, blocks_per_sm, ms, loads / ms / 1e6, loads * 32 / ms / 1e6);
}

template <int K, int kPolicy>
void run_csr(const char* name, const uint2* desc, u64 dn, const u32* pool, u64 pn, u32* out, int blocks_per_sm, int steps)
{
    const int grid = g_sms * blocks_per_sm;
    const u64 stride = (u64) grid * 256;
    chase_csr<K, kPolicy><<<grid, 256>>>(desc, (u32) (dn - 1), pool, (u32) (pn - 1), 2, out, stride);
    g_t->start();
    chase_csr<K, kPolicy><<<grid, 256>>>(desc, (u32) (dn - 1), pool, (u32) (pn - 1), steps, out, stride);
    const float ms = g_t->stop();
    const double st = (double) grid * 256 * K * steps;
    std::printf("  %-28s K=%d bps=%d  %8.3f ms  %7.2f Gstep/s  (x68B = %7.1f GB/s)\n", name, K, blocks_per_sm, ms, st / ms / 1e6, st * 68 / ms / 1e6);
}

int main(int argc, char** argv)
{
    std::vector<u64> sizes_mib;
    for (int i = 1; i < argc; i++) sizes_mib.push_back(std::strtoull(argv[i], nullptr, 10));
    if (sizes_mib.empty()) sizes_mib = {64, 4096};
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, 0));
    g_sms = prop.multiProcessorCount;
    std::printf("device: %s, %d SMs, L2 %d MiB\n", prop.name, g_sms, prop.l2CacheSize >> 20);
    CK(cudaMalloc(&g_sink, 64));
    Timer timer;
    g_t = &timer;
    const int steps = 200;

    for (u64 mib : sizes_mib)
    {
        const u64 bytes = mib << 20;
        const u64 entries = bytes / 4;
        u32* tbl;
        CK(cudaMalloc(&tbl, bytes));
        fill_random<<<g_sms * 8, 256>>>(tbl, entries, 17);
        CK(cudaDeviceSynchronize());
        for (size_t gran : {(size_t) 0, (size_t) 32, (size_t) 64, (size_t) 128})
        {
            if (gran) { cudaError_t e = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran); if (e != cudaSuccess) { std::printf("set limit %zu: %s\n", gran, cudaGetErrorString(e)); cudaGetLastError(); } }
            size_t got = 0;
            cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
            std::printf("== table %llu MiB, MaxL2FetchGranularity requested %zu -> %zu\n", (unsigned long long) mib, gran, got);
            run32<1, kNc>("u32 ld.nc", tbl, entries, 8, steps);
            run32<2, kNc>("u32 ld.nc", tbl, entries, 8, steps);
            run32<4, kNc>("u32 ld.nc", tbl, entries, 8, steps);
            run32<8, kNc>("u32 ld.nc", tbl, entries, 8, steps);
            run32<4, kNc>("u32 ld.nc", tbl, entries, 4, steps);
            run32<8, kNc>("u32 ld.nc", tbl, entries, 4, steps);
            run32<4, kCa>("u32 ld.ca", tbl, entries, 8, steps);
            run32<4, kCg>("u32 ld.cg", tbl, entries, 8, steps);
            run32<4, kCs>("u32 ld.cs", tbl, entries, 8, steps);
            run32<4, kNoAlloc>("u32 L1::no_allocate", tbl, entries, 8, steps);
            run32<4, kEvictFirst>("u32 no_alloc+L2::evict_first", tbl, entries, 8, steps);
            run128<1, kNc>("u128 ld.nc", (const uint4*) tbl, entries / 4, 8, steps);
            run128<2, kNc>("u128 ld.nc", (const uint4*) tbl, entries / 4, 8, steps);
            run128<4, kNc>("u128 ld.nc", (const uint4*) tbl, entries / 4, 8, steps);
            run128<4, kCg>("u128 ld.cg", (const uint4*) tbl, entries / 4, 8, steps);
            run128<4, kEvictFirst>("u128 no_alloc+evict_first", (const uint4*) tbl, entries / 4, 8, steps);
        }
        // CSR-like two-level walk step: desc = first 1/8 of the table (as uint2), pool = the whole table
        {
            const u64 dn = entries / 16;       // uint2 entries in the first eighth
            u32* out;
            const u64 out_entries = (u64) g_sms * 8 * 256 * 4 * 64;
            CK(cudaMalloc(&out, out_entries * 4));
            for (size_t gran : {(size_t) 32, (size_t) 64})
            {
                cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, gran);
                cudaGetLastError();
                std::printf("== csr-like step, table %llu MiB, granularity %zu\n", (unsigned long long) mib, gran);
                run_csr<1, kNc>("desc+pool ld.nc", (const uint2*) tbl, dn, tbl, entries, out, 8, 64);
                run_csr<2, kNc>("desc+pool ld.nc", (const uint2*) tbl, dn, tbl, entries, out, 8, 64);
                run_csr<4, kNc>("desc+pool ld.nc", (const uint2*) tbl, dn, tbl, entries, out, 8, 64);
                run_csr<4, kCg>("desc+pool ld.cg", (const uint2*) tbl, dn, tbl, entries, out, 8, 64);
            }
            CK(cudaFree(out));
        }
        CK(cudaFree(tbl));
    }
    return 0;
}
