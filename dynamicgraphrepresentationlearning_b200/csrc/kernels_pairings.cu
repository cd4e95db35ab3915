// Szudzik pair / unpair / triplet over device arrays: walks/pairings.h:27,88,100,122 as fused elementwise ops.
#include "engine.hpp"
#include "pairings.cuh"

namespace wharf
{
namespace
{
constexpr int kBlock = 256;

__global__ void pair_u64_kernel(const u64* __restrict__ x, const u64* __restrict__ y, u64* __restrict__ out, u64 count, unsigned long long* overflow)
{
    u32 bad = 0;
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        bool o;
        out[i] = szudzik_pair(x[i], y[i], &o);
        bad += o;
    }
    block_add_u64(overflow, bad);
}

__global__ void unpair_u64_kernel(const u64* __restrict__ z, u64* __restrict__ x, u64* __restrict__ y, u64 count)
{
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        u64 a, b;
        szudzik_unpair(z[i], &a, &b);
        x[i] = a;
        y[i] = b;
    }
}

__global__ void pair_u32_kernel(const u32* __restrict__ x, const u32* __restrict__ y, u32* __restrict__ out, u64 count, unsigned long long* overflow)
{
    u32 bad = 0;
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        bool o;
        out[i] = szudzik_pair32(x[i], y[i], &o);
        bad += o;
    }
    block_add_u64(overflow, bad);
}

__global__ void unpair_u32_kernel(const u32* __restrict__ z, u32* __restrict__ x, u32* __restrict__ y, u64 count)
{
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        u32 a, b;
        szudzik_unpair32(z[i], &a, &b);
        x[i] = a;
        y[i] = b;
    }
}

// pair_triplet(a, b, c) = pair(pair(a, b), c)   (pairings.h:88-91)
__global__ void pair_triplet_u64_kernel(const u64* __restrict__ a, const u64* __restrict__ b, const u64* __restrict__ c, u64* __restrict__ out,
                                        u64 count, unsigned long long* overflow)
{
    u32 bad = 0;
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        bool o1, o2;
        const u64 ab = szudzik_pair(a[i], b[i], &o1);
        out[i] = szudzik_pair(ab, c[i], &o2);
        bad += (o1 || o2);
    }
    block_add_u64(overflow, bad);
}

// unpair_triplet (pairings.h:122-128)
__global__ void unpair_triplet_u64_kernel(const u64* __restrict__ z, u64* __restrict__ a, u64* __restrict__ b, u64* __restrict__ c, u64 count)
{
    for (u64 i = (u64) blockIdx.x * kBlock + threadIdx.x; i < count; i += (u64) gridDim.x * kBlock)
    {
        u64 f, s, x, y;
        szudzik_unpair(z[i], &f, &s);
        szudzik_unpair(f, &x, &y);
        a[i] = x;
        b[i] = y;
        c[i] = s;
    }
}

unsigned grid(const LaunchCtx& lc, u64 count) { return grid_for(count, kBlock, 8, lc.sm_count); }
}  // namespace

#define WHARF_LAUNCH_PAIR(kernel, ...)                                   \
    do {                                                                 \
        if (count == 0) return;                                          \
        kernel<<<grid(lc, count), kBlock, 0, lc.stream>>>(__VA_ARGS__);  \
        lc.count();                                                      \
        WHARF_CUDA(cudaGetLastError());                                  \
    } while (0)

void szudzik_pair_u64(const LaunchCtx& lc, const u64* x, const u64* y, u64* out, u64 count, unsigned long long* d_overflow)
{ WHARF_LAUNCH_PAIR(pair_u64_kernel, x, y, out, count, d_overflow); }
void szudzik_unpair_u64(const LaunchCtx& lc, const u64* z, u64* x, u64* y, u64 count)
{ WHARF_LAUNCH_PAIR(unpair_u64_kernel, z, x, y, count); }
void szudzik_pair_u32(const LaunchCtx& lc, const u32* x, const u32* y, u32* out, u64 count, unsigned long long* d_overflow)
{ WHARF_LAUNCH_PAIR(pair_u32_kernel, x, y, out, count, d_overflow); }
void szudzik_unpair_u32(const LaunchCtx& lc, const u32* z, u32* x, u32* y, u64 count)
{ WHARF_LAUNCH_PAIR(unpair_u32_kernel, z, x, y, count); }
void szudzik_pair_triplet_u64(const LaunchCtx& lc, const u64* a, const u64* b, const u64* c, u64* out, u64 count, unsigned long long* d_overflow)
{ WHARF_LAUNCH_PAIR(pair_triplet_u64_kernel, a, b, c, out, count, d_overflow); }
void szudzik_unpair_triplet_u64(const LaunchCtx& lc, const u64* z, u64* a, u64* b, u64* c, u64 count)
{ WHARF_LAUNCH_PAIR(unpair_triplet_u64_kernel, z, a, b, c, count); }

}  // namespace wharf
