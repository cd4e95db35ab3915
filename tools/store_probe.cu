// Streaming-store probe for B200: how fast can one GPU WRITE to HBM, and does the store flavour matter?  The corpus of a
// generate pass is written once (107 GB at scale-25), so this rate - not the copy rate - is the floor of the emit
// kernels.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/_build/store_probe tools/store_probe.cu
//   tools/_build/store_probe [GiB=8]
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { std::fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); std::exit(1); } } while (0)

typedef uint32_t u32;
typedef uint64_t u64;

enum Mode { kPlain = 0, kCs = 1, kCg = 2, kWt = 3, kEvictFirst = 4, kV8 = 5, kNoAllocHint = 6 };

template <int kMode>
__device__ __forceinline__ void store16(uint4* p, uint4 v)
{
    if (kMode == kPlain) *p = v;
    else if (kMode == kCs) __stcs(p, v);
    else if (kMode == kCg) __stcg(p, v);
    else if (kMode == kWt) __stwt(p, v);
    else if (kMode == kEvictFirst)
    {
        u64 pol;
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
    }
    else if (kMode == kNoAllocHint)
    {
        u64 pol;
        asm volatile("createpolicy.fractional.L2::evict_unchanged.b64 %0, 1.0;" : "=l"(pol));
        asm volatile("st.global.L2::cache_hint.v4.u32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w), "l"(pol) : "memory");
    }
}

// Every thread writes `per` consecutive-in-the-warp 16-byte pieces (grid-stride): fully coalesced.
template <int kMode>
__global__ void __launch_bounds__(256) store_kernel(uint4* out, u64 count, u32 salt)
{
    const u64 stride = (u64) gridDim.x * 256;
    for (u64 i = (u64) blockIdx.x * 256 + threadIdx.x; i < count; i += stride) store16<kMode>(out + i, make_uint4((u32) i, salt, salt, (u32) i));
}

// 256-bit stores (st.global.v8.b32, sm_100).
__global__ void __launch_bounds__(256) store_v8_kernel(uint4* out, u64 count32, u32 salt)
{
    const u64 stride = (u64) gridDim.x * 256;
    for (u64 i = (u64) blockIdx.x * 256 + threadIdx.x; i < count32; i += stride)
        asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(out + 2 * i), "r"((u32) i), "r"(salt), "r"(salt), "r"(salt), "r"(salt), "r"(salt),
                     "r"(salt), "r"((u32) i)
                     : "memory");
}

// The emit pattern: thread = column, stores one u32 into each of `rows` rows that are `row_stride` words apart.
template <int kRows, bool kStream>
__global__ void __launch_bounds__(256) rows_kernel(u32* out, u64 columns, u64 row_stride, u32 salt)
{
    const u64 j = (u64) blockIdx.x * 256 + threadIdx.x;
    if (j >= columns) return;
#pragma unroll
    for (int r = 0; r < kRows; r++)
    {
        if (kStream) __stcs(out + (u64) r * row_stride + j, salt + r);
        else out[(u64) r * row_stride + j] = salt + r;
    }
}

// The same with kVec adjacent columns per thread: one 8- or 16-byte store per row (a warp covers 256 / 512 bytes of a row).
template <int kRows, int kVec>
__global__ void __launch_bounds__(256) rowsv_kernel(u32* out, u64 columns, u64 row_stride, u32 salt)
{
    const u64 j = ((u64) blockIdx.x * 256 + threadIdx.x) * kVec;
    if (j >= columns) return;
#pragma unroll
    for (int r = 0; r < kRows; r++)
    {
        u32* p = out + (u64) r * row_stride + j;
        if (kVec == 4) __stcs(reinterpret_cast<uint4*>(p), make_uint4(salt, salt + r, salt, salt));
        else __stcs(reinterpret_cast<uint2*>(p), make_uint2(salt, salt + r));
    }
}

// TMA bulk stores: a block fills 16 KiB of shared memory once and streams it out with cp.async.bulk (1-D).
__global__ void __launch_bounds__(128) bulk_kernel(char* out, u64 bytes, u32 salt)
{
    constexpr u32 kTile = 16384;
    __shared__ __align__(128) char tile[kTile];
    for (u32 i = threadIdx.x; i < kTile / 4; i += 128) reinterpret_cast<u32*>(tile)[i] = salt + i;
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0)
    {
        const u32 smem = (u32) __cvta_generic_to_shared(tile);
        u32 pending = 0;
        for (u64 off = (u64) blockIdx.x * kTile; off + kTile <= bytes; off += (u64) gridDim.x * kTile)
        {
            asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(out + off), "r"(smem), "r"(kTile) : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
            if (++pending >= 8) { asm volatile("cp.async.bulk.wait_group.read 4;" ::: "memory"); pending = 4; }
        }
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

template <class F>
double best_ms(F&& f, int reps = 5)
{
    cudaEvent_t a, b;
    CK(cudaEventCreate(&a));
    CK(cudaEventCreate(&b));
    double best = 1e30;
    for (int i = 0; i < reps; i++)
    {
        CK(cudaEventRecord(a));
        f();
        CK(cudaEventRecord(b));
        CK(cudaEventSynchronize(b));
        float ms;
        CK(cudaEventElapsedTime(&ms, a, b));
        if (ms < best) best = ms;
    }
    CK(cudaGetLastError());
    return best;
}

int main(int argc, char** argv)
{
    const u64 gib = argc > 1 ? std::atoll(argv[1]) : 8;
    const u64 bytes = gib << 30;
    char* buf;
    CK(cudaMalloc(&buf, bytes));
    int sms = 148;
    CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
    auto report = [&](const char* name, double ms, u64 b) { std::printf("%-44s %8.3f ms  %8.1f GB/s\n", name, ms, b / ms / 1e6); std::fflush(stdout); };
    uint4* v = reinterpret_cast<uint4*>(buf);
    const u64 n16 = bytes / 16;
    report("cudaMemsetAsync", best_ms([&] { CK(cudaMemsetAsync(buf, 1, bytes)); }), bytes);
    for (int bps : {4, 8, 16, 32})
    {
        const unsigned grid = sms * bps;
        char name[96];
        std::snprintf(name, sizeof name, "st.v4 plain, %d blocks/SM grid-stride", bps);
        report(name, best_ms([&] { store_kernel<kPlain><<<grid, 256>>>(v, n16, 1); }), bytes);
    }
    const unsigned grid = sms * 8;
    report("st.v4 plain, one piece per thread", best_ms([&] { store_kernel<kPlain><<<(unsigned) (n16 / 256), 256>>>(v, n16, 2); }), bytes);
    report("st.cs.v4", best_ms([&] { store_kernel<kCs><<<grid, 256>>>(v, n16, 3); }), bytes);
    report("st.cg.v4", best_ms([&] { store_kernel<kCg><<<grid, 256>>>(v, n16, 4); }), bytes);
    report("st.wt.v4", best_ms([&] { store_kernel<kWt><<<grid, 256>>>(v, n16, 5); }), bytes);
    report("st.v4 L2::evict_first", best_ms([&] { store_kernel<kEvictFirst><<<grid, 256>>>(v, n16, 6); }), bytes);
    report("st.v4 L2::evict_unchanged", best_ms([&] { store_kernel<kNoAllocHint><<<grid, 256>>>(v, n16, 7); }), bytes);
    report("st.v8 (256-bit)", best_ms([&] { store_v8_kernel<<<grid, 256>>>(v, bytes / 32, 8); }), bytes);
    report("cp.async.bulk 16 KiB tiles, 2 blocks/SM", best_ms([&] { bulk_kernel<<<sms * 2, 128>>>(buf, bytes, 9); }), bytes);
    report("cp.async.bulk 16 KiB tiles, 4 blocks/SM", best_ms([&] { bulk_kernel<<<sms * 4, 128>>>(buf, bytes, 9); }), bytes);
    // emit pattern: thread = column(s), one store into each of R rows that are a whole row apart
    {
        u32* w = reinterpret_cast<u32*>(buf);
#define ROWS_CASE(R)                                                                                                                          \
        {                                                                                                                                     \
            const u64 columns = (bytes / 4 / R) & ~(u64) 1023;                                                                                \
            const u64 total = columns * R * 4;                                                                                                \
            report("rows: " #R " x u32 per thread, st", best_ms([&] { rows_kernel<R, false><<<(unsigned) (columns / 256), 256>>>(w, columns, columns, 1); }), total);      \
            report("rows: " #R " x u32 per thread, st.cs", best_ms([&] { rows_kernel<R, true><<<(unsigned) (columns / 256), 256>>>(w, columns, columns, 1); }), total);    \
            report("rows: " #R " x uint2 per thread, st.cs", best_ms([&] { rowsv_kernel<R, 2><<<(unsigned) (columns / 512), 256>>>(w, columns, columns, 1); }), total);    \
            report("rows: " #R " x uint4 per thread, st.cs", best_ms([&] { rowsv_kernel<R, 4><<<(unsigned) (columns / 1024), 256>>>(w, columns, columns, 1); }), total);   \
        }
        ROWS_CASE(1) ROWS_CASE(4) ROWS_CASE(8) ROWS_CASE(16) ROWS_CASE(32) ROWS_CASE(80)
#undef ROWS_CASE
    }
    // reads for comparison
    CK(cudaFree(buf));
    return 0;
}
