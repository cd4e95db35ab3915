// Random draws of the walk path.
//
// Deterministic mode (config::deterministic_mode, config/globals.h:29): the reference seeds one
// utility::Random per walk with walk_id / n and restarts it at every re-walk (graph/wharfmh.h:289-291,
// 814-816), so the whole corpus consumes only the table X[r][k] = k-th lrand() of Random(r),
// r < walks_per_vertex, k < walk_length.  The table is produced on the host with the reference's exact
// arithmetic (signed-shift splitmix64 seeding + xoroshiro128+, utils/utility.h:152-206).
//
// Keyed MH mode: the reference draws from one unsynchronised global generator seeded by time()
// (config/globals.h:26), so nothing there can be reproduced; this engine keys every draw by what it is for.
#pragma once

#include "common.cuh"

#include <vector>

namespace wharf
{
#ifdef __CUDACC__
#define WHARF_HD __host__ __device__ __forceinline__
#else
#define WHARF_HD inline
#endif

// utility::Random::Random(seed) - utils/utility.h:157-170.  z is a signed long long there, so the three
// right shifts are ARITHMETIC.
inline void random_seed(u64 seed, u64 s[2])
{
    for (int i = 0; i < 2; i++)
    {
        seed += 0x9E3779B97F4A7C15ull;
        i64 z = (i64) seed;
        z = (i64) ((u64) (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull);
        z = (i64) ((u64) (z ^ (z >> 27)) * 0x94D049BB133111EBull);
        s[i] = (u64) (z ^ (z >> 31));
    }
}

// utility::Random::lrand - utils/utility.h:194-206 (xoroshiro128+, rotations 55/14/36).
inline u64 random_next(u64 s[2])
{
    const u64 s0 = s[0];
    u64 s1 = s[1];
    const u64 r = s0 + s1;
    s1 ^= s0;
    s[0] = ((s0 << 55) | (s0 >> 9)) ^ s1 ^ (s1 << 14);
    s[1] = (s1 << 36) | (s1 >> 28);
    return r;
}

inline std::vector<u64> make_draw_table(u32 replicas, u32 length)
{
    std::vector<u64> t((size_t) replicas * length);
    for (u32 r = 0; r < replicas; r++)
    {
        u64 s[2];
        random_seed(r, s);
        for (u32 k = 0; k < length; k++) t[(size_t) r * length + k] = random_next(s);
    }
    return t;
}

// ---- keyed counter generator (MH mode) -------------------------------------------------------------
WHARF_HD u64 sm64(u64 z)
{
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}

enum : u64 { kDomStep = 1, kDomState = 2, kDomSampler = 3, kDomChain = 4 };

WHARF_HD u64 keyed_draw(u64 seed, u64 dom, u64 a, u64 b, u64 c)
{
    u64 h = sm64(seed ^ (dom * 0xD6E8FEB86659FD93ull));
    h = sm64(h ^ a);
    h = sm64(h ^ b);
    h = sm64(h ^ c);
    return h;
}

// utility::Random::drand - utils/utility.h:208-218 applied to a 64-bit draw.
WHARF_HD double unit_double(u64 x)
{
    const u64 bits = (0x3FFull << 52) | (x >> 12);
#ifdef __CUDA_ARCH__
    return __longlong_as_double((long long) bits) - 1.0;
#else
    double d;
    __builtin_memcpy(&d, &bits, 8);
    return d - 1.0;
#endif
}

}  // namespace wharf
