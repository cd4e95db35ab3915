// Szudzik pairing (walks/pairings.h:27-128) with an exact integer square root.
//
// The reference unpairs with floor(sqrt((double) z)) and a single fix-down (pairings.h:102-105), which is
// exact on the domain its tests exercise (operands < 2^16 for 32-bit codes, < 2^32 for 64-bit codes with
// z < 2^53) and can be off by one above it; these functions are exact on the whole 64-bit range and equal
// to the reference wherever the reference is right.
#pragma once

#include "common.cuh"

namespace wharf
{
#ifdef __CUDACC__
#define WHARF_PAIR_HD __host__ __device__ __forceinline__
#else
#define WHARF_PAIR_HD inline
#endif

// floor(sqrt(z)) exactly.
WHARF_PAIR_HD u64 isqrt_u64(u64 z)
{
#ifdef __CUDA_ARCH__
    u64 s = (u64) __dsqrt_rz(__ull2double_rz(z));
#else
    u64 s = (u64) __builtin_sqrt((double) z);
#endif
    if (s > 0xFFFFFFFFull) s = 0xFFFFFFFFull;
    while (s * s > z) s--;
    while (s < 0xFFFFFFFFull && (s + 1) * (s + 1) <= z) s++;
    return s;
}

// pair(x, y) = y >= x ? y*(y+1) + x : x*x + y   (pairings.h:27-86).  *overflow is set when the code
// does not fit 64 bits (the reference wraps silently).
WHARF_PAIR_HD u64 szudzik_pair(u64 x, u64 y, bool* overflow)
{
    const u64 big = y >= x ? y : x;
    if (overflow) *overflow = big > 0xFFFFFFFFull;       // big^2 + big + small < 2^64  <=>  big < 2^32
    return y >= x ? y * (y + 1) + x : x * x + y;
}

// unpair (pairings.h:100-114)
WHARF_PAIR_HD void szudzik_unpair(u64 z, u64* x, u64* y)
{
    const u64 s = isqrt_u64(z);
    const u64 t = z - s * s;
    if (t < s) { *x = s; *y = t; } else { *x = t - s; *y = s; }
}

WHARF_PAIR_HD u32 szudzik_pair32(u32 x, u32 y, bool* overflow)
{
    const u32 big = y >= x ? y : x;
    if (overflow) *overflow = big > 0xFFFFu;
    return y >= x ? y * (y + 1) + x : x * x + y;
}

WHARF_PAIR_HD void szudzik_unpair32(u32 z, u32* x, u32* y)
{
    const u32 s = (u32) isqrt_u64(z);
    const u32 t = z - s * s;
    if (t < s) { *x = s; *y = t; } else { *x = t - s; *y = s; }
}

}  // namespace wharf
