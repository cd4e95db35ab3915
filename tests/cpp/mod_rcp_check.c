/* CPU restatement of mod_u64_u32_rcp (csrc/common.cuh): the device intrinsics __drcp_rn, __dmul_rz, __ull2double_rz, __double2ull_rz
 * become plain double arithmetic under fesetround(); checked against x % d on 30 M cases incl. the edge divisors. */
#include <fenv.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
typedef uint64_t u64; typedef uint32_t u32; typedef int64_t i64;
static u64 s = 88172645463325252ull;
static u64 rnd(void) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
static u32 mod_rcp(u64 x, u32 d, double inv)
{
    u32 hi = (u32) (x >> 32);
    volatile double p1 = (double) hi * inv;          /* FE_TOWARDZERO set: dmul_rz */
    u32 q1 = (u32) p1;
    i64 r1 = (i64) hi - (i64) ((u64) q1 * d);
    if (r1 < 0) r1 += d;
    if (r1 >= (i64) d) r1 -= d;
    if (r1 < 0 || r1 >= (i64) d) return 0xFFFFFFFFu;
    u64 t = ((u64) r1 << 32) | (u32) x;
    volatile double td = (double) t;                 /* ull2double_rz */
    volatile double p2 = td * inv;
    u64 q = (u64) p2;
    i64 rem = (i64) (t - q * (u64) d);
    if (rem < 0) rem += d;
    if (rem >= (i64) d) rem -= d;
    if (rem < 0 || rem >= (i64) d) return 0xFFFFFFFFu;
    return (u32) rem;
}
int main(void)
{
    u64 bad = 0, n = 0;
    u32 ds[] = {1, 2, 3, 5, 7, 16, 17, 255, 256, 257, 65535, 65536, 65537, 1000003, 0x7FFFFFFF, 0x80000000u, 0x80000001u, 0xFFFFFFFEu, 0xFFFFFFFFu};
    for (int pass = 0; pass < 30000000; pass++)
    {
        u32 d = pass < 19 * 1000 ? ds[pass % 19] : (u32) (rnd() >> (rnd() % 32)) ;
        if (d == 0) d = 1;
        u64 x = rnd();
        if (pass % 7 == 0) x = ~0ull - (rnd() & 0xFF);
        if (pass % 11 == 0) x = (u64) d * (rnd() >> 32) + (pass % 3 == 0 ? d - 1 : 0);
        fesetround(FE_TONEAREST);
        volatile double inv = 1.0 / (double) d;      /* drcp_rn */
        fesetround(FE_TOWARDZERO);
        u32 got = mod_rcp(x, d, inv);
        n++;
        if (got != (u32) (x % d)) { if (bad < 5) printf("BAD x=%llu d=%u got=%u want=%u\n", (unsigned long long) x, d, got, (u32) (x % d)); bad++; }
    }
    printf("%llu cases, %llu bad\n", (unsigned long long) n, (unsigned long long) bad);
    return bad != 0;
}
