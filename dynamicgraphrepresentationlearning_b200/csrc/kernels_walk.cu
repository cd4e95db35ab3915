// Walk kernels: bulk generation, re-walk of affected walks, corpus scan, retrieval.
//
// Replaces graph/wharfmh.h:275-326 (generate_initial_random_walks hot loop A), :761-859
// (batch_walk_update insert_job), walks/models/{deepwalk,node2vec}.h, walks/metropolis_hastings_sampler.h
// and the retrieval of :365-427.
//
// Layout: the corpus is a position-major matrix walks[L][W_local] of u32 vertex ids, so that the
// one-walker-per-thread kernels store fully coalesced (lane i writes walks[pos][w0 + i]); a walker that
// stopped (isolated start / sink) leaves kNoVertex in the remaining cells.
#include <mutex>
#include <stdexcept>
#include "engine.hpp"
#include "rng.cuh"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include <cub/device/device_scan.cuh>
#include <cub/device/device_select.cuh>
#include <cub/iterator/counting_input_iterator.cuh>
#include <cub/iterator/transform_input_iterator.cuh>

namespace wharf
{
namespace
{
constexpr int kWalkBlock = 256;

struct DevWalkParams
{
    const RowDesc* __restrict__ desc;
    const u32* __restrict__ pool;
    const u32* __restrict__ version;
    u32* pool_rw;                     // the pool again, writable: aux words interleaved with the neighbours (MH node2vec)
    const u64* __restrict__ draw;
    u32* walks;
    u64 n;
    u64 w_local;
    u64 first;                        // first local walk of this launch (bulk generation in chunks)
    ShardMap map;                     // local walk -> (replica, start vertex)
    u32 L;
    u32 shift;                        // neighbour i of a row sits at pool[begin + (i << shift)] (1: interleaved with aux words)
    u32 coop_search;                  // 1: membership searches in rows of >= kCoopMin neighbours are done by the whole warp
    u32 spec_desc;                    // 1: load the proposal's descriptor before the accept decision, 0: after it (to_dev picks by graph size)
    u32 tri;                          // 1: the filter half of an aux word is the TRIANGLE block of its slot (kernels_graph.cu); 0: per-row blocks
    u32 init_strategy;
    float inv_p, inv_q;
    u64 seed;
};

__device__ __forceinline__ RowDesc load_desc(const RowDesc* __restrict__ desc, u32 v)
{
    const uint2 t = __ldg(reinterpret_cast<const uint2*>(desc) + v);
    return RowDesc{t.x, t.y};
}

// Neighbour i and its aux word in one 8-byte load.  The cache half of the aux word may be written by another walker
// while this kernel runs; a stale copy only means the value is computed again (it is a memo of a pure function), so the
// load may be served by L1 (WHARF_PAIR_LOAD: 0 = L2 only).
template <int kMode>
__device__ __forceinline__ uint2 load_pair_as(const u32* __restrict__ pool, RowDesc d, u32 i)
{
    const uint2* ptr = reinterpret_cast<const uint2*>(pool + d.begin + 2 * (u64) i);
    if constexpr (kMode == 0) return __ldcg(ptr);
    else if constexpr (kMode == 1) return *ptr;
    else return __ldg(ptr);
}

// Neighbour i of a row; with aux words (shift 1, MH node2vec) the aux word of neighbour i follows it.
__device__ __forceinline__ const u32* nbr_ptr(const DevWalkParams& p, RowDesc d, u32 i) { return p.pool + d.begin + ((u64) i << p.shift); }
__device__ __forceinline__ u32* aux_ptr(const DevWalkParams& p, RowDesc d, u32 i) { return p.pool_rw + d.begin + 2 * (u64) i + 1; }
#ifndef WHARF_PAIR_LOAD
#define WHARF_PAIR_LOAD 0
#endif
__device__ __forceinline__ uint2 load_pair(const DevWalkParams& p, RowDesc d, u32 i) { return load_pair_as<WHARF_PAIR_LOAD>(p.pool, d, i); }

// ---- deterministic rule (graph/wharfmh.h:304-309, :836-838) ----------------------------------------
// next = N(cur)[ X[walk_id / n][pos - start_pos] % deg(cur) ]
//
// kRewalk = false: thread t owns local walk t, starts at position 0 on its start vertex.
// kRewalk = true : thread t owns affected[t], resumes at positions[t] on the vertex stored there.
template <bool kRewalk>
__global__ void __launch_bounds__(kWalkBlock) walk_deterministic_kernel(DevWalkParams p, const u32* __restrict__ affected,
                                                                        const u32* __restrict__ positions, u64 count,
                                                                        unsigned long long* steps_out)
{
    const u64 t = (u64) blockIdx.x * kWalkBlock + threadIdx.x;
    u32 steps = 0;
    if (t < count)
    {
        u64 w;
        u32 pos, cur;
        if (kRewalk)
        {
            w   = __ldg(affected + t);
            pos = __ldg(positions + t);
            cur = p.walks[(u64) pos * p.w_local + w];
        }
        else
        {
            w   = p.first + t;
            pos = 0;
        }
        u32 replica, start;                                         // = walk_id / n, walk_id % n  (wharfmh.h:291-292)
        p.map.locate(w, replica, start);
        if (!kRewalk)
        {
            cur = start;
            p.walks[w] = cur;
        }
        const u64* __restrict__ X = p.draw + (u64) replica * p.L;   // stream restarts at the re-walk point (:814-816)
        u32* out = p.walks + (u64) (pos + 1) * p.w_local + w;
        const u32 first = pos;
        const u32 last = p.L - 1;
        bool stopped = false;
        for (; pos < last; ++pos, out += p.w_local)
        {
            const RowDesc d = load_desc(p.desc, cur);
            if (d.degree == 0) { stopped = true; break; }            // isolated start (:277-286) or sink
            const u32 j = mod_u64_u32(__ldg(X + (pos - first)), d.degree);
            cur = __ldg(nbr_ptr(p, d, j));
            *out = cur;
        }
        if (!stopped) stopped = load_desc(p.desc, cur).degree == 0;  // last cell: only the degree matters
        steps = pos - first + (stopped ? 0u : 1u);                   // the draw at the last position is unobservable (:315-317)
        for (++pos; pos <= last; ++pos, out += p.w_local) *out = kNoVertex;
    }
    block_add_u64(steps_out, steps);
}

// Bulk generation with kWalks independent walkers per thread (deterministic rule).  A walker step is two
// dependent random loads (row descriptor, then one neighbour): with one walker per thread the SM holds at
// most 2048 loads in flight, which leaves HBM latency-bound on graphs larger than L2.  Interleaving kWalks
// walkers per thread multiplies the loads in flight without more threads.  Thread t owns local walks
// t, t + T, t + 2T ... (T = threads in the grid) so every store stays coalesced.
template <int kWalks, int kBlock, bool kStreamStore>
__global__ void __launch_bounds__(kBlock) walk_deterministic_ilp_kernel(DevWalkParams p, u64 stride, unsigned long long* steps_out)
{
    const u64 t = (u64) blockIdx.x * kBlock + threadIdx.x;
    u32 cur[kWalks];
    u32 rep_off[kWalks];       // replica * L: offset of the walker's draw row
    bool alive[kWalks];
    u32 steps = 0;
#pragma unroll
    for (int k = 0; k < kWalks; k++)
    {
        const u64 w = t + (u64) k * stride;
        alive[k] = w < p.w_local;
        u32 replica = 0, start = 0;
        if (alive[k]) p.map.locate(w, replica, start);
        rep_off[k] = replica * p.L;
        cur[k] = start;
        if (alive[k]) p.walks[w] = cur[k];
    }
    bool live[kWalks];          // walker still moving (alive and not stopped on a degree-0 vertex)
#pragma unroll
    for (int k = 0; k < kWalks; k++) live[k] = alive[k];
    const u32 last = p.L - 1;
    u32* out = p.walks + p.w_local + t;
    for (u32 pos = 0; pos < last; ++pos, out += p.w_local)
    {
        RowDesc d[kWalks];
#pragma unroll
        for (int k = 0; k < kWalks; k++) d[k] = live[k] ? load_desc(p.desc, cur[k]) : RowDesc{0, 0};
        u32 nxt[kWalks];
#pragma unroll
        for (int k = 0; k < kWalks; k++)
        {
            nxt[k] = kNoVertex;
            if (d[k].degree)
            {
                const u32 j = mod_u64_u32(__ldg(p.draw + rep_off[k] + pos), d[k].degree);
                nxt[k] = __ldg(nbr_ptr(p, d[k], j));
            }
        }
#pragma unroll
        for (int k = 0; k < kWalks; k++)
        {
            if (alive[k])
            {
                if (kStreamStore) __stcs(out + (u64) k * stride, nxt[k]); else out[(u64) k * stride] = nxt[k];
            }
            steps += live[k] && d[k].degree;
            live[k] = live[k] && d[k].degree;
            cur[k] = nxt[k];
        }
    }
#pragma unroll
    for (int k = 0; k < kWalks; k++)
    {
        if (live[k]) live[k] = load_desc(p.desc, cur[k]).degree != 0;
        steps += live[k];                                      // the unobservable draw at the last position (:315-317)
    }
    block_add_u64(steps_out, steps);
}

// ---- Metropolis-Hastings mode ------------------------------------------------------------------------
//
// Sampler cache.  The sampler of state (cur, prev) lives in cur's SamplerManager (graph/vertex.h:12) and,
// because find() hands out copies (wharfmh.h:301,833), never advances past its init value
// (metropolis_hastings_sampler.h:69-108), a pure function of (seed, cur, prev, version[cur]) and the rows of cur
// and prev.  A walker in state (cur, prev) has just traversed the edge prev -> cur, whose slot in the row of
// prev it knows (it picked it), so the cache of the state is the low half of aux word `slot` of the row of prev
// (aux words: kernels_graph.cu):
//     bit  15      valid
//     bits 13..14  weight class of the init value with respect to prev (node2vec.h:74-88)
//     bit  12      0: bits 0..11 = slot of the init value in the row of cur (rows up to 4096 neighbours)
//                  1: bits 0..11 = index of the sampler draw that proposed it (the slot is that draw % degree)
// The cache dies with the row of prev (fresh tails are zero) and is emptied by aux_invalidate_kernel whenever
// cur is a batch source (row and version[cur] change, wharfmh.h:504->539).
constexpr u32 kNoSlot     = 0xFFFFFFFFu;
constexpr u32 kCacheValid = 0x8000u;
constexpr u32 kCacheByDraw = 0x1000u;
constexpr u32 kCacheDirectMax = 4096;     // largest degree whose slots fit bits 0..11
enum : u32 { kClassReturn = 0, kClassNeighbour = 1, kClassFar = 2 };   // 1/p, 1, 1/q

struct InitValue { u32 cls, slot, draw; };   // weight class, slot in the row of cur, index of the proposing draw

__device__ __forceinline__ u32 cache_encode(InitValue v, u32 degree_cur)
{
    return kCacheValid | (v.cls << 13) | (degree_cur <= kCacheDirectMax ? v.slot : (kCacheByDraw | v.draw));
}

__device__ __forceinline__ float class_weight(const DevWalkParams& p, u32 cls)
{
    return cls == kClassReturn ? p.inv_p : (cls == kClassNeighbour ? 1.0f : p.inv_q);
}

// Index of the last element <= key of an ascending row (0 when there is none); branch-free halving, one load per
// level.  *hit = whether that element equals key (std::binary_search of node2vec.h:112-119).
__device__ __forceinline__ u32 row_find(const u32* __restrict__ row, u32 n, u32 key, bool* hit, u32 shift)
{
    u32 base = 0;
    *hit = false;
    if (n == 0) return 0;
    while (n > 1)
    {
        const u32 half = n >> 1;
        if (__ldg(row + ((u64) (base + half) << shift)) <= key) base += half;
        n -= half;
    }
    *hit = __ldg(row + ((u64) base << shift)) == key;
    return base;
}

// Node2Vec::weight as a class - walks/models/node2vec.h:74-88; state = (cur, prev), dprev = row of prev, x = a neighbour
// of cur and x_aux the aux word of the slot cur -> x (it came with the 8-byte load that fetched x).
// Triangle blocks (p.tri; symmetric graphs): the filter half of x_aux is a Bloom block over N(cur) n N(x) hashed by member.
// prev is a neighbour of cur (the walker came from there), so a clear bit of hash(prev) proves prev is no neighbour of x,
// i.e. x is no neighbour of prev - decided by the load that was made anyway, without touching the row of prev.  Only a
// "maybe" (true common neighbours and ~1.5 % false positives) goes on to the exact search (node2vec.h:112-119).
// Per-row blocks (directed graphs): one 16-bit block of the row of prev, hashed by x, is loaded instead.
constexpr u32 kClassPending = 3;     // the filter could not rule x out: the exact search decides

__device__ __forceinline__ u32 n2v_prefilter(const DevWalkParams& p, u32 prev, RowDesc dprev, u32 x, u32 x_aux)
{
    if (x == prev) return kClassReturn;
    if (dprev.degree == 0) return kClassFar;
    if (p.tri)
    {
        const u32 need = filter_bits(filter_hash(prev));
        if (((x_aux >> 16) & need) != need) return kClassFar;                                               // proven absent
    }
    else
    {
        const u32 h = filter_hash(x);
        const u32 need = filter_bits(h);
        if (((__ldcg(aux_ptr(p, dprev, filter_block(h, dprev.degree))) >> 16) & need) != need) return kClassFar;   // proven absent
    }
    return kClassPending;
}

__device__ __forceinline__ u32 n2v_class(const DevWalkParams& p, u32 prev, RowDesc dprev, u32 x, u32 x_aux)
{
    const u32 cls = n2v_prefilter(p, prev, dprev, x, x_aux);
    if (cls != kClassPending) return cls;
    bool hit;
    row_find(p.pool + dprev.begin, dprev.degree, x, &hit, p.shift);
    return hit ? kClassNeighbour : kClassFar;
}

// std::binary_search(N(prev), x) (node2vec.h:112-119) by the WHOLE warp: 32 probes per round cut the range 33-fold, so a
// row of 10^5 neighbours takes 3 rounds + one final 32-wide compare instead of 17 dependent loads of one lane.  Every lane
// calls it with the same arguments.  Used for rows of kCoopMin neighbours and more when WHARF_COOP_SEARCH=1 (A/B:
// profiles/r02*_coop_search*.txt); the queries of a warp are then served one after the other.
constexpr u32 kCoopMin = 1024;
__device__ __forceinline__ bool row_find_warp(const u32* __restrict__ row, u32 n, u32 key, u32 shift, u32 lane)
{
    u32 lo = 0, hi = n;                                   // the element equal to key, if any, sits in [lo, hi)
    while (hi - lo > 32)
    {
        const u32 span = hi - lo;
        const u32 pos = lo + (u32) (((u64) (lane + 1) * span) / 33);            // lo < pos < hi
        const bool le = __ldg(row + ((u64) pos << shift)) <= key;
        const u32 k = __popc(__ballot_sync(0xffffffffu, le));                   // ascending row: the first k probes are <= key
        const u32 nlo = k ? lo + (u32) (((u64) k * span) / 33) : lo;
        const u32 nhi = k < 32 ? lo + (u32) (((u64) (k + 1) * span) / 33) : hi;
        lo = nlo;
        hi = nhi;
    }
    const u32 pos = lo + lane;
    const bool eq = pos < hi && __ldg(row + ((u64) pos << shift)) == key;
    return __any_sync(0xffffffffu, eq);
}

// MetropolisHastingsSampler::accept - walks/metropolis_hastings_sampler.h:118-122.
__device__ __forceinline__ bool mh_accept(float w_old, float w_new, u64 draw)
{
    if (w_old < w_new) return true;
    return unit_double(draw) <= (double) w_new / (double) w_old;
}

// MetropolisHastingsSampler::init, strategies RANDOM and BURNIN (metropolis_hastings_sampler.h:72-85), one lane;
// draws keyed by (seed, cur, prev, version[cur], i).
__device__ InitValue n2v_init_serial(const DevWalkParams& p, u32 cur, RowDesc dcur, u32 prev, RowDesc dprev)
{
    const u64 key_b = ((u64) prev << 32) | __ldg(p.version + cur);
    u64 i = 0;
    u32 draw = 0;
    u32 slot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, key_b, i++), dcur.degree);                 // RANDOM :72-75
    uint2 pair = load_pair(p, dcur, slot);
    u32 cls = n2v_class(p, prev, dprev, pair.x, pair.y);
    if (p.init_strategy == WMH_INIT_BURNIN)                                                                // :77-85
    {
        for (int it = 0; it < 100; it++)
        {
            const u32 cdraw = (u32) i;
            const u32 cslot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, key_b, i++), dcur.degree);
            pair = load_pair(p, dcur, cslot);
            const u32 ccls = n2v_class(p, prev, dprev, pair.x, pair.y);
            const float wn = class_weight(p, ccls), wo = class_weight(p, cls);
            bool acc = wo < wn;
            if (!acc) acc = unit_double(keyed_draw(p.seed, kDomSampler, cur, key_b, i++)) <= (double) wn / (double) wo;
            if (acc) { slot = cslot; cls = ccls; draw = cdraw; }
        }
    }
    return InitValue{cls, slot, draw};
}

// MetropolisHastingsSampler::init, strategy WEIGHT (:87-107): the best of 1 + 20 proposals, a strict '>' keeps
// the first maximum.  Warp-cooperative: the 21 proposals are independent keyed draws (i = 0..20), so a group of 8
// lanes serves one walker - lane `sub` evaluates proposals sub, sub + 8, sub + 16 (three independent load chains:
// neighbour, then filter block) and a butterfly over the group picks (largest weight, smallest i).  The four
// groups of a warp serve four walkers at once.  Every lane calls this with the state of its group's walker.
__device__ __forceinline__ InitValue n2v_init_weight_group(const DevWalkParams& p, u32 sub, bool has, u32 cur, RowDesc dcur, u32 prev, RowDesc dprev)
{
    float w = -1.0f;
    u32 word = 0, idx = 0xFFFFFFFFu;      // word = (class << 30) | slot
    if (has)
    {
        const u64 key_b = ((u64) prev << 32) | __ldg(p.version + cur);
        u32 slot[3], x[3], fh[3], fw[3];
        const double inv = reciprocal_of(dcur.degree);       // three proposals, one divisor
#pragma unroll
        for (u32 k = 0; k < 3; k++) slot[k] = mod_u64_u32_rcp(keyed_draw(p.seed, kDomSampler, cur, key_b, sub + 8 * k), dcur.degree, inv);
        if (p.tri)
        {
            // triangle blocks: the filter word of proposal x is the aux word of the slot cur -> x, tested with hash(prev)
            const u32 hp = filter_hash(prev);
#pragma unroll
            for (u32 k = 0; k < 3; k++)
            {
                const uint2 pair = load_pair(p, dcur, slot[k]);
                x[k] = pair.x;
                fw[k] = pair.y >> 16;
                fh[k] = hp;
            }
        }
        else
        {
#pragma unroll
            for (u32 k = 0; k < 3; k++) x[k] = __ldg(nbr_ptr(p, dcur, slot[k]));
#pragma unroll
            for (u32 k = 0; k < 3; k++)
            {
                fh[k] = filter_hash(x[k]);
                fw[k] = dprev.degree ? __ldcg(aux_ptr(p, dprev, filter_block(fh[k], dprev.degree))) >> 16 : 0u;
            }
        }
        // classes: return / far are decided by the filter alone; the few proposals the filter cannot rule out (common
        // neighbours and 2.5 % false positives) are searched afterwards, one per lane and trip - the search is a
        // chain of dependent loads, and a warp that branched into it once per proposal ran it three times per round
        u32 cls0 = kClassFar, cls1 = kClassFar, cls2 = kClassFar, pending = 0;
#pragma unroll
        for (u32 k = 0; k < 3; k++)
        {
            if (sub + 8 * k >= 21) break;
            const u32 need = filter_bits(fh[k]);
            if (x[k] == prev) { if (k == 0) cls0 = kClassReturn; else if (k == 1) cls1 = kClassReturn; else cls2 = kClassReturn; }
            else if (dprev.degree && (fw[k] & need) == need) pending |= 1u << k;
        }
        while (pending)
        {
            const u32 k = (u32) __ffs(pending) - 1u;
            pending &= pending - 1u;
            const u32 xk = k == 0 ? x[0] : (k == 1 ? x[1] : x[2]);
            bool hit;
            row_find(p.pool + dprev.begin, dprev.degree, xk, &hit, p.shift);
            if (hit) { if (k == 0) cls0 = kClassNeighbour; else if (k == 1) cls1 = kClassNeighbour; else cls2 = kClassNeighbour; }
        }
#pragma unroll
        for (u32 k = 0; k < 3; k++)
        {
            const u32 i = sub + 8 * k;
            if (i >= 21) break;
            const u32 cls = k == 0 ? cls0 : (k == 1 ? cls1 : cls2);
            const float wk = class_weight(p, cls);
            if (wk > w) { w = wk; word = (cls << 30) | slot[k]; idx = i; }
        }
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1)
    {
        const float ow = __shfl_xor_sync(0xffffffffu, w, o);
        const u32 oword = __shfl_xor_sync(0xffffffffu, word, o);
        const u32 oidx = __shfl_xor_sync(0xffffffffu, idx, o);
        if (ow > w || (ow == w && oidx < idx)) { w = ow; word = oword; idx = oidx; }
    }
    return InitValue{word >> 30, word & 0x3FFFFFFFu, idx};
}

__device__ __forceinline__ u32 warp_max_u32(u32 v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// node2vec in MH mode (node2vec.h:42-119 + metropolis_hastings_sampler.h:31-46), one walker per thread.
// The warp stays converged over a warp-uniform trip count (lanes without work are predicated off) so that the
// WEIGHT sampler init can be served by all 32 lanes.  Per step a walker issues: the proposal load, the cache
// word (address known from the previous step), one row search for the proposal's weight (per-thread binary
// search over the ascending row of prev; 32 independent chains per warp keep more loads in flight than a
// k-ary cooperative search that serialises the warp's queries), and the descriptor of the proposal
// (speculative: it is the next cur unless the proposal is rejected).  The descriptor of prev is carried.
// 64 registers -> 4 blocks of 256 per SM.  Capping registers for 5 or 6 resident blocks spills and is slower
// (measured at scale-20: 35.7 ms at 4 blocks, 37.7 ms at 5, 43.4 ms at 6).
template <bool kRewalk>
__global__ void __launch_bounds__(kWalkBlock, 4) walk_n2v_kernel(DevWalkParams p, const u32* __restrict__ affected,
                                                              const u32* __restrict__ positions, u64 count,
                                                              unsigned long long* steps_out)
{
    const u64 t = (u64) blockIdx.x * kWalkBlock + threadIdx.x;
    const u32 lane = threadIdx.x & 31;
    const bool owner = t < count;
    const u32 last_pos = p.L - 1;
    u64 w = 0, walk_id = 0;
    u32 pos = 0, cur = 0, prev = 0, slot = kNoSlot;        // slot: index of cur in the row of prev
    RowDesc d{0, 0}, dprev{0, 0};
    if (owner)
    {
        if (kRewalk) { w = __ldg(affected + t); pos = __ldg(positions + t); } else { w = p.first + t; }
        u32 replica, start;
        p.map.locate(w, replica, start);
        walk_id = (u64) replica * p.n + start;                       // global id keys the draws
        if (kRewalk) cur = p.walks[(u64) pos * p.w_local + w];
        else { cur = start; p.walks[w] = cur; }
        d = load_desc(p.desc, cur);
        prev = cur;
        if (pos > 0) prev = p.walks[(u64) (pos - 1) * p.w_local + w];                                  // wharfmh.h:819-823
        else if (d.degree) prev = __ldg(nbr_ptr(p, d, mod_u64_u32(keyed_draw(p.seed, kDomState, walk_id, 0, 0), d.degree)));   // node2vec.h:42-49
        dprev = load_desc(p.desc, prev);
        bool hit;
        const u32 i = row_find(p.pool + dprev.begin, dprev.degree, cur, &hit, p.shift);   // absent in a directed graph or after a delete
        if (hit) slot = i;
    }
    const u32 first = pos;
    u32* out = p.walks + (u64) (pos + 1) * p.w_local + w;
    const u32 trips = warp_max_u32(owner ? last_pos - pos : 0u);
    bool stopped = false, have_carried = false;
    u32 carried = 0;                                       // aux word that came with the load of cur
    for (u32 k = 0; k < trips; k++)
    {
        bool step = owner && !stopped && pos < last_pos;
        if (step && d.degree == 0) { stopped = true; step = false; }              // isolated start (:277-286) or sink
        u32 cand = 0, cslot = 0, cache = 0, cls_new = kClassFar, last_slot = kNoSlot, cand_aux = 0;
        RowDesc dcand{0, 0};
        if (step)
        {
            // cache of the state (cur, prev): the aux word next to cur in the row of prev - it came with the 8-byte load
            // that picked cur one step ago; only a walker's first step (or a re-walk's) has to fetch it
            if (slot != kNoSlot) cache = (have_carried ? carried : __ldcg(aux_ptr(p, dprev, slot))) & 0xFFFFu;
            cslot = mod_u64_u32(keyed_draw(p.seed, kDomStep, walk_id, pos, 0), d.degree);   // propose_vertex - node2vec.h:97-105
            const uint2 pair = load_pair(p, d, cslot);
            cand = pair.x;
            cand_aux = pair.y;
            // the accept decision needs the proposal's class, not its descriptor, and 70 % of the proposals are rejected: on
            // graphs whose descriptors do not fit the L2 it is loaded once the next vertex is known (to_dev)
            if (p.spec_desc) dcand = load_desc(p.desc, cand);
            cls_new = n2v_prefilter(p, prev, dprev, cand, cand_aux);
            if (cls_new == kClassPending && !(p.coop_search && dprev.degree >= kCoopMin))
            {
                bool hit;                                   // per-thread search: 32 independent chains per warp
                row_find(p.pool + dprev.begin, dprev.degree, cand, &hit, p.shift);
                cls_new = hit ? kClassNeighbour : kClassFar;
            }
        }
        if (p.coop_search)
        {
            u32 pending = __ballot_sync(0xffffffffu, step && cls_new == kClassPending);     // long rows, one query at a time
            while (pending)
            {
                const int src = __ffs(pending) - 1;
                pending &= pending - 1;
                const u32 b = __shfl_sync(0xffffffffu, dprev.begin, src), nn = __shfl_sync(0xffffffffu, dprev.degree, src);
                const u32 key = __shfl_sync(0xffffffffu, cand, src);
                const bool hit = row_find_warp(p.pool + b, nn, key, p.shift, lane);
                if ((int) lane == src) cls_new = hit ? kClassNeighbour : kClassFar;
            }
        }
        const bool need = step && !(cache & kCacheValid);
        InitValue iv{kClassFar, 0, 0};
        if (p.init_strategy == WMH_INIT_WEIGHT)
        {
            u32 pending = __ballot_sync(0xffffffffu, need);
            while (pending)                                   // four walkers per round, one per 8-lane group
            {
                u32 mine = pending;                           // pending with the first (lane / 8) bits dropped
                for (u32 g = lane >> 3; g && mine; --g) mine &= mine - 1;
                const bool has = mine != 0;
                const int src = has ? __ffs(mine) - 1 : 0;
                const u32 s_cur = __shfl_sync(0xffffffffu, cur, src), s_prev = __shfl_sync(0xffffffffu, prev, src);
                const RowDesc s_d{__shfl_sync(0xffffffffu, d.begin, src), __shfl_sync(0xffffffffu, d.degree, src)};
                const RowDesc s_dp{__shfl_sync(0xffffffffu, dprev.begin, src), __shfl_sync(0xffffffffu, dprev.degree, src)};
                const InitValue r = n2v_init_weight_group(p, lane & 7, has, s_cur, s_d, s_prev, s_dp);
                // a served walker is the k-th pending lane (k < 4): its result sits in every lane of group k
                const u32 my_rank = __popc(pending & ((1u << lane) - 1u));
                const int from = (int) ((my_rank & 3u) * 8u);
                const InitValue got{__shfl_sync(0xffffffffu, r.cls, from), __shfl_sync(0xffffffffu, r.slot, from), __shfl_sync(0xffffffffu, r.draw, from)};
                if (((pending >> lane) & 1u) && my_rank < 4) iv = got;
#pragma unroll
                for (int k = 0; k < 4; k++) pending &= pending - 1;
            }
        }
        else if (need) iv = n2v_init_serial(p, cur, d, prev, dprev);
        if (need)
        {
            cache = cache_encode(iv, d.degree);
            last_slot = iv.slot;
            if (slot != kNoSlot) __stcg(reinterpret_cast<unsigned short*>(aux_ptr(p, dprev, slot)), (unsigned short) cache);   // low half only
        }
        if (step)
        {
            const float wn = class_weight(p, cls_new);                                       // sample :31-46
            const float wo = class_weight(p, (cache >> 13) & 3u);
            bool acc = wo < wn;
            if (!acc) acc = unit_double(keyed_draw(p.seed, kDomStep, walk_id, pos, 1)) <= (double) wn / (double) wo;
            u32 next = cand;
            if (!acc)
            {
                if (last_slot == kNoSlot)
                {
                    last_slot = cache & 0xFFFu;
                    if (cache & kCacheByDraw)                                                 // big row: re-derive the slot from its draw
                        last_slot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, ((u64) prev << 32) | __ldg(p.version + cur), last_slot), d.degree);
                }
                cslot = last_slot;
                const uint2 pair = load_pair(p, d, cslot);
                next = pair.x;
                cand_aux = pair.y;
            }
            if (!acc || !p.spec_desc) dcand = load_desc(p.desc, next);
            __stcs(out, next);
            prev = cur; dprev = d; cur = next; d = dcand; slot = cslot;
            carried = cand_aux;
            have_carried = true;
            ++pos;
            out += p.w_local;
        }
    }
    u32 steps = 0;
    if (owner)
    {
        // the draw at the last position is unobservable (:315-317) but counts as a step when the walker can move
        steps = pos - first + ((!stopped && d.degree) ? 1u : 0u);
        for (++pos; pos <= last_pos; ++pos, out += p.w_local) __stcs(out, kNoVertex);
    }
    block_add_u64(steps_out, steps);
}

// MetropolisHastingsSampler::sample as a CHAIN (metropolis_hastings_sampler.h:31-46; fixture tests/sampler.cpp:86-127): one
// sampler object of the fixed state (cur, prev) sampled `iters` times, its stored vertex advancing with every accepted
// proposal (the walk kernels sample from copies, whose stored vertex never moves: see "Sampler cache" above).  One thread:
// the chain is sequential by definition.  Membership by plain binary search, so it runs on either row layout.
// counts[j] = samples that returned neighbour j of cur; *init_out = the init value (all three strategies, :69-108).
__global__ void sampler_chain_kernel(DevWalkParams p, u32 model, u32 cur, u32 prev, u64 iters, u32* __restrict__ init_out,
                                     unsigned long long* __restrict__ counts)
{
    if (blockIdx.x || threadIdx.x) return;
    const RowDesc d = load_desc(p.desc, cur), dprev = load_desc(p.desc, prev);
    if (d.degree == 0) { *init_out = kNoVertex; return; }
    auto weight_of = [&](u32 x) -> float {                                     // RandomWalkModel::weight - deepwalk.h:67-70, node2vec.h:74-88
        if (model == WMH_DEEPWALK) return 1.0f;
        if (x == prev) return p.inv_p;
        bool hit;
        row_find(p.pool + dprev.begin, dprev.degree, x, &hit, p.shift);
        return hit ? 1.0f : p.inv_q;
    };
    const u64 key_b = ((u64) prev << 32) | __ldg(p.version + cur);
    u64 i = 0;
    u32 slot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, key_b, i++), d.degree);                       // RANDOM :72-75
    if (p.init_strategy == WMH_INIT_BURNIN)                                                                    // :77-85
    {
        for (int it = 0; it < 100; it++)
        {
            const u32 cslot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, key_b, i++), d.degree);
            const float wn = weight_of(__ldg(nbr_ptr(p, d, cslot))), wo = weight_of(__ldg(nbr_ptr(p, d, slot)));
            bool acc = wo < wn;
            if (!acc) acc = unit_double(keyed_draw(p.seed, kDomSampler, cur, key_b, i++)) <= (double) wn / (double) wo;
            if (acc) slot = cslot;
        }
    }
    else if (p.init_strategy == WMH_INIT_WEIGHT)                                                               // :87-107
    {
        float best = weight_of(__ldg(nbr_ptr(p, d, slot)));
        for (int it = 0; it < 20; it++)
        {
            const u32 cslot = mod_u64_u32(keyed_draw(p.seed, kDomSampler, cur, key_b, i++), d.degree);
            const float w = weight_of(__ldg(nbr_ptr(p, d, cslot)));
            if (w > best) { best = w; slot = cslot; }
        }
    }
    *init_out = __ldg(nbr_ptr(p, d, slot));
    for (u64 k = 0; k < iters; k++)                                                                            // sample :31-46
    {
        const u32 cslot = mod_u64_u32(keyed_draw(p.seed, kDomChain, cur, prev, 2 * k), d.degree);
        const float wn = weight_of(__ldg(nbr_ptr(p, d, cslot))), wo = weight_of(__ldg(nbr_ptr(p, d, slot)));
        bool acc = wo < wn;
        if (!acc) acc = unit_double(keyed_draw(p.seed, kDomChain, cur, prev, 2 * k + 1)) <= (double) wn / (double) wo;
        if (acc) slot = cslot;
        counts[slot] += 1;
    }
}

// DeepWalk in MH mode: weight == 1 (deepwalk.h:67-70) so accept() is drand() <= 1.0, always true, and the walk
// is the proposal chain (deepwalk.h:79-87) with keyed draws.
template <bool kRewalk>
__global__ void __launch_bounds__(kWalkBlock) walk_deepwalk_mh_kernel(DevWalkParams p, const u32* __restrict__ affected,
                                                                      const u32* __restrict__ positions, u64 count,
                                                                      unsigned long long* steps_out)
{
    const u64 t = (u64) blockIdx.x * kWalkBlock + threadIdx.x;
    u32 steps = 0;
    if (t < count)
    {
        u64 w;
        u32 pos = 0, cur;
        if (kRewalk) { w = __ldg(affected + t); pos = __ldg(positions + t); } else { w = p.first + t; }
        u32 replica, start;
        p.map.locate(w, replica, start);
        const u64 walk_id = (u64) replica * p.n + start;
        if (kRewalk) cur = p.walks[(u64) pos * p.w_local + w];
        else { cur = start; p.walks[w] = cur; }
        u32* out = p.walks + (u64) (pos + 1) * p.w_local + w;
        const u32 first = pos;
        const u32 last_pos = p.L - 1;
        RowDesc d = load_desc(p.desc, cur);
        for (; pos < last_pos && d.degree; ++pos, out += p.w_local)
        {
            cur = __ldg(nbr_ptr(p, d, mod_u64_u32(keyed_draw(p.seed, kDomStep, walk_id, pos, 0), d.degree)));
            __stcs(out, cur);
            d = load_desc(p.desc, cur);
        }
        steps = pos - first + (d.degree ? 1u : 0u);
        for (++pos; pos <= last_pos; ++pos, out += p.w_local) __stcs(out, kNoVertex);
    }
    block_add_u64(steps_out, steps);
}

// ---- cooperative sampler init over peer memory (MH node2vec, world > 1) ----------------------------------
//
// The init value of the sampler of state (cur, prev) is a pure function of the replicated graph, so the box computes
// each one once: states are numbered in row order (state e = eoff[prev] + slot of cur in the row of prev), rank g
// owns the states of the vertex range [split[g], split[g + 1]) (equal edge counts), and its kernel stores the 16-bit
// cache words straight into the exchange buffer of every peer (32-byte vector stores over NVLink: 16 consecutive
// states) and into its own aux words.  After the barrier every rank copies the words it received into its own.

struct DegreeOf
{
    const RowDesc* desc;
    u64 n;
    __device__ u32 operator()(u32 v) const { return v < n ? desc[v].degree : 0u; }
};

// split[g] = first vertex whose state offset is >= g * m / world (g = 0..world), so hub rows are never cut.
__global__ void share_splits_kernel(const u32* __restrict__ eoff, u32 n, u32 world, u32* __restrict__ split)
{
    const u32 g = threadIdx.x;
    if (g > world) return;
    const u32 m = eoff[n];
    const u32 target = (u32) ((u64) m * g / world);
    u32 lo = 0, hi = n;                                  // lower_bound over eoff[0..n)
    while (lo < hi)
    {
        const u32 mid = (lo + hi) >> 1;
        if (eoff[mid] < target) lo = mid + 1; else hi = mid;
    }
    split[g] = g == world ? n : lo;
}

constexpr u32 kShareChunk = 64;      // states per warp round: four 8-lane groups x 16 consecutive states

__global__ void __launch_bounds__(kWalkBlock) n2v_share_init_kernel(DevWalkParams p, const u32* __restrict__ eoff, const u32* __restrict__ split,
                                                                     PeerPtrs peers, u32 world, u32 rank)
{
    const u32 lane = threadIdx.x & 31, sub = lane & 7, group = lane >> 3;
    const u32 vlo = split[rank], vhi = split[rank + 1];
    if (vlo >= vhi) return;
    const u64 elo = eoff[vlo], ehi = eoff[vhi];
    const u64 base = elo & ~(u64) (kShareChunk - 1);
    const u64 warps = ((u64) gridDim.x * kWalkBlock) >> 5;
    unsigned short* const remote = (sub < world && sub != rank) ? reinterpret_cast<unsigned short*>(peers.buf[sub] + kPeerFlagBytes) : nullptr;
    for (u64 chunk = ((u64) blockIdx.x * kWalkBlock + threadIdx.x) >> 5; base + chunk * kShareChunk < ehi; chunk += warps)
    {
        const u64 e0 = base + chunk * kShareChunk + group * 16;          // this group's 16 states
        // row of the group's first state: largest v in [vlo, vhi) with eoff[v] <= max(e0, elo)
        u32 v = vlo;
        {
            const u64 key = e0 < elo ? elo : e0;
            u32 lo = vlo, hi = vhi;                                      // upper_bound - 1
            while (lo < hi)
            {
                const u32 mid = (lo + hi) >> 1;
                if (__ldg(eoff + mid) <= key) lo = mid + 1; else hi = mid;
            }
            v = lo > vlo ? lo - 1 : vlo;
        }
        u32 next_off = v + 1 <= vhi ? __ldg(eoff + v + 1) : 0xFFFFFFFFu;
        RowDesc dprev = load_desc(p.desc, v);
        u32 row_off = __ldg(eoff + v);
        u32 acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll 1
        for (u32 r = 0; r < 16; r++)
        {
            const u64 e = e0 + r;
            const bool mine = e >= elo && e < ehi;
            u32 cur = 0, slot = 0;
            RowDesc d{0, 0};
            if (mine)
            {
                while (e >= next_off)                                    // next row (skips empty rows)
                {
                    ++v;
                    row_off = next_off;
                    next_off = __ldg(eoff + v + 1);
                    dprev = load_desc(p.desc, v);
                }
                slot = (u32) (e - row_off);
                cur = __ldg(nbr_ptr(p, dprev, slot));
                d = load_desc(p.desc, cur);
            }
            const bool has = mine && d.degree != 0;                      // a walker on a sink never samples
            InitValue iv{kClassFar, 0, 0};
            if (p.init_strategy == WMH_INIT_WEIGHT) iv = n2v_init_weight_group(p, sub, has, cur, d, v, dprev);
            else
            {
                if (has && sub == 0) iv = n2v_init_serial(p, cur, d, v, dprev);
                const int from = (int) (lane & ~7u);
                iv = InitValue{__shfl_sync(0xffffffffu, iv.cls, from), __shfl_sync(0xffffffffu, iv.slot, from), __shfl_sync(0xffffffffu, iv.draw, from)};
            }
            const u32 word = has ? cache_encode(iv, d.degree) : 0u;
            acc[r >> 1] |= word << ((r & 1) * 16);
            if (mine && sub == 0)
                __stcg(reinterpret_cast<unsigned short*>(aux_ptr(p, dprev, slot)), (unsigned short) word);   // own aux word, low half
        }
        if (remote)
        {
            if (e0 >= elo && e0 + 16 <= ehi)                             // whole window owned: one sector
            {
                uint4* dst = reinterpret_cast<uint4*>(remote + e0);
                dst[0] = make_uint4(acc[0], acc[1], acc[2], acc[3]);
                dst[1] = make_uint4(acc[4], acc[5], acc[6], acc[7]);
            }
            else
            {
                for (u32 r = 0; r < 16; r++)
                    if (e0 + r >= elo && e0 + r < ehi) remote[e0 + r] = (unsigned short) (acc[r >> 1] >> ((r & 1) * 16));
            }
        }
    }
}

// Eight lanes per row: tails of the rows other ranks computed <- the words they stored into the own exchange buffer.
// A barrier that gave up waiting (error word set) means the window may hold words of an earlier pass or only part of
// this one: the caches are then left EMPTY (0 = lazy first-touch init in the walk kernel) instead of imported.  After a
// barrier that succeeded every word read here is fresh: each exchange overwrites every state of the current graph in
// every window (state e is stored by the rank that owns it), so nothing of an earlier pass can survive into an import.
// (Zeroing each word after reading it, as a second line of defence, cost 0.5 ms at two ranks and 0.9 ms at eight: the
// owner's 2-byte stores into its peer-mapped window are slow; measured with tools/r02_import_probe.sh.)  The group takes
// the first kImportHead elements of its row; longer rows are listed and finished by a block each.
constexpr u32 kImportHead = 256;

__device__ __forceinline__ void import_words(const unsigned short* __restrict__ src, unsigned short* __restrict__ halves, u32 first, u32 last, u32 lane,
                                             u32 lanes, bool failed)
{
    for (u32 i = first + lane; i < last; i += lanes)
        halves[4 * (u64) i + 2] = failed ? (unsigned short) 0 : src[i];                        // low half of aux word i = halves[4 i + 2]
}

__global__ void __launch_bounds__(256) n2v_share_import_kernel(const RowDesc* __restrict__ desc, u32* __restrict__ pool, const u32* __restrict__ eoff,
                                                               const u32* __restrict__ split, const unsigned short* __restrict__ received, u32 n, u32 rank,
                                                               const u32* __restrict__ error_word, u32* __restrict__ long_rows, u32* __restrict__ n_long)
{
    const u64 t = (u64) blockIdx.x * 256 + threadIdx.x;
    const u32 v = (u32) (t >> 3), sub = threadIdx.x & 7;
    if (v >= n || (v >= split[rank] && v < split[rank + 1])) return;
    const bool failed = *reinterpret_cast<const volatile u32*>(error_word) != 0;
    const RowDesc d = load_desc(desc, v);
    import_words(received + __ldg(eoff + v), reinterpret_cast<unsigned short*>(pool + d.begin), 0, min(d.degree, kImportHead), sub, 8, failed);
    if (sub == 0 && d.degree > kImportHead) long_rows[atomicAdd(n_long, 1u)] = v;
}

__global__ void __launch_bounds__(256) n2v_share_import_long_kernel(const RowDesc* __restrict__ desc, u32* __restrict__ pool, const u32* __restrict__ eoff,
                                                                    const unsigned short* __restrict__ received, const u32* __restrict__ error_word,
                                                                    const u32* __restrict__ long_rows, const u32* __restrict__ n_long)
{
    const bool failed = *reinterpret_cast<const volatile u32*>(error_word) != 0;
    const u32 count = *n_long;
    for (u32 j = blockIdx.x; j < count; j += gridDim.x)
    {
        const u32 v = long_rows[j];
        const RowDesc d = load_desc(desc, v);
        import_words(received + __ldg(eoff + v), reinterpret_cast<unsigned short*>(pool + d.begin), kImportHead, d.degree, threadIdx.x, 256, failed);
    }
}

// Flag barrier in peer memory: lane g tells rank g "rank `rank` reached epoch", then waits for rank g's word in the
// own flag page.  Kernel boundaries order the data: the stores of every kernel enqueued before this one are performed
// before the release store, and kernels enqueued after it start after the acquire loads saw every peer.
__global__ void peer_barrier_kernel(PeerPtrs peers, u32 world, u32 rank, u64 epoch, u64 timeout_ns)
{
    const u32 g = threadIdx.x;
    if (g >= world) return;
    __threadfence_system();
    unsigned long long* there = reinterpret_cast<unsigned long long*>(peers.buf[g]) + rank;
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(there), "l"(epoch) : "memory");
    const unsigned long long* here = reinterpret_cast<const unsigned long long*>(peers.buf[rank]) + g;
    unsigned long long t0, now, seen;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;)
    {
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(here) : "memory");
        if (seen >= epoch) break;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (now - t0 > timeout_ns)
        {
            *reinterpret_cast<u32*>(peers.buf[rank] + kPeerErrorOffset) = 1u;
            break;
        }
        __nanosleep(200);
    }
    __threadfence_system();
}

// ---- corpus scan: earliest position standing on a touched vertex (MapOfChanges, wharfmh.h:519-537) ----
__global__ void __launch_bounds__(256) scan_touched_kernel(const u32* __restrict__ walks, u64 w_local, u32 L,
                                                            const u32* __restrict__ bitmap, u32* __restrict__ min_pos)
{
    const u64 w = (u64) blockIdx.x * 256 + threadIdx.x;
    if (w >= w_local) return;
    const u32* col = walks + w;
    u32 found = kNoPos;
    // positions in chunks of 8 independent coalesced loads, then test in order
    for (u32 base = 0; base < L && found == kNoPos; base += 8)
    {
        u32 v[8];
#pragma unroll
        for (int i = 0; i < 8; i++) v[i] = (base + i < L) ? __ldcs(col + (u64) (base + i) * w_local) : kNoVertex;
#pragma unroll
        for (int i = 0; i < 8; i++)
        {
            if (found != kNoPos) break;
            if (v[i] == kNoVertex) { base = L; break; }                  // walker stopped: nothing further
            if ((__ldg(bitmap + (v[i] >> 5)) >> (v[i] & 31)) & 1u) found = base + i;
        }
    }
    min_pos[w] = found;
}

struct IsAffected
{
    const u32* min_pos;
    __device__ bool operator()(u32 w) const { return min_pos[w] != kNoPos; }
};

__global__ void gather_positions_kernel(const u32* __restrict__ min_pos, const u32* __restrict__ ids, u64 count, u32* __restrict__ positions)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) positions[i] = min_pos[ids[i]];
}

// ---- retrieval -----------------------------------------------------------------------------------------
// Tile transpose [L][W] -> rows [count][L] for local walks first..first+count.
__global__ void __launch_bounds__(256) transpose_rows_kernel(const u32* __restrict__ walks, u64 w_local, u32 L, u64 first, u64 count,
                                                              u32* __restrict__ out)
{
    __shared__ u32 tile[32][33];
    const u64 w0 = (u64) blockIdx.x * 32;
    const u32 p0 = blockIdx.y * 32;
    const u32 tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    for (u32 r = ty; r < 32; r += 8)
    {
        const u32 pos = p0 + r;
        const u64 w = w0 + tx;
        tile[r][tx] = (pos < L && w < count) ? walks[(u64) pos * w_local + first + w] : kNoVertex;
    }
    __syncthreads();
    for (u32 r = ty; r < 32; r += 8)
    {
        const u64 w = w0 + r;
        const u32 pos = p0 + tx;
        if (w < count && pos < L) out[w * L + pos] = tile[tx][r];
    }
}

__global__ void gather_rows_kernel(const u32* __restrict__ walks, u64 w_local, u32 L, const u32* __restrict__ ids, u64 count,
                                   u32* __restrict__ out)
{
    // one warp per walk, lanes over positions
    const u64 warp = ((u64) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 lane = threadIdx.x & 31;
    if (warp >= count) return;
    const u64 w = ids[warp];
    for (u32 pos = lane; pos < L; pos += 32) out[warp * L + pos] = walks[(u64) pos * w_local + w];
}

// Sum over walks of the FNV-1a-64 of each walk's text form "v0 v1 ... " (graph/wharfmh.h:365-394).
__global__ void __launch_bounds__(256) fingerprint_kernel(const u32* __restrict__ walks, u64 w_local, u32 L, unsigned long long* sum)
{
    const u64 w = (u64) blockIdx.x * 256 + threadIdx.x;
    u64 h = 0;
    if (w < w_local)
    {
        h = 0xcbf29ce484222325ull;
        for (u32 pos = 0; pos < L; pos++)
        {
            u32 v = walks[(u64) pos * w_local + w];
            if (v == kNoVertex) break;
            char buf[10];
            int len = 0;
            do { buf[len++] = (char) ('0' + v % 10); v /= 10; } while (v);
            while (len) { h ^= (u8) buf[--len]; h *= 0x100000001b3ull; }
            h ^= (u8) ' ';
            h *= 0x100000001b3ull;
        }
    }
    // wrapping 64-bit sum: warp reduce then one atomic per warp
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) h += __shfl_xor_sync(0xffffffffu, h, o);
    if ((threadIdx.x & 31) == 0 && h) atomicAdd(sum, (unsigned long long) h);
}

// Global walk id (replica * n + start vertex) of local walks: the ids insert/delete_edges_batch return (wharfmh.h:548,698).
__global__ void global_ids_kernel(ShardMap map, const u32* __restrict__ local_ids, u64 count, u64* __restrict__ out)
{
    const u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) out[i] = map.global_id(local_ids[i]);
}

__global__ void fill_u32_kernel(u32* __restrict__ d, u64 count, u32 value)
{
    for (u64 i = (u64) blockIdx.x * blockDim.x + threadIdx.x; i < count; i += (u64) gridDim.x * blockDim.x) d[i] = value;
}

DevWalkParams to_dev(const WalkParams& wp)
{
    DevWalkParams p;
    p.desc = wp.desc; p.pool = wp.pool; p.version = wp.version; p.pool_rw = wp.pool_rw; p.draw = wp.draw; p.walks = wp.walks;
    p.n = wp.n; p.w_local = wp.w_local; p.map = wp.map; p.L = wp.L;
    p.init_strategy = wp.init_strategy;
    p.shift = wp.shift;
    p.tri = wp.tri;
    {
        const char* coop = std::getenv("WHARF_COOP_SEARCH");
        p.coop_search = coop && std::atoi(coop) != 0;
        // descriptor of the proposal before the accept decision (hides its latency; wasted for the 70 % that are rejected) while
        // the descriptor array sits in the L2 anyway, after it beyond that: scale-20 (8 MB) 23.07 vs 23.46 ms, LiveJournal shape
        // (39 MB) 111.6 vs 112.2 ms, scale-23 (64 MB) 245.1 vs 238.9 ms - profiles/r02k_spec_desc_ab.txt.  WHARF_SPEC_DESC=0|1 forces.
        const char* spec = std::getenv("WHARF_SPEC_DESC");
        p.spec_desc = spec ? (std::atoi(spec) != 0) : (wp.n * sizeof(RowDesc) <= (48ull << 20));
    }
    p.inv_p = 1 / wp.p;                                                 // node2vec.h:78
    p.inv_q = 1 / wp.q;                                                 // node2vec.h:86
    p.seed = wp.seed;
    p.first = 0;
    return p;
}

// Tuning knobs of the bulk deterministic kernel: walkers per thread / block size / streaming stores.
// Defaults are the measured best on B200 (profiles/); WHARF_DET_VARIANT="walks,block,stream" overrides for sweeps.
struct DetVariant { int walks = 1, block = 256, stream = 0; };

DetVariant det_variant()
{
    DetVariant v;
    if (const char* e = std::getenv("WHARF_DET_VARIANT")) std::sscanf(e, "%d,%d,%d", &v.walks, &v.block, &v.stream);
    return v;
}

template <int kWalks, int kBlock, bool kStream>
void launch_det_ilp(const LaunchCtx& lc, const DevWalkParams& p, unsigned long long* d_steps)
{
    const u64 threads = (p.w_local + kWalks - 1) / kWalks;
    const unsigned grid = div_up(threads, kBlock);
    const u64 stride = (u64) grid * kBlock;
    walk_deterministic_ilp_kernel<kWalks, kBlock, kStream><<<grid, kBlock, 0, lc.stream>>>(p, stride, d_steps);
}

void launch_generate_deterministic(const LaunchCtx& lc, const DevWalkParams& p, unsigned long long* d_steps)
{
    const DetVariant v = det_variant();
#define WHARF_DET_CASE(W, B, S) if (v.walks == W && v.block == B && v.stream == S) { launch_det_ilp<W, B, S != 0>(lc, p, d_steps); lc.count(); WHARF_CUDA(cudaGetLastError()); return; }
    WHARF_DET_CASE(1, 256, 0) WHARF_DET_CASE(1, 256, 1) WHARF_DET_CASE(1, 128, 0) WHARF_DET_CASE(1, 512, 0) WHARF_DET_CASE(1, 1024, 0)
    WHARF_DET_CASE(2, 256, 0) WHARF_DET_CASE(2, 256, 1) WHARF_DET_CASE(2, 128, 1) WHARF_DET_CASE(2, 512, 1)
    WHARF_DET_CASE(4, 256, 0) WHARF_DET_CASE(4, 256, 1) WHARF_DET_CASE(4, 128, 1) WHARF_DET_CASE(4, 512, 1)
    WHARF_DET_CASE(8, 256, 1) WHARF_DET_CASE(8, 128, 1)
#undef WHARF_DET_CASE
    throw CudaError(cudaErrorInvalidValue, "unsupported WHARF_DET_VARIANT");
}

template <bool kRewalk>
void launch_walk(const LaunchCtx& lc, const WalkParams& wp, const u32* ids, const u32* positions, u64 first, u64 count, unsigned long long* d_steps)
{
    if (count == 0) return;
    DevWalkParams p = to_dev(wp);
    p.first = first;
    const unsigned grid = div_up(count, kWalkBlock);
    if (wp.deterministic && !kRewalk && first == 0 && count == wp.w_local && det_variant().walks != 0)
    {
        launch_generate_deterministic(lc, p, d_steps);
        return;
    }
    if (wp.deterministic)
        walk_deterministic_kernel<kRewalk><<<grid, kWalkBlock, 0, lc.stream>>>(p, ids, positions, count, d_steps);
    else if (wp.model == WMH_NODE2VEC)
        walk_n2v_kernel<kRewalk><<<grid, kWalkBlock, 0, lc.stream>>>(p, ids, positions, count, d_steps);
    else
        walk_deepwalk_mh_kernel<kRewalk><<<grid, kWalkBlock, 0, lc.stream>>>(p, ids, positions, count, d_steps);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}
}  // namespace

void walks_generate(const LaunchCtx& lc, const WalkParams& wp, unsigned long long* d_steps)
{
    launch_walk<false>(lc, wp, nullptr, nullptr, 0, wp.w_local, d_steps);
}

void walks_generate_range(const LaunchCtx& lc, const WalkParams& wp, u64 first, u64 count, unsigned long long* d_steps)
{
    launch_walk<false>(lc, wp, nullptr, nullptr, first, count, d_steps);
}

void walks_rewalk(const LaunchCtx& lc, const WalkParams& wp, const u32* d_affected, const u32* d_positions, u64 count,
                  unsigned long long* d_steps)
{
    launch_walk<true>(lc, wp, d_affected, d_positions, 0, count, d_steps);
}

void walks_scan_touched(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, const u32* d_bitmap, u32* d_min_pos)
{
    if (w_local == 0) return;
    scan_touched_kernel<<<div_up(w_local, 256), 256, 0, lc.stream>>>(d_walks, w_local, L, d_bitmap, d_min_pos);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

// ---- the same scan with its scratch in static device memory (wmh_scan_touched_no_alloc) ----------------
// For callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc): the touched-vertex
// bitmap the scan needs lives in a file-scope device array instead of a stream-ordered allocation.  Every GPU that loads
// the library holds these 4 MiB until the process exits.  All calls on a device share the array: a call clears the words it
// needs before marking, so it never sees an earlier call's bits, but two calls that overlap on the GPU race.
static __device__ __align__(256) u32 g_static_touched_bitmap[kNoAllocMaxVertices / 32];

__global__ void __launch_bounds__(256) mark_touched_static_kernel(const u32* __restrict__ vertices, u64 count, u64 n, u32* __restrict__ bitmap)
{
    const u64 i = (u64) blockIdx.x * 256 + threadIdx.x;
    if (i >= count) return;
    const u32 v = vertices[i];
    if (v < n) atomicOr(bitmap + (v >> 5), 1u << (v & 31));      // ids beyond the graph name no walker's vertex: ignored
}

void walks_scan_touched_no_alloc(const LaunchCtx& lc, int device, const u32* d_walks, u64 w_local, u32 L, u64 n, const u32* d_vertices, u64 count,
                                 u32* d_min_pos)
{
    if (n > kNoAllocMaxVertices) throw std::length_error("graph has more vertices than the static scratch of the no-alloc scan holds");
    // In host code the array's name is a host-side stand-in, not a device address: the address the kernels get is looked up
    // once per device, with that device current (the caller bound it), and kept.
    static std::mutex lock;
    static u32* address[64] = {};
    if (device < 0 || device >= 64) throw std::length_error("device index beyond the no-alloc scan's table");
    u32* bitmap = nullptr;
    {
        std::lock_guard<std::mutex> guard(lock);
        if (!address[device])
        {
            void* p = nullptr;
            WHARF_CUDA(cudaGetSymbolAddress(&p, g_static_touched_bitmap));
            address[device] = static_cast<u32*>(p);
        }
        bitmap = address[device];
    }
    if (w_local == 0) return;
    WHARF_CUDA(cudaMemsetAsync(bitmap, 0, ((n + 31) / 32) * sizeof(u32), lc.stream));     // the last call's bits
    if (count)
    {
        mark_touched_static_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(d_vertices, count, n, bitmap);
        lc.count();
    }
    scan_touched_kernel<<<div_up(w_local, 256), 256, 0, lc.stream>>>(d_walks, w_local, L, bitmap, d_min_pos);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

u64 walks_compact_affected(const LaunchCtx& lc, const u32* d_min_pos, u64 w_local, u32* d_ids, u32* d_positions)
{
    if (w_local == 0) return 0;
    Scratch<u64> d_count(lc.stream, 1);
    cub::CountingInputIterator<u32> iota(0);
    IsAffected pred{d_min_pos};
    size_t temp_bytes = 0;
    WHARF_CUDA(cub::DeviceSelect::If(nullptr, temp_bytes, iota, d_ids, d_count.ptr, (i64) w_local, pred, lc.stream));
    Scratch<u8> temp(lc.stream, temp_bytes);
    WHARF_CUDA(cub::DeviceSelect::If(temp.ptr, temp_bytes, iota, d_ids, d_count.ptr, (i64) w_local, pred, lc.stream));
    lc.count_lib();
    u64 count = 0;
    WHARF_CUDA(cudaMemcpyAsync(&count, d_count.ptr, sizeof(u64), cudaMemcpyDeviceToHost, lc.stream));
    WHARF_CUDA(cudaStreamSynchronize(lc.stream));
    if (count)
    {
        gather_positions_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(d_min_pos, d_ids, count, d_positions);
        lc.count();
        WHARF_CUDA(cudaGetLastError());
    }
    return count;
}

void walks_transpose_rows(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, u64 first, u64 count, u32* d_out)
{
    if (count == 0) return;
    dim3 grid(div_up(count, 32), div_up(L, 32));
    transpose_rows_kernel<<<grid, 256, 0, lc.stream>>>(d_walks, w_local, L, first, count, d_out);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void walks_gather_rows(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, const u32* d_ids, u64 count, u32* d_out)
{
    if (count == 0) return;
    gather_rows_kernel<<<div_up(count * 32, 256), 256, 0, lc.stream>>>(d_walks, w_local, L, d_ids, count, d_out);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void walks_fingerprint(const LaunchCtx& lc, const u32* d_walks, u64 w_local, u32 L, unsigned long long* d_sum)
{
    if (w_local == 0) return;
    fingerprint_kernel<<<div_up(w_local, 256), 256, 0, lc.stream>>>(d_walks, w_local, L, d_sum);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void sampler_chain(const LaunchCtx& lc, const WalkParams& wp, u32 cur, u32 prev, u64 iters, u32* d_init, unsigned long long* d_counts)
{
    sampler_chain_kernel<<<1, 32, 0, lc.stream>>>(to_dev(wp), wp.model, cur, prev, iters, d_init, d_counts);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void walks_global_ids(const LaunchCtx& lc, const ShardMap& map, const u32* d_local_ids, u64 count, u64* d_out)
{
    if (count == 0) return;
    global_ids_kernel<<<div_up(count, 256), 256, 0, lc.stream>>>(map, d_local_ids, count, d_out);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void fill_u32(const LaunchCtx& lc, u32* d, u64 count, u32 value)
{
    if (count == 0) return;
    fill_u32_kernel<<<grid_for(count, 256, 8, lc.sm_count), 256, 0, lc.stream>>>(d, count, value);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void peer_barrier(const LaunchCtx& lc, PeerLink& link)
{
    ++link.epoch;
    peer_barrier_kernel<<<1, 32, 0, lc.stream>>>(link.ptrs, link.world, link.rank, link.epoch, (u64) WMH_PEER_TIMEOUT_MS * 1000000ull);
    lc.count();
    WHARF_CUDA(cudaGetLastError());
}

void n2v_share_sampler_init(const LaunchCtx& lc, const WalkParams& wp, u64 m, PeerLink& link)
{
    const u32 n = (u32) wp.n;
    Trace tr(lc.stream);
    Scratch<u32> eoff(lc.stream, (u64) n + 1), split(lc.stream, kPeerMaxWorld + 1);
    {
        cub::CountingInputIterator<u32> iota(0);
        cub::TransformInputIterator<u32, DegreeOf, cub::CountingInputIterator<u32>> degrees(iota, DegreeOf{wp.desc, wp.n});
        size_t temp_bytes = 0;
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, temp_bytes, degrees, eoff.ptr, (int) (n + 1), lc.stream));
        Scratch<u8> temp(lc.stream, temp_bytes);
        WHARF_CUDA(cub::DeviceScan::ExclusiveSum(temp.ptr, temp_bytes, degrees, eoff.ptr, (int) (n + 1), lc.stream));
        lc.count_lib();
    }
    share_splits_kernel<<<1, 32, 0, lc.stream>>>(eoff.ptr, n, link.world, split.ptr);
    const DevWalkParams p = to_dev(wp);
    const u64 chunks = (m / link.world + kShareChunk) / kShareChunk + 2;
    const unsigned grid = (unsigned) std::min<u64>(div_up(chunks * 32, kWalkBlock), (u64) lc.sm_count * 8);
    n2v_share_init_kernel<<<grid, kWalkBlock, 0, lc.stream>>>(p, eoff.ptr, split.ptr, link.ptrs, link.world, link.rank);
    lc.count(2);
    tr.mark("share: scan + splits + init kernel");
    peer_barrier(lc, link);                                    // every rank's words have arrived
    tr.mark("share: barrier (words arrived)");
    {
        Scratch<u32> long_rows(lc.stream, m / kImportHead + 2);                 // rows longer than kImportHead: at most m / kImportHead of them
        WHARF_CUDA(cudaMemsetAsync(long_rows.ptr, 0, 4, lc.stream));            // [0] = their count
        const unsigned short* received = reinterpret_cast<const unsigned short*>(link.own() + kPeerFlagBytes);
        const u32* error_word = reinterpret_cast<const u32*>(link.own() + kPeerErrorOffset);
        n2v_share_import_kernel<<<div_up((u64) n * 8, 256), 256, 0, lc.stream>>>(wp.desc, wp.pool_rw, eoff.ptr, split.ptr, received, n, link.rank, error_word,
                                                                                 long_rows.ptr + 1, long_rows.ptr);
        n2v_share_import_long_kernel<<<lc.sm_count * 8, 256, 0, lc.stream>>>(wp.desc, wp.pool_rw, eoff.ptr, received, error_word, long_rows.ptr + 1, long_rows.ptr);
        lc.count(2);
    }
    tr.mark("share: import");
    peer_barrier(lc, link);                                    // nobody overwrites a buffer that is still being read
    tr.mark("share: barrier (windows free)");
    WHARF_CUDA(cudaGetLastError());
    ++link.exchanges;
}

}  // namespace wharf
