/*
 * TEST INFRASTRUCTURE ONLY.  CPU restatement ("oracle") of the WharfMH hot path in plain C.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this
 * library, and only as the checker.  The product path (dynamicgraphrepresentationlearning_b200/) never
 * links, imports or executes it.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks this file against
 *   - the golden vectors the survey took from the unmodified reference (Random KATs, the
 *     6-vertex corpus before/after an insert and a delete, the cora corpus hashes), and
 *   - fixtures produced here by oracle/_ref/wharf_ref (the unmodified reference, oracle/ref_driver.cpp)
 *     with tests/golden/make_golden.py (RMAT graphs, insert/delete sequences, batch generator output).
 * The keyed Metropolis-Hastings mode has NO bit-level counterpart in the reference (its MH mode is not
 * reproducible): that mode is "parity unpinned" at the bit level and is checked
 * distributionally against oracle/_ref (tests/test_mh_distribution.py).
 *
 * Every function cites the reference file:line (relative to /root/reference) it restates.
 *
 * Deliberate, documented deviations (the reference's behaviour there is a crash or corruption):
 *   - 64-bit index keys (reference: u32 walk_id*L+pos wraps for n > 5.37 M vertices, inverted_index.h:14,76).
 *   - A walker standing on a degree-0 vertex stops: the remaining positions hold WO_NO_VERTEX
 *     (reference: SIGFPE via lrand() % 0, utils/utility.h:220; or a dangling sentinel entry,
 *     graph/wharfmh.h:800-811).
 *   - The minimum affected position per walk is the true minimum (reference: racy check-then-act,
 *     graph/wharfmh.h:524-536; equal to the true minimum when run_seq=true).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define WO_NO_VERTEX 0xFFFFFFFFu
#define WO_API __attribute__((visibility("default")))

typedef uint64_t u64;
typedef uint32_t u32;
typedef int64_t  i64;

/* ------------------------------------------------------------------------------------------------
 * Minimal fork-join helper (stands in for pbbs parallel_for, libs/compressed_trees/pbbslib/parallel.h).
 * Thread count: env WO_THREADS, default = online cores.  Chunks are claimed with an atomic counter.
 * ---------------------------------------------------------------------------------------------- */
#include <pthread.h>
#include <unistd.h>

typedef void (*wo_range_fn)(u64 lo, u64 hi, void* ctx, u64* acc);
typedef struct { wo_range_fn fn; void* ctx; u64 n, chunk; u64 next; u64 acc; pthread_mutex_t mu; } wo_pf;

WO_API int wo_threads(void)
{
    const char* s = getenv("WO_THREADS");
    long t = s ? atol(s) : sysconf(_SC_NPROCESSORS_ONLN);
    return t < 1 ? 1 : (t > 1024 ? 1024 : (int) t);
}

static void* wo_pf_worker(void* arg)
{
    wo_pf* p = (wo_pf*) arg;
    u64 local = 0;
    for (;;)
    {
        u64 lo = __atomic_fetch_add(&p->next, p->chunk, __ATOMIC_RELAXED);
        if (lo >= p->n) break;
        u64 hi = lo + p->chunk < p->n ? lo + p->chunk : p->n;
        p->fn(lo, hi, p->ctx, &local);
    }
    pthread_mutex_lock(&p->mu);
    p->acc += local;
    pthread_mutex_unlock(&p->mu);
    return NULL;
}

/* Runs fn over [0,n) in chunks; returns the sum of the per-thread accumulators. */
static u64 wo_parallel_for(u64 n, u64 chunk, wo_range_fn fn, void* ctx)
{
    wo_pf p = { fn, ctx, n, chunk ? chunk : 1, 0, 0, PTHREAD_MUTEX_INITIALIZER };
    int t = wo_threads();
    if (t == 1 || n <= p.chunk) { u64 acc = 0; if (n) fn(0, n, ctx, &acc); return acc; }
    pthread_t th[1024];
    int started = 0;
    for (int i = 0; i < t - 1; i++) if (pthread_create(&th[started], NULL, wo_pf_worker, &p) == 0) started++;
    wo_pf_worker(&p);
    for (int i = 0; i < started; i++) pthread_join(th[i], NULL);
    return p.acc;
}

/* ------------------------------------------------------------------------------------------------
 * utility::Random  (utils/utility.h:152-223)
 * ---------------------------------------------------------------------------------------------- */

/* Seeding: two rounds of splitmix64 in which z is a SIGNED long long, so every `>>` is arithmetic
 * (utils/utility.h:157-170). */
WO_API void wo_random_seed(u64 seed, u64 s[2])
{
    for (int i = 0; i < 2; i++)
    {
        seed += UINT64_C(0x9E3779B97F4A7C15);
        i64 z = (i64) seed;
        z = (i64) ((u64) (z ^ (z >> 30)) * UINT64_C(0xBF58476D1CE4E5B9));
        z = (i64) ((u64) (z ^ (z >> 27)) * UINT64_C(0x94D049BB133111EB));
        s[i] = (u64) (z ^ (z >> 31));
    }
}

static inline u64 rotl64(u64 x, int k) { return (x << k) | (x >> (64 - k)); }

/* xoroshiro128+ with constants (55, 14, 36)  (utils/utility.h:194-206). */
WO_API u64 wo_lrand(u64 s[2])
{
    const u64 s0 = s[0];
    u64 s1 = s[1];
    const u64 result = s0 + s1;
    s1 ^= s0;
    s[0] = rotl64(s0, 55) ^ s1 ^ (s1 << 14);
    s[1] = rotl64(s1, 36);
    return result;
}

/* (utils/utility.h:208-218) */
static inline double bits_to_unit_double(u64 x)
{
    union { u64 i; double d; } a;
    a.i = (UINT64_C(0x3FF) << 52) | (x >> 12);
    return a.d - 1.0;
}

WO_API double wo_drand(u64 s[2]) { return bits_to_unit_double(wo_lrand(s)); }

/* irand(int max) = lrand() % max: the int is converted to u64 for the modulo (utils/utility.h:220). */
static inline u64 irand_from(u64 x, u64 degree) { return x % (u64) (i64) (int) degree; }

/* X[r][k] = k-th lrand() of Random(r): the only randomness of deterministic mode
 * (graph/wharfmh.h:289-291,307 and 814-816,838). */
WO_API void wo_draw_table(u32 replicas, u32 length, u64* out)
{
    for (u32 r = 0; r < replicas; r++)
    {
        u64 s[2];
        wo_random_seed(r, s);
        for (u32 k = 0; k < length; k++) out[(u64) r * length + k] = wo_lrand(s);
    }
}

/* ------------------------------------------------------------------------------------------------
 * pairings::Szudzik<T>  (walks/pairings.h:27-128)
 * ---------------------------------------------------------------------------------------------- */

WO_API u32 wo_pair32(u32 x, u32 y) { return y >= x ? y * (y + 1) + x : x * x + y; }   /* pairings.h:27-86 */
WO_API u64 wo_pair64(u64 x, u64 y) { return y >= x ? y * (y + 1) + x : x * x + y; }

/* unpair: floor(sqrt((double) z)), one fix-down, then the two cases (pairings.h:100-114). */
WO_API void wo_unpair32(u32 z, u32* x, u32* y)
{
    u32 s = (u32) floor(sqrt((double) z));
    if (s * s > z) s--;
    u32 t = z - s * s;
    if (t < s) { *x = s; *y = t; } else { *x = t - s; *y = s; }
}

WO_API void wo_unpair64(u64 z, u64* x, u64* y)
{
    u64 s = (u64) floor(sqrt((double) z));
    if (s * s > z) s--;
    u64 t = z - s * s;
    if (t < s) { *x = s; *y = t; } else { *x = t - s; *y = s; }
}

WO_API u32 wo_pair_triplet32(u32 a, u32 b, u32 c) { return wo_pair32(wo_pair32(a, b), c); }  /* pairings.h:88-91 */
WO_API u64 wo_pair_triplet64(u64 a, u64 b, u64 c) { return wo_pair64(wo_pair64(a, b), c); }

WO_API void wo_unpair_triplet32(u32 z, u32* a, u32* b, u32* c)                              /* pairings.h:122-128 */
{
    u32 f, s;
    wo_unpair32(z, &f, &s);
    wo_unpair32(f, a, b);
    *c = s;
}

WO_API void wo_unpair_triplet64(u64 z, u64* a, u64* b, u64* c)
{
    u64 f, s;
    wo_unpair64(z, &f, &s);
    wo_unpair64(f, a, b);
    *c = s;
}

/* CompressedWalks::find_next (walks/compressed_walks.h:52-86): linear scan over the codes stored for one
 * vertex; each code = pair(walk_id*L+pos, next).  Returns WO_NO_VERTEX when absent. */
WO_API u32 wo_compressed_find_next(const u64* codes, u64 count, u64 walk_id, u32 position, u32 walk_length)
{
    for (u64 i = 0; i < count; i++)
    {
        u64 key, next;
        wo_unpair64(codes[i], &key, &next);
        u64 w = key / walk_length;
        u64 p = key - w * walk_length;
        if (w == walk_id && p == position) return (u32) next;
    }
    return WO_NO_VERTEX;
}

/* ------------------------------------------------------------------------------------------------
 * RMAT batch generator  (utils/utility.h:55-146, libs/compressed_trees/rmat_util.h:7-43,
 * libs/compressed_trees/pbbslib/utilities.h:108-147, pbbslib/random.h:38-51)
 * ---------------------------------------------------------------------------------------------- */

WO_API u32 wo_hash32(u32 a)                         /* pbbslib/utilities.h:108-116 */
{
    a = (a + 0x7ed55d16) + (a << 12);
    a = (a ^ 0xc761c23c) ^ (a >> 19);
    a = (a + 0x165667b1) + (a << 5);
    a = (a + 0xd3a2646c) ^ (a << 9);
    a = (a + 0xfd7046c5) + (a << 3);
    a = (a ^ 0xb55a4f09) ^ (a >> 16);
    return a;
}

WO_API u64 wo_hash64(u64 u)                         /* pbbslib/utilities.h:136-147 */
{
    u64 v = u * UINT64_C(3935559000370003845) + UINT64_C(2691343689449507681);
    v ^= v >> 21;
    v ^= v << 37;
    v ^= v >> 4;
    v *= UINT64_C(4768777513237032717);
    v ^= v << 20;
    v ^= v >> 41;
    v ^= v << 5;
    return v;
}

static u64 log2_up_u64(u64 i)                       /* pbbslib/utilities.h:286-291 */
{
    u64 a = 0, b = i - 1;
    while (b > 0) { b >>= 1; a++; }
    return a;
}

/* rMat<unsigned int>::operator()(i): the recursion of rmat_util.h:26-36 unrolled from the deepest level
 * (which consumes randStart + (levels-1)*stride... the recursion evaluates hashDouble(randStart) at the
 * OUTERMOST level, adding nn/2, and hashDouble(randStart + j*stride) at depth j). */
static void rmat_edge(u32 n_pow2, u32 h, double a, double ab, double abc, u32 i, u32* src, u32* dst)
{
    u32 rand_start  = wo_hash32((u32) (2 * i) * h);
    u32 rand_stride = wo_hash32((u32) (2 * i + 1) * h);
    u32 x = 0, y = 0;
    u32 r = rand_start;
    for (u32 nn = n_pow2; nn > 1; nn /= 2)
    {
        double d = (double) wo_hash32(r) / (double) 0xFFFFFFFFu;   /* hashDouble, rmat_util.h:21-24 */
        if (d < a) { }
        else if (d < ab)  { y += nn / 2; }
        else if (d < abc) { x += nn / 2; }
        else              { x += nn / 2; y += nn / 2; }
        r += rand_stride;
    }
    *src = x;
    *dst = y;
}

typedef struct { u64* keys; u64 edges_number; u32 n_pow2, h; double a, ab, abc; int directed; } rmat_ctx;
static void rmat_range(u64 lo, u64 hi, void* ctx, u64* acc)
{
    rmat_ctx* c = (rmat_ctx*) ctx; (void) acc;
    for (u64 i = lo; i < hi; i++)
    {
        u32 s, d;
        rmat_edge(c->n_pow2, c->h, c->a, c->ab, c->abc, (u32) i, &s, &d);
        c->keys[i] = ((u64) s << 32) | d;
        if (!c->directed) c->keys[i + c->edges_number] = ((u64) d << 32) | s;   /* utility.h:91-97 */
    }
}

static int cmp_u64(const void* p, const void* q)
{
    u64 a = *(const u64*) p, b = *(const u64*) q;
    return a < b ? -1 : a > b;
}

/* generate_batch_of_edges (utils/utility.h:55-146).  out_pairs must hold 2*edges_number*(directed?1:2) u32.
 * Returns the number of edges kept (sorted by (src,dst), duplicates and - unless self_loops - loops dropped). */
WO_API u64 wo_generate_batch(u64 edges_number, u64 vertices_number, u64 batch_seed, int self_loops, int directed,
                             double a, double b, double c, u32* out_pairs)
{
    u64 n_pow2 = UINT64_C(1) << (log2_up_u64(vertices_number) - 1);          /* utility.h:78 */
    u32 rmat_seed = (u32) wo_hash64(0 + batch_seed);                         /* rand.ith_rand(0) -> intT (u32) */
    u32 h = wo_hash32(rmat_seed);                                            /* rmat_util.h:15 */
    u64 total = directed ? edges_number : 2 * edges_number;
    u64* keys = (u64*) malloc((total ? total : 1) * sizeof(u64));
    rmat_ctx rc = { keys, edges_number, (u32) n_pow2, h, a, a + b, a + b + c, directed };
    wo_parallel_for(edges_number, 1 << 16, rmat_range, &rc);
    qsort(keys, total, sizeof(u64), cmp_u64);                                /* lexicographic (src,dst) */
    u64 kept = 0;
    for (u64 i = 0; i < total; i++)                                          /* utility.h:121-129 */
    {
        u32 s = (u32) (keys[i] >> 32), d = (u32) keys[i];
        if (!self_loops && s == d) continue;
        if (i > 0 && keys[i] == keys[i - 1]) continue;
        out_pairs[2 * kept] = s;
        out_pairs[2 * kept + 1] = d;
        kept++;
    }
    free(keys);
    return kept;
}

/* ------------------------------------------------------------------------------------------------
 * Engine restatement
 * ---------------------------------------------------------------------------------------------- */

enum { WO_DEEPWALK = 0, WO_NODE2VEC = 1 };
enum { WO_INIT_RANDOM = 0, WO_INIT_BURNIN = 1, WO_INIT_WEIGHT = 2 };

typedef struct
{
    u32    walks_per_vertex;   /* config::walks_per_vertex  (config/globals.h:7)  */
    u32    walk_length;        /* config::walk_length       (config/globals.h:10) */
    u32    model;              /* config::random_walk_model (config/globals.h:14) */
    u32    deterministic;      /* config::deterministic_mode(config/globals.h:29) */
    float  param_p;            /* config::paramP */
    float  param_q;            /* config::paramQ */
    u32    init_strategy;      /* config::sampler_init_strategy */
    u32    reserved;
    u64    seed;               /* keyed-MH seed (no reference counterpart: the reference seeds from time()) */
} wo_config;

typedef struct
{
    wo_config cfg;
    u64  n;                    /* number_of_vertices() (graph/wharfmh.h:117-123) */
    u64  m;                    /* number_of_edges()    (graph/wharfmh.h:130-133) */
    u64* row_ptr;              /* n+1 */
    u32* col;                  /* ascending, unique per row: what flatten_graph yields (wharfmh.h:175-208) */
    u32* version;              /* per-vertex sampler-manager generation (graph/vertex.h:12; reset at wharfmh.h:504->539) */
    u32* walks;                /* [n*wpv][L], WO_NO_VERTEX padded */
    u64* draw;                 /* X[wpv][L] */
    u64* affected;             /* last update: sorted affected walk ids */
    u32* affected_pos;         /* last update: earliest affected position per affected walk */
    u64  n_affected;
    u64  steps_last;           /* walker steps executed by the last generate / update */
} wo_engine;

static inline u64 deg_of(const wo_engine* e, u32 v) { return e->row_ptr[v + 1] - e->row_ptr[v]; }

static int has_edge(const wo_engine* e, u32 from, u32 to)                    /* node2vec.h:112-119 */
{
    u64 lo = e->row_ptr[from], hi = e->row_ptr[from + 1];
    while (lo < hi)
    {
        u64 mid = lo + (hi - lo) / 2;
        u32 c = e->col[mid];
        if (c == to) return 1;
        if (c < to) lo = mid + 1; else hi = mid;
    }
    return 0;
}

/* --- keyed counter RNG for MH mode (engine-defined; mirrored bit for bit by csrc/wharf_rng.cuh) --- */
static inline u64 sm64(u64 z)
{
    z += UINT64_C(0x9E3779B97F4A7C15);
    z = (z ^ (z >> 30)) * UINT64_C(0xBF58476D1CE4E5B9);
    z = (z ^ (z >> 27)) * UINT64_C(0x94D049BB133111EB);
    return z ^ (z >> 31);
}
enum { DOM_STEP = 1, DOM_STATE = 2, DOM_SAMPLER = 3 };
WO_API u64 wo_keyed(u64 seed, u64 dom, u64 a, u64 b, u64 c)
{
    u64 h = sm64(seed ^ (dom * UINT64_C(0xD6E8FEB86659FD93)));
    h = sm64(h ^ a);
    h = sm64(h ^ b);
    h = sm64(h ^ c);
    return h;
}

/* RandomWalkModel::weight (deepwalk.h:67-70, node2vec.h:74-88); state = (cur, prev). */
static float model_weight(const wo_engine* e, u32 prev, u32 x)
{
    if (e->cfg.model == WO_DEEPWALK) return 1.0f;
    if (x == prev) return 1 / e->cfg.param_p;
    if (has_edge(e, prev, x)) return 1;
    return 1 / e->cfg.param_q;
}

/* MetropolisHastingsSampler::init (metropolis_hastings_sampler.h:69-108) with draws keyed
 * (seed, cur, aux, version[cur], i).  Because SamplerManager::find returns a COPY
 * (libcuckoo cuckoohash_map.hh:596-605; call sites wharfmh.h:301,833) the stored chain state never moves
 * past this value: it is a pure function of its key. */
static u32 sampler_init_value(const wo_engine* e, u32 cur, u32 aux)
{
    const u64 d = deg_of(e, cur);
    const u64* row = NULL; (void) row;
    const u32* nb = e->col + e->row_ptr[cur];
    const u32 prev = aux;
    const u64 ver = e->version[cur];
    u64 i = 0;
    u32 last = nb[irand_from(wo_keyed(e->cfg.seed, DOM_SAMPLER, cur, ((u64) aux << 32) | ver, i++), d)];   /* RANDOM :72-75 */
    if (e->cfg.init_strategy == WO_INIT_BURNIN)                                                   /* :77-85 */
    {
        for (int it = 0; it < 100; it++)
        {
            u32 cand = nb[irand_from(wo_keyed(e->cfg.seed, DOM_SAMPLER, cur, ((u64) aux << 32) | ver, i++), d)];
            float wn = model_weight(e, prev, cand), wo = model_weight(e, prev, last);
            int acc = wo < wn;
            if (!acc) acc = bits_to_unit_double(wo_keyed(e->cfg.seed, DOM_SAMPLER, cur, ((u64) aux << 32) | ver, i++)) <= (double) wn / (double) wo;
            if (acc) last = cand;
        }
    }
    else if (e->cfg.init_strategy == WO_INIT_WEIGHT)                                              /* :87-107 */
    {
        u32 best = last;
        float best_w = model_weight(e, prev, best);
        for (int it = 0; it < 20; it++)
        {
            u32 cand = nb[irand_from(wo_keyed(e->cfg.seed, DOM_SAMPLER, cur, ((u64) aux << 32) | ver, i++), d)];
            float w = model_weight(e, prev, cand);
            if (w > best_w) { best_w = w; best = cand; }
        }
        last = best;
    }
    return last;
}

/* MetropolisHastingsSampler::sample used as a CHAIN (metropolis_hastings_sampler.h:31-46; fixture tests/sampler.cpp:86-127):
 * one sampler object of a fixed state (cur, aux) sampled `iters` times, its stored vertex advancing with every accepted
 * proposal - unlike the copies generate_initial_random_walks samples from.  Starts from the init value; draws keyed
 * (seed, DOM_CHAIN, cur, aux, 2 i) for the proposal and (..., 2 i + 1) for the acceptance test.
 * counts[j] += 1 for every sample that returns neighbour j of cur.  Returns the init value. */
enum { DOM_CHAIN = 4 };
WO_API u32 wo_sampler_chain(const wo_engine* e, u32 cur, u32 aux, u64 iters, u64* counts)
{
    const u64 d = deg_of(e, cur);
    if (d == 0) return WO_NO_VERTEX;
    const u32* nb = e->col + e->row_ptr[cur];
    const u32 init = sampler_init_value(e, cur, aux);
    u32 last = init;
    u64 last_slot = 0;
    while (nb[last_slot] != last) last_slot++;
    for (u64 i = 0; i < iters; i++)
    {
        const u64 cslot = irand_from(wo_keyed(e->cfg.seed, DOM_CHAIN, cur, aux, 2 * i), d);
        const u32 cand = nb[cslot];
        const float wn = model_weight(e, aux, cand), wo = model_weight(e, aux, last);
        int acc = wo < wn;
        if (!acc) acc = bits_to_unit_double(wo_keyed(e->cfg.seed, DOM_CHAIN, cur, aux, 2 * i + 1)) <= (double) wn / (double) wo;
        if (acc) { last = cand; last_slot = cslot; }
        counts[last_slot]++;
    }
    return init;
}

/* One walker, positions [start_pos, L): the loop body of wharfmh.h:294-325 (init) and :825-856 (re-walk).
 * row[start_pos] must already hold the vertex the walker stands on.  `prev` is only read in MH node2vec. */
static u64 walk_from(wo_engine* e, u64 walk_id, u32 start_pos, u32 prev)
{
    const u32 L = e->cfg.walk_length;
    u32* row = e->walks + walk_id * L;
    const u64* X = e->draw + (walk_id / e->n) * L;            /* Random(walk_id / n), restarted at draw 0 */
    u32 cur = row[start_pos];
    u64 steps = 0;
    for (u32 pos = start_pos; pos < L; pos++)
    {
        const u64 d = deg_of(e, cur);
        if (d == 0)                                           /* documented deviation: stop at a sink */
        {
            for (u32 q = pos + 1; q < L; q++) row[q] = WO_NO_VERTEX;
            break;
        }
        steps++;
        u32 next;
        const u32* nb = e->col + e->row_ptr[cur];
        if (e->cfg.deterministic)
        {
            next = nb[irand_from(X[pos - start_pos], d)];     /* wharfmh.h:307 / :838 */
        }
        else
        {
            /* MetropolisHastingsSampler::sample (metropolis_hastings_sampler.h:31-46) on a copy whose
             * last_sampled_vertex is the init value; accept (:118-122). */
            u32 aux  = e->cfg.model == WO_NODE2VEC ? prev : cur;
            u32 cand = nb[irand_from(wo_keyed(e->cfg.seed, DOM_STEP, walk_id, pos, 0), d)];
            u32 last = sampler_init_value(e, cur, aux);
            float wn = model_weight(e, aux, cand);
            float wo = model_weight(e, aux, last);
            int acc = wo < wn;
            if (!acc) acc = bits_to_unit_double(wo_keyed(e->cfg.seed, DOM_STEP, walk_id, pos, 1)) <= (double) wn / (double) wo;
            next = acc ? cand : last;
        }
        if (pos + 1 < L) row[pos + 1] = next;                 /* the last sample is unobservable (:315-317) */
        prev = cur;
        cur = next;
    }
    return steps;
}

WO_API wo_engine* wo_create(u64 n, u64 m, const u64* offsets_n, const u32* edges, const wo_config* cfg)
{
    wo_engine* e = (wo_engine*) calloc(1, sizeof(wo_engine));
    e->cfg = *cfg;
    e->n = n;
    e->m = m;
    e->row_ptr = (u64*) malloc((n + 1) * sizeof(u64));
    memcpy(e->row_ptr, offsets_n, n * sizeof(u64));           /* n offsets, last degree = m - offsets[n-1] (wharfmh.h:74-75) */
    e->row_ptr[n] = m;
    e->col = (u32*) malloc((m ? m : 1) * sizeof(u32));
    memcpy(e->col, edges, m * sizeof(u32));
    e->version = (u32*) calloc(n ? n : 1, sizeof(u32));
    e->draw = (u64*) malloc((u64) cfg->walks_per_vertex * cfg->walk_length * sizeof(u64) + 8);
    wo_draw_table(cfg->walks_per_vertex, cfg->walk_length, e->draw);
    return e;
}

WO_API void wo_destroy(wo_engine* e)
{
    if (!e) return;
    free(e->row_ptr); free(e->col); free(e->version); free(e->walks); free(e->draw);
    free(e->affected); free(e->affected_pos); free(e);
}

static void generate_range(u64 lo, u64 hi, void* ctx, u64* acc)
{
    wo_engine* e = (wo_engine*) ctx;
    const u32 L = e->cfg.walk_length;
    for (u64 w = lo; w < hi; w++)
    {
        u32* row = e->walks + w * L;
        u32 start = (u32) (w % e->n);                         /* :277,292 */
        row[0] = start;
        if (deg_of(e, start) == 0)                            /* :277-286: single sentinel entry */
        {
            for (u32 q = 1; q < L; q++) row[q] = WO_NO_VERTEX;
            continue;
        }
        u32 prev = start;
        if (!e->cfg.deterministic && e->cfg.model == WO_NODE2VEC)            /* node2vec.h:42-49 */
            prev = e->col[e->row_ptr[start] + irand_from(wo_keyed(e->cfg.seed, DOM_STATE, w, 0, 0), deg_of(e, start))];
        *acc += walk_from(e, w, 0, prev);
    }
}

/* generate_initial_random_walks (graph/wharfmh.h:250-356). */
WO_API u64 wo_generate(wo_engine* e)
{
    const u64 W = e->n * e->cfg.walks_per_vertex;
    const u32 L = e->cfg.walk_length;
    free(e->walks);
    e->walks = (u32*) malloc((W * L ? W * L : 1) * sizeof(u32));
    u64 steps = wo_parallel_for(W, 1024, generate_range, e);
    e->steps_last = steps;
    return steps;
}

typedef struct { wo_engine* e; const unsigned char* touched; u32* minpos; } scan_ctx;
static void scan_range(u64 lo, u64 hi, void* ctx, u64* acc)
{
    scan_ctx* c = (scan_ctx*) ctx;
    const u32 L = c->e->cfg.walk_length;
    for (u64 w = lo; w < hi; w++)
    {
        const u32* row = c->e->walks + w * L;
        u32 mp = WO_NO_VERTEX;
        for (u32 p = 0; p < L && row[p] != WO_NO_VERTEX; p++)
            if (c->touched[row[p]]) { mp = p; break; }
        c->minpos[w] = mp;
        *acc += mp != WO_NO_VERTEX;
    }
}

static void rewalk_range(u64 lo, u64 hi, void* ctx, u64* acc)
{
    wo_engine* e = (wo_engine*) ctx;
    const u32 L = e->cfg.walk_length;
    for (u64 j = lo; j < hi; j++)
    {
        u64 w = e->affected[j];
        u32 p = e->affected_pos[j];
        u32* row = e->walks + w * L;
        u32 prev = row[p];
        if (e->cfg.model == WO_NODE2VEC && !e->cfg.deterministic)
        {
            if (p > 0) prev = row[p - 1];                                    /* :819-823 */
            else if (deg_of(e, row[0]) > 0)                                  /* initial_state again (:817) */
                prev = e->col[e->row_ptr[row[0]] + irand_from(wo_keyed(e->cfg.seed, DOM_STATE, w, 0, 0), deg_of(e, row[0]))];
        }
        *acc += walk_from(e, w, p, prev);
    }
}

static u64 sort_unique_keys(u64* keys, u64 cnt)
{
    qsort(keys, cnt, sizeof(u64), cmp_u64);
    u64 k = 0;
    for (u64 i = 0; i < cnt; i++)
        if (i == 0 || keys[i] != keys[i - 1]) keys[k++] = keys[i];
    return k;
}

/* insert_edges_batch / delete_edges_batch (graph/wharfmh.h:439-576 / 588-726) followed by
 * batch_walk_update (:733-923).  `pairs` = cnt x (src,dst).  Returns 0, or -1 when an id is >= n. */
WO_API int wo_update(wo_engine* e, const u32* pairs, u64 cnt, int is_delete, int remove_dups, int apply_walk_updates)
{
    const u64 n = e->n;
    const u32 L = e->cfg.walk_length;
    for (u64 i = 0; i < 2 * cnt; i++) if (pairs[i] >= n) return -1;

    /* sort by (src,dst) (:452, :1056-1104); self-loops dropped only with remove_dups (:456-470).
     * Rows are sets, so duplicates collapse either way. */
    u64* keys = (u64*) malloc((cnt ? cnt : 1) * sizeof(u64));
    u64 k = 0;
    for (u64 i = 0; i < cnt; i++)
    {
        if (remove_dups && pairs[2 * i] == pairs[2 * i + 1]) continue;
        keys[k++] = ((u64) pairs[2 * i] << 32) | pairs[2 * i + 1];
    }
    k = sort_unique_keys(keys, k);

    /* touched = distinct sources (:475-481); each gets a fresh SamplerManager (:504 -> :539). */
    unsigned char* touched = (unsigned char*) calloc(n ? n : 1, 1);
    for (u64 i = 0; i < k; i++)
    {
        u32 s = (u32) (keys[i] >> 32);
        if (!touched[s]) { touched[s] = 1; e->version[s]++; }
    }

    /* row-wise sorted-set union (tree_plus uniont, :511) or difference old \ batch (:659). */
    u64* new_ptr = (u64*) malloc((n + 1) * sizeof(u64));
    u32* new_col = (u32*) malloc((e->m + k + 1) * sizeof(u32));
    u64 out = 0, b = 0;
    for (u64 v = 0; v < n; v++)
    {
        new_ptr[v] = out;
        u64 a = e->row_ptr[v], a_end = e->row_ptr[v + 1];
        u64 b_end = b;
        while (b_end < k && (keys[b_end] >> 32) == v) b_end++;
        if (!is_delete)
        {
            while (a < a_end || b < b_end)
            {
                u32 x;
                if (b >= b_end) x = e->col[a++];
                else if (a >= a_end) x = (u32) keys[b++];
                else
                {
                    u32 ca = e->col[a], cb = (u32) keys[b];
                    if (ca < cb) { x = ca; a++; } else if (cb < ca) { x = cb; b++; } else { x = ca; a++; b++; }
                }
                new_col[out++] = x;
            }
        }
        else
        {
            for (; a < a_end; a++)
            {
                u32 ca = e->col[a];
                while (b < b_end && (u32) keys[b] < ca) b++;
                if (b < b_end && (u32) keys[b] == ca) continue;
                new_col[out++] = ca;
            }
            b = b_end;
        }
    }
    new_ptr[n] = out;
    free(e->row_ptr); free(e->col);
    e->row_ptr = new_ptr; e->col = new_col; e->m = out;
    free(keys);

    free(e->affected); free(e->affected_pos);
    e->affected = NULL; e->affected_pos = NULL; e->n_affected = 0; e->steps_last = 0;
    if (!e->walks) { free(touched); return 0; }

    /* Map of changes (:507, :519-537): the index of a touched vertex v holds exactly the keys (w,p) with
     * walk[w][p] == v, so MoC[w] = min { p : walk[w][p] touched }. */
    const u64 W = n * e->cfg.walks_per_vertex;
    u32* minpos = (u32*) malloc((W ? W : 1) * sizeof(u32));
    scan_ctx sc = { e, touched, minpos };
    u64 na = wo_parallel_for(W, 4096, scan_range, &sc);
    free(touched);
    e->affected = (u64*) malloc((na ? na : 1) * sizeof(u64));
    e->affected_pos = (u32*) malloc((na ? na : 1) * sizeof(u32));
    for (u64 w = 0, j = 0; w < W; w++)
        if (minpos[w] != WO_NO_VERTEX) { e->affected[j] = w; e->affected_pos[j] = minpos[w]; j++; }
    e->n_affected = na;
    free(minpos);

    if (!apply_walk_updates) return 0;

    /* batch_walk_update (:761-859): drop the old suffix, re-walk from min_pos with the stream restarted. */
    e->steps_last = wo_parallel_for(na, 256, rewalk_range, e);
    return 0;
}

WO_API u64        wo_num_vertices(const wo_engine* e)  { return e->n; }
WO_API u64        wo_num_edges(const wo_engine* e)     { return e->m; }
WO_API const u64* wo_row_ptr(const wo_engine* e)       { return e->row_ptr; }
WO_API const u32* wo_col(const wo_engine* e)           { return e->col; }
WO_API const u32* wo_walks(const wo_engine* e)         { return e->walks; }
WO_API u64        wo_num_affected(const wo_engine* e)  { return e->n_affected; }
WO_API const u64* wo_affected(const wo_engine* e)      { return e->affected; }
WO_API const u32* wo_affected_pos(const wo_engine* e)  { return e->affected_pos; }
WO_API u64        wo_steps_last(const wo_engine* e)    { return e->steps_last; }
WO_API u32        wo_sampler_init_value(const wo_engine* e, u32 cur, u32 aux) { return sampler_init_value(e, cur, aux); }

/* Checksum of checksums: the wrapping sum over walks of the FNV-1a-64 of each walk's own text form "v0 v1 ... "
 * (graph/wharfmh.h:365-394, no line break).  A sum does not depend on the order of the walks, so the fingerprints of
 * the shards of a partitioned corpus add up to the fingerprint of the whole (tests compare wmh_corpus_fingerprint). */
WO_API u64 wo_corpus_fingerprint(const u32* walks, u64 n_walks, u32 L)
{
    u64 sum = 0;
    char buf[16];
    for (u64 w = 0; w < n_walks; w++)
    {
        u64 h = UINT64_C(0xcbf29ce484222325);
        for (u32 p = 0; p < L; p++)
        {
            u32 v = walks[w * L + p];
            if (v == WO_NO_VERTEX) break;
            int len = 0;
            do { buf[len++] = (char) ('0' + v % 10); v /= 10; } while (v);
            while (len) { h ^= (unsigned char) buf[--len]; h *= UINT64_C(0x100000001b3); }
            h ^= (unsigned char) ' '; h *= UINT64_C(0x100000001b3);
        }
        sum += h;
    }
    return sum;
}

/* FNV-1a-64 over the text form the reference's walk() emits ("v0 v1 ... " + '\n', wharfmh.h:365-394):
 * the corpus fingerprint used by the golden vectors of SURVEY.md 8c. */
WO_API u64 wo_corpus_fnv1a(const u32* walks, u64 n_walks, u32 L)
{
    u64 h = UINT64_C(0xcbf29ce484222325);
    char buf[16];
    for (u64 w = 0; w < n_walks; w++)
    {
        for (u32 p = 0; p < L; p++)
        {
            u32 v = walks[w * L + p];
            if (v == WO_NO_VERTEX) break;
            int len = 0;
            do { buf[len++] = (char) ('0' + v % 10); v /= 10; } while (v);
            while (len) { h ^= (unsigned char) buf[--len]; h *= UINT64_C(0x100000001b3); }
            h ^= (unsigned char) ' '; h *= UINT64_C(0x100000001b3);
        }
        h ^= (unsigned char) '\n'; h *= UINT64_C(0x100000001b3);
    }
    return h;
}
