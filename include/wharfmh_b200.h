/*
 * wharfmh_b200.h - C ABI of the B200-native WharfMH hot path (libwharfmh_b200.so).
 *
 * Drop-in boundary: the reference has no FFI; its boundary is the public C++ surface of
 * dygrl::WharfMH (graph/wharfmh.h:21) plus the mutable globals of namespace config
 * (config/globals.h:7-29).  Each entry point below names the reference member it replaces
 * (file:line relative to the reference root).  include/wharfmh_b200.hpp rebuilds the C++ class on top
 * of these functions; INTEGRATION.md shows the binding a maintainer of the reference would add.
 *
 * Conventions
 *   - plain pointers and sizes only; every function returns a wmh_status (0 = OK) unless noted;
 *     wmh_last_error() returns the message of the last failure on the calling thread
 *     (reference: std::exit(1) / abort() / exceptions / SIGFPE).
 *   - one host thread drives one engine; an engine owns one CUDA stream on one device
 *     (reference: process-global state, one engine per process).
 *   - `host` pointers are ordinary (pageable or pinned) host memory; `_device` variants take device
 *     pointers on the engine's device and never touch the host.
 *   - walk ids are GLOBAL ids (walk_id = replica * n + start_vertex, graph/wharfmh.h:275-292); with
 *     world > 1 an engine owns whole replicas first and, of the remaining replicas, the walks of one start
 *     vertex out of every block of `world` consecutive ids (wmh_walk_owner, wmh_local_walk_ids); the graph
 *     is replicated.
 *   - there is no CPU fallback: every call fails with WMH_ERR_CUDA when no sm_100-class device is usable.
 */
#ifndef WHARFMH_B200_H
#define WHARFMH_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WMH_API __attribute__((visibility("default")))

/* Vertex value used for "no vertex" in retrieved walks (positions after a walker stopped). */
#define WMH_NO_VERTEX 0xFFFFFFFFu
/* The reference's end-of-walk sentinel as stored in its index (graph/wharfmh.h:283,317). */
#define WMH_REF_SENTINEL 0xFFFFFFFEu

typedef enum
{
    WMH_OK = 0,
    WMH_ERR_INVALID = 1,      /* bad argument (null pointer, id >= n, walk_length == 0 ...) */
    WMH_ERR_CUDA = 2,         /* CUDA runtime failure or no usable device */
    WMH_ERR_OOM = 3,          /* device allocation failed */
    WMH_ERR_STATE = 4,        /* call not valid in the current state (e.g. walks not generated yet) */
    WMH_ERR_DOMAIN = 5,       /* value outside the representable domain (pairing overflow, > 2^32 pool) */
    WMH_ERR_RANGE = 6         /* caller buffer too small / id out of range */
} wmh_status;

/* types::RandomWalkModelType (config/types.h:31) */
typedef enum { WMH_DEEPWALK = 0, WMH_NODE2VEC = 1 } wmh_model;
/* types::SamplerInitStartegy (config/types.h:34) */
typedef enum { WMH_INIT_RANDOM = 0, WMH_INIT_BURNIN = 1, WMH_INIT_WEIGHT = 2 } wmh_sampler_init;
/* How affected walks are found after an edge batch (walks/inverted_index.h + graph/wharfmh.h:519-537). */
typedef enum
{
    WMH_SEARCH_AUTO = 0,      /* index when few index entries are touched, corpus scan otherwise */
    WMH_SEARCH_INDEX = 1,     /* device radix-sorted inverted index + segmented search */
    WMH_SEARCH_SCAN = 2       /* stream the walk matrix against the touched-vertex bitmap */
} wmh_search_mode;

typedef enum
{
    WMH_GENERATE_AUTO = 0,       /* coalesced when the adjacency does not fit the L2, one walker per thread otherwise */
    WMH_GENERATE_PER_WALKER = 1, /* one walker per thread, two dependent random loads per step */
    WMH_GENERATE_COALESCED = 2   /* step-synchronous: walkers of a replica that met are advanced once (deterministic mode only) */
} wmh_generate_mode;

/* The knobs of namespace config (config/globals.h:7-29) plus placement.  Read at call time, like the
 * reference reads its globals (graph/wharfmh.h:254-272). */
typedef struct
{
    uint32_t walks_per_vertex;   /* config::walks_per_vertex (uint8_t in the reference: 1..255) */
    uint32_t walk_length;        /* config::walk_length      (uint8_t in the reference: 1..255) */
    uint32_t model;              /* wmh_model: config::random_walk_model */
    uint32_t deterministic;      /* config::deterministic_mode: 1 = per-walk xoroshiro stream, bit-exact vs reference */
    float    param_p;            /* config::paramP */
    float    param_q;            /* config::paramQ */
    uint32_t init_strategy;      /* wmh_sampler_init: config::sampler_init_strategy */
    uint32_t search_mode;        /* wmh_search_mode */
    uint64_t seed;               /* keyed-MH seed (reference: utility::Random(std::time(nullptr)), config/globals.h:26) */
    int32_t  device;             /* CUDA device ordinal */
    uint32_t rank;               /* shard of the walk corpus owned by this engine */
    uint32_t world;              /* number of shards (>= 1) */
    uint32_t generate_mode;      /* wmh_generate_mode: how deterministic bulk generation is scheduled (same corpus either way) */
} wmh_config;

typedef struct wmh_engine wmh_engine;

/* Device-memory accounting: the counterpart of WharfMH::memory_footprint (graph/wharfmh.h:928-998). */
typedef struct
{
    uint64_t graph_bytes;        /* row descriptors + neighbour pool (in use) */
    uint64_t graph_pool_bytes;   /* neighbour pool capacity */
    uint64_t walks_bytes;        /* position-major walk matrix */
    uint64_t index_bytes;        /* inverted index (0 when not built) */
    uint64_t scratch_bytes;      /* per-batch scratch currently held */
    uint64_t total_bytes;
} wmh_memory;

/* The four global timers of config/config.h:10-14, in seconds, accumulated since the last reset. */
typedef enum
{
    WMH_TIMER_GRAPH_UPDATE_INSERT = 0,
    WMH_TIMER_WALK_UPDATE_INSERT = 1,
    WMH_TIMER_GRAPH_UPDATE_DELETE = 2,
    WMH_TIMER_WALK_UPDATE_DELETE = 3,
    WMH_TIMER_GENERATE = 4        /* device time of the last generate_initial_random_walks */
} wmh_timer;

/* ---- lifecycle ---------------------------------------------------------------------------------- */

WMH_API void wmh_default_config(wmh_config* cfg);                        /* config/globals.h defaults */
WMH_API const char* wmh_last_error(void);
WMH_API const char* wmh_version(void);

/* WharfMH(long n, long m, uintE* offsets, uintV* edges, bool free_memory) - graph/wharfmh.h:58.
 * `offsets` has n entries (no n+1-th; the last degree is m - offsets[n-1], wharfmh.h:74-75); rows must be
 * sorted and unique.  The arrays are copied, never freed (the C++ shim frees them when free_memory). */
WMH_API int wmh_create(wmh_engine** out, uint64_t n, uint64_t m, const uint64_t* offsets, const uint32_t* edges,
                       const wmh_config* cfg);
/* Same, CSR already resident on the engine's device. */
WMH_API int wmh_create_device(wmh_engine** out, uint64_t n, uint64_t m, const uint64_t* d_offsets,
                              const uint32_t* d_edges, const wmh_config* cfg);
/* WharfMH(long n, long m) - graph/wharfmh.h:26: n isolated vertices. */
WMH_API int wmh_create_empty(wmh_engine** out, uint64_t n, const wmh_config* cfg);
/* WharfMH::destroy - graph/wharfmh.h:228 (also releases the handle). */
WMH_API void wmh_destroy(wmh_engine* e);
/* WharfMH::destroy_index - graph/wharfmh.h:237: drops the walk corpus and its inverted index. */
WMH_API int wmh_destroy_index(wmh_engine* e);
/* Assigning config::* between calls (experiments/src/throughput-latency.cpp:19-75). */
WMH_API int wmh_set_config(wmh_engine* e, const wmh_config* cfg);
WMH_API int wmh_get_config(const wmh_engine* e, wmh_config* cfg);

/* ---- graph -------------------------------------------------------------------------------------- */

WMH_API uint64_t wmh_number_of_vertices(const wmh_engine* e);            /* graph/wharfmh.h:117 */
WMH_API uint64_t wmh_number_of_edges(const wmh_engine* e);               /* graph/wharfmh.h:130 */
/* WharfMH::flatten_graph - graph/wharfmh.h:175 / FlatGraph (graph/snapshot.h:108-194):
 * row_ptr has n+1 entries, col has number_of_edges() entries, rows ascending. */
WMH_API int wmh_flatten_graph(wmh_engine* e, uint64_t* row_ptr, uint32_t* col);
WMH_API int wmh_degree(wmh_engine* e, uint32_t vertex, uint64_t* degree);
WMH_API int wmh_neighbors(wmh_engine* e, uint32_t vertex, uint32_t* out, uint64_t capacity, uint64_t* degree);

/* ---- walks -------------------------------------------------------------------------------------- */

/* WharfMH::generate_initial_random_walks - graph/wharfmh.h:250. */
WMH_API int wmh_generate_initial_random_walks(wmh_engine* e);
/* The same pass with the corpus streamed out while it is generated (what vertex-classification.cpp:167-178 does
 * after generation, walk by walk): out receives number_of_local_walks rows of walk_length vertices in local walk
 * order, WMH_NO_VERTEX after a stopped walker.  The device keeps the corpus as after
 * wmh_generate_initial_random_walks.  Pinned host memory lets the copies overlap the walk kernels. */
WMH_API int wmh_generate_initial_random_walks_to_host(wmh_engine* e, uint32_t* out, uint64_t capacity_walks);
/* Walker steps executed by the last generate / batch update (a step = one iteration of wharfmh.h:294). */
WMH_API uint64_t wmh_last_steps(const wmh_engine* e);
WMH_API uint64_t wmh_number_of_walks(const wmh_engine* e);               /* global: n * walks_per_vertex */
WMH_API uint64_t wmh_number_of_local_walks(const wmh_engine* e);         /* owned by this rank */

/* WharfMH::insert_edges_batch / delete_edges_batch - graph/wharfmh.h:439 / 588.
 * `edges` = m pairs (src, dst) in host memory.  When !sorted the caller's array is sorted in place by
 * (src, dst), as the reference does (wharfmh.h:452).  remove_dups also drops self-loops (wharfmh.h:456-470).
 * `nn` is accepted for signature parity (sort-key width in the reference) and ignored.
 * *n_affected receives the number of affected walks owned by this rank; fetch them with wmh_affected_walks. */
WMH_API int wmh_insert_edges_batch(wmh_engine* e, uint64_t m, uint32_t* edges, int sorted, int remove_dups,
                                   uint64_t nn, int apply_walk_updates, uint64_t* n_affected);
WMH_API int wmh_delete_edges_batch(wmh_engine* e, uint64_t m, uint32_t* edges, int sorted, int remove_dups,
                                   uint64_t nn, int apply_walk_updates, uint64_t* n_affected);
/* Same with the batch resident on the device (read-only; not sorted in place). */
WMH_API int wmh_insert_edges_batch_device(wmh_engine* e, uint64_t m, const uint32_t* d_edges, int remove_dups,
                                          int apply_walk_updates, uint64_t* n_affected);
WMH_API int wmh_delete_edges_batch_device(wmh_engine* e, uint64_t m, const uint32_t* d_edges, int remove_dups,
                                          int apply_walk_updates, uint64_t* n_affected);

/* The pbbs::sequence<WalkID> returned by insert/delete_edges_batch (wharfmh.h:548,698) as global ids in
 * ascending order (the reference returns cuckoo iteration order: compare as sets), with the earliest
 * affected position of each (types::MapOfChanges, config/types.h:37).  Either output may be NULL. */
WMH_API int wmh_affected_walks(wmh_engine* e, uint64_t* ids, uint32_t* positions, uint64_t capacity);
/* Device views of the same (local walk index -> use wmh_local_to_global_walk): valid until the next batch. */
WMH_API int wmh_affected_walks_device(wmh_engine* e, const uint32_t** d_local_ids, const uint32_t** d_positions,
                                      uint64_t* count);
/* MetropolisHastingsSampler::sample - walks/metropolis_hastings_sampler.h:31-46 - as the reference's own test drives it
 * (tests/sampler.cpp:86-127): ONE sampler object of the fixed state (cur, prev), initialised with the engine's
 * sampler_init_strategy (:69-108) and sampled `iters` times, its stored vertex advancing with every accepted proposal.
 * counts[j] receives the number of samples that returned neighbour j of `cur` (capacity >= degree(cur), row order);
 * *init_value the vertex the sampler started from.  Model, p, q and seed are the engine's; draws are keyed
 * (seed, cur, prev, sample number), so the chain is reproducible.  For DeepWalk the weight is 1 and `prev` is unused. */
WMH_API int wmh_sampler_chain(wmh_engine* e, uint32_t cur, uint32_t prev, uint64_t iters, uint32_t* init_value,
                              uint64_t* counts, uint64_t capacity);

/* The same ids (global, ascending) written to DEVICE memory (capacity >= the affected count): what a multi-rank caller
 * hands to its all-gather without a trip through the host.  Complete when the call returns. */
WMH_API int wmh_affected_walks_global_device(wmh_engine* e, uint64_t* d_ids, uint64_t capacity);
/* Distinct source vertices of the last batch (the vertices whose walks were affected), ascending. */
WMH_API int wmh_touched_vertices(wmh_engine* e, uint32_t* out, uint64_t capacity, uint64_t* count);

/* No counterpart in the reference: added for callers that cannot allocate (code under
 * CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  The corpus scan of the affected-walk search
 * (types::MapOfChanges) with its scratch - the touched-vertex bitmap - in a static device array instead of an allocation:
 * d_min_pos[w] (device, wmh_number_of_local_walks entries) receives the first position at which local walk w stands on
 * one of the `count` vertices in d_vertices (device; ids >= number_of_vertices are ignored), 0xFFFFFFFF when it never does -
 * for the source vertices of a batch that is exactly what wmh_affected_walks reports.  Calls no cudaMalloc / cudaFree / pool or
 * library allocation; runs on the engine's stream and is complete when it returns.  Limits: at most 2^25 = 33554432 vertices
 * (WMH_ERR_RANGE beyond; the 4 MiB array is held by every GPU that loads the library until the process exits); all calls
 * on a device share the array, so calls that overlap on the GPU (two engines, two threads) race - serialise them. */
WMH_API int wmh_scan_touched_no_alloc(wmh_engine* e, const uint32_t* d_vertices, uint64_t count, uint32_t* d_min_pos);

/* WharfMH::walk - graph/wharfmh.h:365: vertices of one walk; *length <= walk_length. */
WMH_API int wmh_walk(wmh_engine* e, uint64_t walk_id, uint32_t* vertices, uint32_t* length);
/* The string form walk() returns: "v0 v1 ... v79 " (trailing space).  Returns needed size in *needed. */
WMH_API int wmh_walk_string(wmh_engine* e, uint64_t walk_id, char* buf, size_t capacity, size_t* needed);
/* WharfMH::vertex_at_walk - graph/wharfmh.h:404. */
WMH_API int wmh_vertex_at_walk(wmh_engine* e, uint64_t walk_id, uint32_t position, uint32_t* vertex);
/* Bulk retrieval of local walks [first, first+count) (local order = ascending global id), walk-major
 * rows of walk_length vertices padded with WMH_NO_VERTEX. */
WMH_API int wmh_walks_copy(wmh_engine* e, uint64_t first, uint64_t count, uint32_t* out);
/* Bulk retrieval of selected local walks (e.g. the affected ones: vertex-classification.cpp:167-178). */
WMH_API int wmh_walks_gather(wmh_engine* e, const uint32_t* local_ids, uint64_t count, uint32_t* out);
WMH_API uint64_t wmh_local_to_global_walk(const wmh_engine* e, uint64_t local_walk);
WMH_API int64_t  wmh_global_to_local_walk(const wmh_engine* e, uint64_t walk_id);   /* -1 when not owned */
/* Sharding rule (config.world > 1).  The walks_per_vertex replicas of the corpus (replica = walk_id / n,
 * graph/wharfmh.h:291) are independent: with q = walks_per_vertex / world every rank owns q replicas entirely
 * (replicas [rank * q, rank * q + q)); the remaining walks_per_vertex - world * q replicas are sharded by start
 * vertex: vertices are taken in blocks of `world` consecutive ids and every rank owns one vertex of each block,
 * rotated from block to block so that dense and sparse id residues spread evenly.
 * wmh_walk_owner = the rank that owns walk_id; wmh_local_walk_ids = the global ids of the walks this rank owns, in
 * local walk order (ascending); returns their number. */
WMH_API uint32_t wmh_walk_owner(const wmh_engine* e, uint64_t walk_id);
WMH_API uint64_t wmh_local_walk_ids(const wmh_engine* e, uint64_t* out, uint64_t capacity);
/* The same rule without an engine (pure host arithmetic, no device needed): a launcher can decide placement, and a
 * consumer can route a walk id to its owner, from (n, walks_per_vertex, world) alone. */
WMH_API uint32_t wmh_partition_owner(uint64_t n, uint32_t walks_per_vertex, uint32_t world, uint64_t walk_id);
WMH_API uint64_t wmh_partition_local_walk_ids(uint64_t n, uint32_t walks_per_vertex, uint32_t rank, uint32_t world,
                                              uint64_t* out, uint64_t capacity);
/* FNV-1a-64 fingerprint of the local corpus in walk() text form (computed on the device). */
WMH_API int wmh_corpus_fingerprint(wmh_engine* e, uint64_t* sum_of_row_hashes);

/* ---- inverted index (walks/inverted_index.h) ----------------------------------------------------- */

/* Builds (or rebuilds) the device index now; batch updates build it lazily when search_mode needs it.  No size cap beyond
 * memory (4 B per corpus cell + 8 B x (n+1) per chunk + 16 B per cell of one chunk while sorting): corpora of more than 2^31
 * cells per rank are indexed in chunks of whole positions, so cell ids stay 32-bit (the reference's u32 key walk_id*L+pos
 * wraps there). */
WMH_API int wmh_build_index(wmh_engine* e);
/* InvertedIndex::find_next - walks/inverted_index.h:74-92 on vertex `vertex`: next vertex of walk
 * `walk_id` standing on `vertex` at `position`; WMH_REF_SENTINEL at the last position,
 * *found = 0 when (walk_id, position) is not stored at that vertex. */
WMH_API int wmh_index_find_next(wmh_engine* e, uint32_t vertex, uint64_t walk_id, uint32_t position,
                                uint32_t* next, int* found);
/* Number of index entries stored at `vertex` (InvertedIndex::size()). */
WMH_API int wmh_index_size(wmh_engine* e, uint32_t vertex, uint64_t* entries);
/* The same for every vertex at once (entries[n]); all zero while no corpus exists (what flatten_vertex_tree,
 * graph/wharfmh.h:140, shows of each vertex's inverted_index after construction and after destroy_index). */
WMH_API int wmh_index_sizes(wmh_engine* e, uint64_t* entries);

/* ---- Szudzik pairing (walks/pairings.h) and the paired corpus (walks/compressed_walks.h) --------- */

/* Szudzik<T>::pair / unpair / pair_triplet / unpair_triplet (pairings.h:27,100,88,122) over device arrays.
 * *overflow (host) is set to the number of inputs whose code does not fit T (reference: silent wrap). */
WMH_API int wmh_szudzik_pair_u64(int device, const uint64_t* d_x, const uint64_t* d_y, uint64_t* d_out, uint64_t count, uint64_t* overflow);
WMH_API int wmh_szudzik_unpair_u64(int device, const uint64_t* d_z, uint64_t* d_x, uint64_t* d_y, uint64_t count);
WMH_API int wmh_szudzik_pair_u32(int device, const uint32_t* d_x, const uint32_t* d_y, uint32_t* d_out, uint64_t count, uint64_t* overflow);
WMH_API int wmh_szudzik_unpair_u32(int device, const uint32_t* d_z, uint32_t* d_x, uint32_t* d_y, uint64_t count);
WMH_API int wmh_szudzik_pair_triplet_u64(int device, const uint64_t* d_a, const uint64_t* d_b, const uint64_t* d_c, uint64_t* d_out, uint64_t count, uint64_t* overflow);
WMH_API int wmh_szudzik_unpair_triplet_u64(int device, const uint64_t* d_z, uint64_t* d_a, uint64_t* d_b, uint64_t* d_c, uint64_t count);
/* Host scalar forms (same arithmetic as the device ops; exact integer sqrt). */
WMH_API uint64_t wmh_szudzik_pair(uint64_t x, uint64_t y);
WMH_API void     wmh_szudzik_unpair(uint64_t z, uint64_t* x, uint64_t* y);

/* CompressedWalks (walks/compressed_walks.h:13-90) as flat device arrays: for every vertex v (CSR-style
 * `ptr`, n+1 entries) the ascending codes pair(walk_id*L + position, next) of the steps standing on v,
 * next = WMH_REF_SENTINEL at the last position.  Fused: encodes straight from the walk matrix + index.
 * Requires (global walks)*L < 2^32 (the u64 pairing domain).  Outputs are host arrays; codes may be NULL
 * to query sizes (ptr still filled). */
WMH_API int wmh_compressed_walks(wmh_engine* e, uint64_t* ptr, uint64_t* codes, uint64_t capacity, uint64_t* total);
/* CompressedWalks::find_next - walks/compressed_walks.h:52-86, evaluated on the device over the codes of `vertex`. */
WMH_API int wmh_compressed_find_next(wmh_engine* e, uint32_t vertex, uint64_t walk_id, uint32_t position,
                                     uint32_t* next, int* found);

/* ---- reporting ---------------------------------------------------------------------------------- */

WMH_API double wmh_timer_get(const wmh_engine* e, int which);            /* timer::get_total(), config/config.h:10-14 */
WMH_API void   wmh_timer_reset(wmh_engine* e, int which);                /* timer::reset() */
WMH_API int    wmh_memory_footprint(wmh_engine* e, wmh_memory* out);     /* graph/wharfmh.h:928 */
/* Kernel launches issued by this engine since creation (for bench.py's gpu_launches). */
WMH_API uint64_t wmh_kernel_launches(const wmh_engine* e);
/* Device time (ms) of the dominant kernel of the last generate / update (walk kernel), and its launches. */
WMH_API double   wmh_last_walk_kernel_ms(const wmh_engine* e);
/* Device milliseconds the last call spent on the inverted index (walks/inverted_index.h): its build inside
 * wmh_generate_initial_random_walks (graph/wharfmh.h:312-354) or its maintenance inside a batch (wharfmh.h:861-922);
 * 0 when the call did not touch the index (search_mode SCAN, or AUTO without an index). */
WMH_API double   wmh_last_index_ms(const wmh_engine* e);
/* How the last batch found its affected walks: WMH_SEARCH_INDEX or WMH_SEARCH_SCAN as TAKEN (0: no search ran). */
WMH_API uint32_t wmh_last_search_mode(const wmh_engine* e);
/* How node2vec's "is the proposal a neighbour of prev" test (walks/models/node2vec.h:74-88,112-119) is pre-filtered in
 * Metropolis-Hastings mode: 0 = no aux words yet (no MH node2vec corpus so far), 1 = per-row Bloom blocks (any graph),
 * 2 = triangle blocks keyed on the traversed edge (symmetric graph and symmetric batches only; a directed batch drops
 * the engine back to 1).  The filter only prunes exact searches: corpora are identical in every mode. */
WMH_API uint32_t wmh_neighbour_filter(const wmh_engine* e);
/* Self-check of that filter (tests): the number of edge slots whose filter block lacks a bit it must carry - triangle mode:
 * a member of N(v) n N(c) missing from the block of slot v -> c; per-row mode: a neighbour missing from its row's block.
 * Anything but 0 would make the walk kernel skip an exact search it needs. */
WMH_API int      wmh_filter_check(wmh_engine* e, uint64_t* missing);
/* Raw device pointers for zero-copy consumers (position-major walk matrix [L][local_walks]). */
WMH_API const uint32_t* wmh_device_walks(const wmh_engine* e);
WMH_API void*    wmh_stream(const wmh_engine* e);                        /* cudaStream_t */
WMH_API int      wmh_synchronize(wmh_engine* e);
/* Engines allocate stream-ordered from the library's own memory pool on the device (the host application's default
 * pool is left alone) and leave freed blocks there for the next engine / batch (the analogue of the reference's static allocator pools, graph/wharfmh.h:1041-1047, which are
 * never returned either).  This hands every unused block back to the driver. */
WMH_API int      wmh_trim_memory(int device);

/* ---- peer memory between the ranks of one box (no counterpart in the single-process reference) -- */

/* With config.world > 1 every rank holds a replica of the graph and owns the walks whose start vertex satisfies
 * start % world == rank.  Work that is a pure function of the replicated graph - the first-touch value of every
 * node2vec sampler (SamplerManager, graph/vertex.h:12 / metropolis_hastings_sampler.h:69-108) and, in
 * deterministic mode, the step of every class of walkers - is then computed ONCE per box instead of once per
 * rank: each rank computes a slice and its kernel stores the results straight into the exchange buffers of
 * all its peers over NVLink (peer-mapped device memory, no collective library on the data path).
 *
 *   wmh_peer_create  allocates this rank's exchange buffer (device memory, zeroed) and returns its 64-byte
 *                    CUDA IPC handle; the caller distributes the handles through any host channel
 *                    (torch.distributed all_gather, MPI, a file).
 *   wmh_peer_attach  maps the buffers of all `config.world` ranks: handles = world x 64 bytes, in rank order
 *                    (the own slot is ignored).  Every rank must have returned from wmh_peer_create before any
 *                    rank attaches, and all ranks must issue the same sequence of generate calls afterwards
 *                    (the kernels meet at a flag barrier in peer memory; a rank that never arrives makes the
 *                    others fail with WMH_ERR_CUDA after WMH_PEER_TIMEOUT_MS instead of hanging).
 *   wmh_peer_detach  unmaps the peers and frees the buffer (also done by wmh_destroy).
 * Results never depend on whether peers are attached: the corpus is bit-identical either way. */
#define WMH_PEER_HANDLE_BYTES 64
#define WMH_PEER_MAX_WORLD 8
#define WMH_PEER_TIMEOUT_MS 4000
WMH_API int wmh_peer_create(wmh_engine* e, uint64_t bytes, unsigned char handle_out[WMH_PEER_HANDLE_BYTES]);
WMH_API int wmh_peer_attach(wmh_engine* e, const unsigned char* handles);
WMH_API int wmh_peer_detach(wmh_engine* e);
/* Bytes of exchange buffer the current graph and config need (0 when no exchange applies). */
WMH_API uint64_t wmh_peer_bytes_needed(const wmh_engine* e);
/* Number of generate passes that used the peer exchange so far (0: every pass ran stand-alone). */
WMH_API uint64_t wmh_peer_exchanges(const wmh_engine* e);

/* ---- the generators, host scalar forms (no device needed) ------------------------------------------ */

/* utility::Random(seed) followed by `count` lrand() calls (utils/utility.h:157-170 seeding with arithmetic
 * shifts, :194-206 xoroshiro128+ 55/14/36).  Deterministic mode's draw table is row r = this stream for
 * seed r (graph/wharfmh.h:289-291: one Random(walk_id / n) per walk). */
WMH_API void wmh_random_stream(uint64_t seed, uint64_t count, uint64_t* out);
/* utility::Random::drand applied to one 64-bit draw (utils/utility.h:208-218): [0, 1). */
WMH_API double wmh_unit_double(uint64_t draw);
/* The counter-based generator of keyed MH mode: the draw for (seed, domain, a, b, c); domain 1 = walk step
 * (a = global walk id, b = position, c = draw number), 2 = node2vec initial state (a = walk id),
 * 3 = sampler init (a = cur, b = prev << 32 | generation of cur's sampler manager, c = proposal number). */
WMH_API uint64_t wmh_keyed_draw(uint64_t seed, uint64_t domain, uint64_t a, uint64_t b, uint64_t c);

/* ---- synthetic input (utils/utility.h:55-146 generate_batch_of_edges, rmat_util.h) --------------- */

/* Device restatement of utility::generate_batch_of_edges: RMAT(a,b,c) edges over 2^(log2_up(vertices)-1)
 * ids, optional symmetrisation, sort by (src,dst), duplicate and (unless self_loops) loop removal.
 * d_out must hold 2*edges_number*(directed?1:2) u32; *count receives the number of pairs kept. */
WMH_API int wmh_generate_batch_of_edges_device(int device, uint64_t edges_number, uint64_t vertices_number,
                                               uint64_t batch_seed, int self_loops, int directed,
                                               double a, double b, double c, uint32_t* d_out, uint64_t* count);
/* The same with the pairs returned to host memory (what utility::generate_batch_of_edges hands its callers,
 * throughput-latency.cpp:121-123): out holds `capacity` pairs, capacity >= edges_number*(directed?1:2). */
WMH_API int wmh_generate_batch_of_edges(int device, uint64_t edges_number, uint64_t vertices_number,
                                        uint64_t batch_seed, int self_loops, int directed,
                                        double a, double b, double c, uint32_t* out, uint64_t capacity, uint64_t* count);
/* Builds CSR (d_offsets[n+1] u64) from sorted (src,dst) pairs on the device. */
WMH_API int wmh_pairs_to_csr_device(int device, uint64_t n, const uint32_t* d_pairs, uint64_t count,
                                    uint64_t* d_offsets, uint32_t* d_edges);

/* ---- ingest and export: the text formats either side of the path (host code, no device needed) ---- */

/* Ligra "AdjacencyGraph" text (libs/compressed_trees/common/IO.h:67-106, read_unweighted_graph): the tokens
 * "AdjacencyGraph", n, m, n offsets, m neighbour ids.  Call with offsets = edges = NULL for *n and *m, then
 * with offsets[n] (no n+1-th entry, as the reference) and edges[m].  The token count must be n + m + 3
 * (IO.h:94 asserts it); offsets must be non-decreasing and <= m, neighbour ids < n. */
WMH_API int wmh_parse_adjacency_graph(const char* text, uint64_t bytes, uint64_t* n, uint64_t* m,
                                      uint64_t* offsets, uint32_t* edges);
/* The same format written out (what experiments/bin/SNAPtoAdj produces).  Returns the bytes needed; writes
 * only when capacity is enough (no terminating NUL). */
WMH_API uint64_t wmh_format_adjacency_graph(uint64_t n, uint64_t m, const uint64_t* offsets, const uint32_t* edges,
                                            char* out, uint64_t capacity);
/* SNAP edge list ("u v" per line; blank lines and lines starting with '#' or '%' skipped): the input of
 * SNAPtoAdj and of create_edge_stream (experiments/src/vertex-classification.cpp:5-38).  pairs receives
 * (u, v) in file order; *count = pairs in the text (also when pairs is NULL: call twice), *vertices =
 * largest id + 1.  WMH_ERR_RANGE when pairs is given but holds fewer than *count. */
WMH_API int wmh_parse_edge_list(const char* text, uint64_t bytes, uint32_t* pairs, uint64_t capacity,
                                uint64_t* count, uint64_t* vertices);
/* Rows of walk_length vertices (WMH_NO_VERTEX padded, as wmh_walks_copy / wmh_walks_gather return them) as
 * text: each walk as WharfMH::walk prints it (graph/wharfmh.h:365-394, every vertex followed by a space)
 * and a newline.  Returns the bytes needed; writes only when capacity is enough. */
WMH_API uint64_t wmh_format_walks(const uint32_t* rows, uint64_t count, uint32_t walk_length, char* out,
                                  uint64_t capacity);
/* walks.txt hand-off of the walks a batch changed (vertex-classification.cpp:167-178): gathers the named
 * walks (global ids held by this engine) from the device and writes them with wmh_format_walks. */
WMH_API int wmh_write_walks(wmh_engine* e, const uint64_t* walk_ids, uint64_t count, const char* path, int append,
                            uint64_t* bytes_written);

#ifdef __cplusplus
}
#endif
#endif /* WHARFMH_B200_H */
