// wharfmh_b200.hpp - the reference's C++ surface for the hot path, re-created over the C ABI.
//
// A translation unit that used to `#include <wharfmh.h>` from the reference (graph/wharfmh.h) can include
// this header instead and link libwharfmh_b200.so: the names, argument meaning and ownership rules of
// dygrl::WharfMH, namespace config, the four global timers and the small pbbs / utility helpers the
// reference's drivers use (experiments/src/throughput-latency.cpp:1-200) are kept.  All compute happens in
// the CUDA library behind include/wharfmh_b200.h; this header holds no algorithm.
//
// Differences a caller can observe (see DESIGN.md "Boundary"):
//   - errors throw std::runtime_error instead of std::exit(1) / SIGFPE;
//   - insert/delete_edges_batch return the affected walk ids ascending (reference: cuckoo iteration order);
//   - walk ids are 64-bit (types::WalkID is uint32_t in the reference and wraps beyond 5.37 M vertices).
#ifndef WHARFMH_B200_HPP
#define WHARFMH_B200_HPP

#include "wharfmh_b200.h"

#include <algorithm>
#include <array>
#include <chrono>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <fstream>
#include <iostream>
#include <limits>
#include <sstream>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#define dygrl dynamic_graph_representation_learning_with_metropolis_hastings   // config/config.h:4

using uintV = uint32_t;   // libs/compressed_trees/common/types.h:15-38 (vertex ids)
using uintE = uint64_t;   // -DEDGELONG (edge offsets)
using intV  = int32_t;

// ---- the slice of pbbslib the reference's callers touch ---------------------------------------------
namespace pbbs
{
    template <class T> T* new_array_no_init(size_t n) { return static_cast<T*>(std::malloc((n ? n : 1) * sizeof(T))); }   // utilities.h:158-170 (-DUSEMALLOC)
    template <class T> T* new_array(size_t n) { T* p = new_array_no_init<T>(n); for (size_t i = 0; i < n; i++) new (p + i) T(); return p; }
    inline void free_array(void* p) { std::free(p); }

    template <class T> size_t log2_up(T i) { size_t a = 0; T b = i - 1; while (b > 0) { b = b >> 1; a++; } return a; }     // utilities.h:286-291

    // pbbs::sequence<T>: owning array with size()/operator[]/begin()/end() (pbbslib/seq.h)
    template <class T>
    class sequence
    {
    public:
        sequence() = default;
        explicit sequence(size_t n) : data_(n) {}
        size_t size() const { return data_.size(); }
        T& operator[](size_t i) { return data_[i]; }
        const T& operator[](size_t i) const { return data_[i]; }
        T* begin() { return data_.data(); }
        T* end() { return data_.data() + data_.size(); }
        const T* begin() const { return data_.data(); }
        const T* end() const { return data_.data() + data_.size(); }
        void clear() { data_.clear(); }
    private:
        std::vector<T> data_;
    };
}

namespace pbbs
{
    // pbbslib/seq.h delayed_sequence: element i is f(i), computed on access
    template <class T, class F>
    struct delayed_sequence
    {
        size_t n;
        F f;
        size_t size() const { return n; }
        T operator[](size_t i) const { return f(i); }
    };
    template <class T, class F> delayed_sequence<T, F> delayed_seq(size_t n, F f) { return delayed_sequence<T, F>{n, std::move(f)}; }
}

// pbbslib/parallel.h parallel_for: host-side loops of the callers are not the hot path here - they run in order
template <class F>
inline void parallel_for(size_t start, size_t end, F f, long /*granularity*/ = 0, bool /*conservative*/ = false) { for (size_t i = start; i < end; i++) f(i); }

// pbbslib/get_time.h:9-81
struct timer
{
    double total_time = 0;
    std::chrono::steady_clock::time_point last;
    bool on = false;
    std::string name;
    explicit timer(std::string n = "WharfMH time", bool start_now = true) : name(std::move(n)) { if (start_now) start(); }
    void start() { on = true; last = std::chrono::steady_clock::now(); }
    double stop() { on = false; double d = std::chrono::duration<double>(std::chrono::steady_clock::now() - last).count(); total_time += d; return d; }
    void reset() { total_time = 0; on = false; }
    double get_total() const { return on ? total_time + std::chrono::duration<double>(std::chrono::steady_clock::now() - last).count() : total_time; }
    void reportTotal(const std::string& what) const { std::cout << name << ": " << what << ": " << get_total() << std::endl; }
};

inline int num_workers() { return 1; }   // host threads driving the engine; the parallelism is on the device

namespace types   // config/types.h:4-41
{
    using Vertex   = uintV;
    using Degree   = uintE;
    using WalkID   = uint64_t;                // widened: the reference's uint32_t wraps for n > 5,368,709
    using Position = uint8_t;
    using State    = std::pair<Vertex, Vertex>;
    enum RandomWalkModelType { DEEPWALK, NODE2VEC };
    enum SamplerInitStartegy { RANDOM, BURNIN, WEIGHT };
}

namespace config   // config/globals.h:7-29
{
    inline uint8_t walks_per_vertex   = 10;
    inline uint8_t walk_length        = 80;
    inline auto random_walk_model     = types::RandomWalkModelType::NODE2VEC;
    inline float paramP               = 4.0;
    inline float paramQ               = 1.0;
    inline auto sampler_init_strategy = types::SamplerInitStartegy::WEIGHT;
    inline uint64_t random_seed       = static_cast<uint64_t>(std::time(nullptr));   // config::random = Random(time(nullptr))
    inline auto deterministic_mode    = true;
    // additions of this engine
    inline int device                 = 0;
    inline int search_mode            = WMH_SEARCH_AUTO;
    inline int generate_mode          = WMH_GENERATE_AUTO;      // schedule of deterministic bulk generation (same corpus)
    inline uint32_t rank              = 0;                      // one process per GPU: this process's share of the corpus
    inline uint32_t world             = 1;                      // (whole replicas first, the rest by start vertex; wharfmh_b200.h)
}

namespace dynamic_graph_representation_learning_with_metropolis_hastings
{
    namespace detail
    {
        inline wmh_engine*& current_engine() { static wmh_engine* e = nullptr; return e; }
        inline void check(int rc) { if (rc != WMH_OK) throw std::runtime_error(std::string("wharfmh_b200: ") + wmh_last_error()); }
        inline wmh_config make_config()
        {
            wmh_config c;
            wmh_default_config(&c);
            c.walks_per_vertex = config::walks_per_vertex;
            c.walk_length = config::walk_length;
            c.model = config::random_walk_model == types::DEEPWALK ? WMH_DEEPWALK : WMH_NODE2VEC;
            c.deterministic = config::deterministic_mode ? 1 : 0;
            c.param_p = config::paramP;
            c.param_q = config::paramQ;
            c.init_strategy = config::sampler_init_strategy == types::RANDOM ? WMH_INIT_RANDOM
                            : config::sampler_init_strategy == types::BURNIN ? WMH_INIT_BURNIN : WMH_INIT_WEIGHT;
            c.search_mode = static_cast<uint32_t>(config::search_mode);
            c.generate_mode = static_cast<uint32_t>(config::generate_mode);
            c.seed = config::random_seed;
            c.device = config::device;
            c.rank = config::rank;
            c.world = config::world;
            return c;
        }
    }

    // graph/snapshot.h:108-194: O(1) access to every vertex's ascending neighbour array
    struct FlatGraphEntry
    {
        const types::Vertex* neighbors;
        types::Degree degree;
    };

    class FlatGraph
    {
    public:
        FlatGraph(std::vector<uintE> row_ptr, std::vector<uintV> col) : row_ptr_(std::move(row_ptr)), col_(std::move(col)) {}
        FlatGraphEntry operator[](types::Vertex v) const { return FlatGraphEntry{col_.data() + row_ptr_[v], row_ptr_[v + 1] - row_ptr_[v]}; }
        types::Vertex size() const { return static_cast<types::Vertex>(row_ptr_.size() - 1); }
        size_t size_in_bytes() const { return row_ptr_.size() * sizeof(uintE) + col_.size() * sizeof(uintV); }
        std::tuple<const types::Vertex*, types::Degree, bool> neighbors(types::Vertex v) const { auto e = (*this)[v]; return {e.neighbors, e.degree, false}; }
    private:
        std::vector<uintE> row_ptr_;
        std::vector<uintV> col_;
    };

    // graph/wharfmh.h:140 (flatten_vertex_tree) / graph/vertex.h: what the reference's tests read of a vertex entry -
    // its neighbours, the size of its inverted index and of its sampler manager (tests/wharfmh.cpp:65-99,114-118,134-139).
    struct VertexEntryView
    {
        struct Edges                                                     // tree_plus edge list of the vertex
        {
            const uintV* row = nullptr;
            size_t count = 0;
            size_t degree() const { return count; }
            size_t size() const { return count; }
            uintV* get_edges(uintV /*source*/) const                     // a fresh array the caller may free with pbbs::free_array
            {
                uintV* out = pbbs::new_array_no_init<uintV>(count);
                if (count) std::memcpy(out, row, count * sizeof(uintV));
                return out;
            }
        } compressed_edges;
        struct Index                                                     // walks/inverted_index.h
        {
            const void* root = nullptr;                                  // nullptr <=> empty, as the PAM map's root
            size_t entries = 0;
            size_t size() const { return entries; }
        } inverted_index;
        struct Samplers                                                  // graph/vertex.h:12 SamplerManager
        {
            bool untouched = true;
            // Exact while no Metropolis-Hastings walk has run on the engine (deterministic mode never touches a sampler).
            // Afterwards the states live in the aux words of the device rows, keyed by edge, and are not counted per vertex.
            size_t size() const
            {
                if (!untouched) throw std::logic_error("wharfmh_b200: sampler states are kept per edge on the device, not per vertex");
                return 0;
            }
        } samplers;
        const Samplers* sampler_manager = nullptr;
    };

    class FlatVertexTree
    {
    public:
        FlatVertexTree() = default;
        FlatVertexTree(std::vector<uintE> row_ptr, std::vector<uintV> col, const std::vector<uint64_t>& index_sizes, bool samplers_untouched)
            : row_ptr_(std::move(row_ptr)), col_(std::move(col)), entries_(index_sizes.size())
        {
            for (size_t v = 0; v < entries_.size(); v++)
            {
                auto& e = entries_[v];
                e.compressed_edges.row = col_.data() + row_ptr_[v];
                e.compressed_edges.count = row_ptr_[v + 1] - row_ptr_[v];
                e.inverted_index.entries = index_sizes[v];
                e.inverted_index.root = index_sizes[v] ? static_cast<const void*>(&e) : nullptr;
                e.samplers.untouched = samplers_untouched;
            }
            for (auto& e : entries_) e.sampler_manager = &e.samplers;
        }
        FlatVertexTree(const FlatVertexTree&) = delete;
        FlatVertexTree& operator=(const FlatVertexTree&) = delete;
        FlatVertexTree(FlatVertexTree&&) = default;                      // vectors keep their buffers: the views stay valid
        size_t size() const { return entries_.size(); }
        const VertexEntryView& operator[](size_t v) const { return entries_[v]; }
        size_t size_in_bytes() const { return row_ptr_.size() * sizeof(uintE) + col_.size() * sizeof(uintV) + entries_.size() * sizeof(VertexEntryView); }
    private:
        std::vector<uintE> row_ptr_;
        std::vector<uintV> col_;
        std::vector<VertexEntryView> entries_;
    };

    // graph/wharfmh.h:21
    class WharfMH
    {
    public:
        // graph/wharfmh.h:26
        WharfMH(long graph_vertices, long /*graph_edges*/)
        {
            wmh_config c = detail::make_config();
            detail::check(wmh_create_empty(&engine_, static_cast<uint64_t>(graph_vertices), &c));
            detail::current_engine() = engine_;
        }

        // graph/wharfmh.h:58: n offsets (no n+1-th), rows sorted and unique; takes ownership when free_memory
        WharfMH(long graph_vertices, long graph_edges, uintE* offsets, uintV* edges, bool free_memory = true)
        {
            wmh_config c = detail::make_config();
            int rc = wmh_create(&engine_, static_cast<uint64_t>(graph_vertices), static_cast<uint64_t>(graph_edges), offsets, edges, &c);
            if (free_memory) { pbbs::free_array(offsets); pbbs::free_array(edges); }      // :98-103
            detail::check(rc);
            detail::current_engine() = engine_;
        }

        WharfMH(const WharfMH&) = delete;
        WharfMH& operator=(const WharfMH&) = delete;
        WharfMH(WharfMH&& o) noexcept : engine_(o.engine_), mh_walked_(o.mh_walked_) { o.engine_ = nullptr; }
        ~WharfMH() { destroy(); }

        [[nodiscard]] size_t number_of_vertices() const { return wmh_number_of_vertices(engine_); }   // :117
        [[nodiscard]] size_t number_of_edges() const { return wmh_number_of_edges(engine_); }         // :130

        [[nodiscard]] FlatGraph flatten_graph() const                                                  // :175
        {
            std::vector<uintE> row_ptr(number_of_vertices() + 1);
            std::vector<uintV> col(number_of_edges());
            detail::check(wmh_flatten_graph(engine_, row_ptr.data(), col.data()));
            return FlatGraph(std::move(row_ptr), std::move(col));
        }

        [[nodiscard]] FlatVertexTree flatten_vertex_tree() const                                       // :140
        {
            if (!engine_) return FlatVertexTree();                                                      // destroyed: no vertices (tests/wharfmh.cpp:114-117)
            const size_t n = number_of_vertices();
            std::vector<uintE> row_ptr(n + 1);
            std::vector<uintV> col(number_of_edges());
            std::vector<uint64_t> index_sizes(n);
            detail::check(wmh_flatten_graph(engine_, row_ptr.data(), col.data()));
            detail::check(wmh_index_sizes(engine_, index_sizes.data()));
            return FlatVertexTree(std::move(row_ptr), std::move(col), index_sizes, !mh_walked_);
        }

        // :219.  map_function(entry, i) with entry.first = the vertex and entry.second = its entry, in vertex order.
        template <class Function>
        void map_vertices(Function map_function, bool /*run_seq*/ = false, size_t /*granularity*/ = 0) const
        {
            const FlatVertexTree flat = flatten_vertex_tree();
            for (size_t v = 0; v < flat.size(); v++) map_function(std::pair<types::Vertex, const VertexEntryView&>(static_cast<types::Vertex>(v), flat[v]), v);
        }

        void destroy()                                                                                  // :228
        {
            if (!engine_) return;
            if (detail::current_engine() == engine_) detail::current_engine() = nullptr;
            wmh_destroy(engine_);
            engine_ = nullptr;
        }

        void destroy_index() { detail::check(wmh_destroy_index(engine_)); }                             // :237

        void generate_initial_random_walks()                                                            // :250
        {
            sync_config();
            detail::check(wmh_generate_initial_random_walks(engine_));
        }

        // Extension: the same pass with the corpus streamed to `out` ([local walks][walk_length]) while it is generated.
        void generate_initial_random_walks(types::Vertex* out, size_t capacity_walks)
        {
            sync_config();
            detail::check(wmh_generate_initial_random_walks_to_host(engine_, out, capacity_walks));
        }

        std::string walk(types::WalkID walk_id)                                                         // :365
        {
            size_t need = 0;
            detail::check(wmh_walk_string(engine_, walk_id, nullptr, 0, &need));
            std::string s(need, '\0');
            detail::check(wmh_walk_string(engine_, walk_id, s.data(), need, &need));
            s.resize(need - 1);
            return s;
        }

        types::Vertex vertex_at_walk(types::WalkID walk_id, types::Position position)                   // :404
        {
            uint32_t v = 0;
            detail::check(wmh_vertex_at_walk(engine_, walk_id, position, &v));
            return v;
        }

        // :439.  Sorts the caller's array in place when !sorted, never frees it; returns the affected walk ids.
        pbbs::sequence<types::WalkID> insert_edges_batch(size_t m, std::tuple<uintV, uintV>* edges, bool sorted = false, bool remove_dups = false,
                                                         size_t nn = std::numeric_limits<size_t>::max(), bool apply_walk_updates = true, bool run_seq = false)
        { return batch(false, m, edges, sorted, remove_dups, nn, apply_walk_updates, run_seq); }

        // :588
        pbbs::sequence<types::WalkID> delete_edges_batch(size_t m, std::tuple<uintV, uintV>* edges, bool sorted = false, bool remove_dups = false,
                                                         size_t nn = std::numeric_limits<size_t>::max(), bool apply_walk_updates = true, bool run_seq = false)
        { return batch(true, m, edges, sorted, remove_dups, nn, apply_walk_updates, run_seq); }

        // experiments/src/vertex-classification.cpp:167-178: walk(id) + newline for every id, gathered on the device in bulk
        // and written to `path` (the walks.txt the downstream trainer reads).  Returns the bytes written.
        size_t write_walks(const char* path, const pbbs::sequence<types::WalkID>& walk_ids, bool append = false)
        {
            uint64_t written = 0;
            detail::check(wmh_write_walks(engine_, walk_ids.begin(), walk_ids.size(), path, append ? 1 : 0, &written));
            return written;
        }

        // :928 (device memory instead of allocator pools)
        void memory_footprint() const
        {
            wmh_memory mem{};
            if (engine_) detail::check(wmh_memory_footprint(engine_, &mem));      // after destroy(): an empty graph, as the reference prints (tests/wharfmh.cpp:103-106)
            std::cout << "Graph: \n\t" << "Vertices: " << number_of_vertices() << ", Edges: " << number_of_edges() << std::endl;
            std::cout << "Device memory: \n\tgraph = " << mem.graph_bytes << " B (pool " << mem.graph_pool_bytes << " B), walks = " << mem.walks_bytes
                      << " B, index = " << mem.index_bytes << " B, scratch = " << mem.scratch_bytes << " B, total = " << mem.total_bytes << " B" << std::endl;
        }
        void print_memory_pool_stats() const { memory_footprint(); }                                    // :1003

        wmh_engine* handle() const { return engine_; }

        // -- several processes, one GPU each (no counterpart in the single-process reference) ----------------------
        // Global ids of the walks this process owns, ascending; the rank that owns a walk.
        std::vector<types::WalkID> local_walk_ids() const
        {
            std::vector<types::WalkID> ids(wmh_local_walk_ids(engine_, nullptr, 0));
            if (!ids.empty()) wmh_local_walk_ids(engine_, ids.data(), ids.size());
            return ids;
        }
        uint32_t walk_owner(types::WalkID walk_id) const { return wmh_walk_owner(engine_, walk_id); }
        // Peer memory between the processes of one box (wharfmh_b200.h, wmh_peer_*): create the exchange window, hand the
        // returned handle to every other process (MPI_Allgather, a file ...), then attach all of them in rank order.
        std::array<unsigned char, WMH_PEER_HANDLE_BYTES> peer_create(uint64_t bytes = 0)
        {
            std::array<unsigned char, WMH_PEER_HANDLE_BYTES> handle{};
            if (bytes == 0) bytes = wmh_peer_bytes_needed(engine_) * 5 / 4 + (1u << 20);
            detail::check(wmh_peer_create(engine_, bytes, handle.data()));
            return handle;
        }
        void peer_attach(const std::vector<std::array<unsigned char, WMH_PEER_HANDLE_BYTES>>& handles)
        {
            std::vector<unsigned char> blob;
            for (const auto& h : handles) blob.insert(blob.end(), h.begin(), h.end());
            detail::check(wmh_peer_attach(engine_, blob.data()));
        }
        void peer_detach() { detail::check(wmh_peer_detach(engine_)); }

    private:
        wmh_engine* engine_ = nullptr;
        bool mh_walked_ = false;             // a generate or batch ran with deterministic_mode off: sampler states may exist

        void sync_config()
        {
            wmh_config cur, want = detail::make_config();
            detail::check(wmh_get_config(engine_, &cur));
            want.rank = cur.rank; want.world = cur.world; want.device = cur.device;
            detail::check(wmh_set_config(engine_, &want));
            if (!config::deterministic_mode) mh_walked_ = true;      // called by every member that can walk
        }

        pbbs::sequence<types::WalkID> batch(bool is_delete, size_t m, std::tuple<uintV, uintV>* edges, bool sorted, bool remove_dups, size_t nn,
                                            bool apply_walk_updates, bool /*run_seq*/)
        {
            sync_config();
            std::vector<uint32_t> pairs(2 * m);                  // std::tuple's member order is implementation-defined: repack
            for (size_t i = 0; i < m; i++) { pairs[2 * i] = std::get<0>(edges[i]); pairs[2 * i + 1] = std::get<1>(edges[i]); }
            uint64_t affected = 0;
            detail::check((is_delete ? wmh_delete_edges_batch : wmh_insert_edges_batch)(engine_, m, pairs.data(), sorted, remove_dups, nn,
                                                                                      apply_walk_updates, &affected));
            if (!sorted) for (size_t i = 0; i < m; i++) edges[i] = std::make_tuple(pairs[2 * i], pairs[2 * i + 1]);   // in-place sort (:452)
            pbbs::sequence<types::WalkID> ids(affected);
            if (affected) detail::check(wmh_affected_walks(engine_, ids.begin(), nullptr, affected));
            return ids;
        }
    };
}

// ---- the four global timers (config/config.h:10-14): read the engine's CUDA-event accumulators ----------
struct wharf_engine_timer
{
    int which;
    double get_total() const { auto* e = dygrl::detail::current_engine(); return e ? wmh_timer_get(e, which) : 0.0; }
    void reset() const { auto* e = dygrl::detail::current_engine(); if (e) wmh_timer_reset(e, which); }
    void start() const {}
    void stop() const {}
};
inline const wharf_engine_timer graph_update_time_on_insert{WMH_TIMER_GRAPH_UPDATE_INSERT};
inline const wharf_engine_timer walk_update_time_on_insert{WMH_TIMER_WALK_UPDATE_INSERT};
inline const wharf_engine_timer graph_update_time_on_delete{WMH_TIMER_GRAPH_UPDATE_DELETE};
inline const wharf_engine_timer walk_update_time_on_delete{WMH_TIMER_WALK_UPDATE_DELETE};

namespace utility   // utils/utility.h
{
    inline double GB(size_t bytes) { return bytes / 1024.0 / 1024.0 / 1024.0; }     // :15-24
    inline double MB(size_t bytes) { return bytes / 1024.0 / 1024.0; }              // :33-41

    // :55-146.  Same edges, order and count as the reference's generator; produced on the device.
    inline std::pair<std::tuple<uintV, uintV>*, size_t> generate_batch_of_edges(size_t edges_number, size_t vertices_number, size_t batch_seed,
                                                                                bool self_loops = false, bool directed = true, double a = 0.5,
                                                                                double b = 0.2, double c = 0.1, bool /*run_seq*/ = false);
}

namespace detail
{
    inline std::string read_file(const char* fname)
    {
        std::ifstream in(fname, std::ios::in | std::ios::binary);
        if (!in) throw std::runtime_error(std::string("Unable to open file: ") + fname);      // IO.h:54-57 aborts here
        std::ostringstream all;
        all << in.rdbuf();
        return all.str();
    }
}

inline std::pair<std::tuple<uintV, uintV>*, size_t> utility::generate_batch_of_edges(size_t edges_number, size_t vertices_number, size_t batch_seed, bool self_loops,
                                                                                     bool directed, double a, double b, double c, bool /*run_seq*/)
{
    const size_t slots = edges_number * (directed ? 1 : 2);
    std::vector<uint32_t> pairs(2 * slots);
    uint64_t kept = 0;
    dygrl::detail::check(wmh_generate_batch_of_edges(config::device, edges_number, vertices_number, batch_seed, self_loops, directed, a, b, c, pairs.data(), slots, &kept));
    auto* edges = pbbs::new_array_no_init<std::tuple<uintV, uintV>>(kept);      // the caller frees it with pbbs::free_array (throughput-latency.cpp:146)
    for (size_t i = 0; i < kept; i++) new (edges + i) std::tuple<uintV, uintV>(pairs[2 * i], pairs[2 * i + 1]);
    return {edges, static_cast<size_t>(kept)};
}

// libs/compressed_trees/common/IO.h:67-106: Ligra "AdjacencyGraph" text format -> (n, m, offsets[n], edges[m])
inline std::tuple<size_t, size_t, uintE*, uintV*> read_unweighted_graph(const char* fname, bool /*is_symmetric*/, bool /*mmap*/ = false)
{
    const std::string text = detail::read_file(fname);
    uint64_t n = 0, m = 0;
    dygrl::detail::check(wmh_parse_adjacency_graph(text.data(), text.size(), &n, &m, nullptr, nullptr));
    std::cout << "Vertices: " << n << " Edges: " << m << std::endl;
    uintE* offsets = pbbs::new_array_no_init<uintE>(n);
    uintV* edges = pbbs::new_array_no_init<uintV>(m);
    const int rc = wmh_parse_adjacency_graph(text.data(), text.size(), &n, &m, offsets, edges);
    if (rc != WMH_OK)
    {
        pbbs::free_array(offsets);
        pbbs::free_array(edges);
        dygrl::detail::check(rc);
    }
    return {n, m, offsets, edges};
}

// experiments/src/vertex-classification.cpp:5-38: the lines "u v" of a file in partitions of edge_partition_size lines, every
// line as (u, v) and (v, u); each partition is a heap array the caller passes to insert_edges_batch(size, array, false, true).
// (The driver defines create_edge_stream(commandLine&, stream) itself - -f and -eps options - so that signature is left to it.)
inline void create_edge_stream(const char* fname, size_t edge_partition_size, std::vector<std::pair<std::tuple<uintV, uintV>*, uintV>>& stream)
{
    if (edge_partition_size == 0) throw std::invalid_argument("edge partition size must be positive");
    const std::string text = detail::read_file(fname);
    uint64_t lines = 0, vertices = 0;
    dygrl::detail::check(wmh_parse_edge_list(text.data(), text.size(), nullptr, 0, &lines, &vertices));
    std::vector<uint32_t> pairs(2 * lines);
    dygrl::detail::check(wmh_parse_edge_list(text.data(), text.size(), pairs.data(), lines, &lines, &vertices));
    for (size_t first = 0; first < lines; first += edge_partition_size)
    {
        const size_t k = std::min<size_t>(edge_partition_size, lines - first);
        auto* part = new std::tuple<uintV, uintV>[2 * k];
        for (size_t j = 0; j < k; j++)
        {
            const uintV u = pairs[2 * (first + j)], v = pairs[2 * (first + j) + 1];
            part[2 * j] = std::make_tuple(u, v);
            part[2 * j + 1] = std::make_tuple(v, u);
        }
        stream.emplace_back(part, static_cast<uintV>(2 * k));
    }
}

// libs/compressed_trees/pbbslib/parse_command_line.h:30-80
struct commandLine
{
    int argc;
    char** argv;
    std::string comLine;
    commandLine(int c, char** v, std::string s = "") : argc(c), argv(v), comLine(std::move(s)) {}
    bool getOption(const std::string& option) const { for (int i = 1; i < argc; i++) if (option == argv[i]) return true; return false; }
    char* getOptionValue(const std::string& option) const { for (int i = 1; i < argc - 1; i++) if (option == argv[i]) return argv[i + 1]; return nullptr; }
    std::string getOptionValue(const std::string& option, const std::string& def) const { char* v = getOptionValue(option); return v ? std::string(v) : def; }
    long getOptionLongValue(const std::string& option, long def) const { char* v = getOptionValue(option); return v ? std::atol(v) : def; }
    double getOptionDoubleValue(const std::string& option, double def) const { char* v = getOptionValue(option); return v ? std::atof(v) : def; }
};

#endif  // WHARFMH_B200_HPP
