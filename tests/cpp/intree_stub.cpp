// INTEGRATION.md section 2, compile-checked against the REAL reference: one translation unit that includes the reference's
// own graph/wharfmh.h (from /root/reference, unchanged) AND the C ABI, builds the reference's WharfMH on the host, and hands
// its graph to the B200 engine exactly as the b200_attach() stub of the document does.  Both engines in one binary: on a
// machine without a GPU the reference part runs and wmh_create refuses loudly; on the B200 both corpora are printed
// (deterministic mode: they must be equal line by line).
#include <wharfmh.h>          // the reference (include path: /root/reference/graph etc.)
#include <wharfmh_b200.h>     // this repository's C ABI

static wmh_engine* b200_attach(const dygrl::WharfMH& reference)
{
    auto flat = reference.flatten_graph();                                           // graph/wharfmh.h:175
    std::vector<uint64_t> off(flat.size());
    std::vector<uint32_t> adj;
    uint64_t m = 0;
    for (size_t v = 0; v < flat.size(); v++)
    {
        off[v] = m;
        adj.insert(adj.end(), flat[v].neighbors, flat[v].neighbors + flat[v].degree);
        m += flat[v].degree;
    }
    wmh_config c;
    wmh_default_config(&c);
    c.walks_per_vertex = config::walks_per_vertex;
    c.walk_length = config::walk_length;
    c.model = config::random_walk_model == types::DEEPWALK ? WMH_DEEPWALK : WMH_NODE2VEC;
    c.param_p = config::paramP;
    c.param_q = config::paramQ;
    c.deterministic = config::deterministic_mode;
    c.init_strategy = config::sampler_init_strategy;                                 // same enum order (config/types.h:34)
    wmh_engine* engine = nullptr;
    if (wmh_create(&engine, flat.size(), m, off.data(), adj.data(), &c) != WMH_OK)
    {
        std::cerr << "b200: " << wmh_last_error() << std::endl;
        return nullptr;
    }
    return engine;
}

static std::string b200_walk(wmh_engine* engine, size_t w)
{
    size_t need = 0;
    wmh_walk_string(engine, w, nullptr, 0, &need);
    std::string s(need, '\0');
    wmh_walk_string(engine, w, s.data(), need, &need);
    s.resize(need - 1);
    return s;
}

// Every walk of both engines, side by side.  Returns the number of walks whose text differs.
static size_t compare(const char* phase, dygrl::WharfMH& reference, wmh_engine* engine, size_t walks, bool print)
{
    size_t differing = 0;
    for (size_t w = 0; w < walks; w++)
    {
        const std::string a = reference.walk(w), b = b200_walk(engine, w);
        if (print) std::cout << phase << " " << w << ": reference " << a << "| b200 " << b << std::endl;
        differing += a != b;
    }
    std::cout << phase << ": " << walks << " walks, differing " << differing << ", edges " << reference.number_of_edges() << " / "
              << wmh_number_of_edges(engine) << std::endl;
    return differing;
}

// Both engines take the same batch the way the reference's drivers pass it (sorted = false, remove_dups = true, run_seq = true:
// the reference's map of changes is racy otherwise, graph/wharfmh.h:524-536).
static size_t batch_on_both(const char* phase, bool is_delete, dygrl::WharfMH& reference, wmh_engine* engine, const std::vector<std::tuple<uintV, uintV>>& batch,
                            size_t walks, bool print)
{
    auto* for_reference = pbbs::new_array_no_init<std::tuple<uintV, uintV>>(batch.size());
    std::vector<uint32_t> pairs(2 * batch.size());
    for (size_t i = 0; i < batch.size(); i++)
    {
        for_reference[i] = batch[i];
        pairs[2 * i] = std::get<0>(batch[i]);
        pairs[2 * i + 1] = std::get<1>(batch[i]);
    }
    auto affected = is_delete ? reference.delete_edges_batch(batch.size(), for_reference, false, true, std::numeric_limits<size_t>::max(), true, true)
                              : reference.insert_edges_batch(batch.size(), for_reference, false, true, std::numeric_limits<size_t>::max(), true, true);
    pbbs::free_array(for_reference);
    uint64_t k = 0;
    const int rc = (is_delete ? wmh_delete_edges_batch : wmh_insert_edges_batch)(engine, batch.size(), pairs.data(), 0, 1, ~0ull, 1, &k);
    if (rc != WMH_OK) { std::cerr << "b200: " << wmh_last_error() << std::endl; return 1; }
    std::cout << phase << ": affected " << affected.size() << " / " << k << std::endl;
    return (affected.size() != k) + compare(phase, reference, engine, walks, print);
}

int main(int argc, char** argv)
{
    size_t n, m;
    uintE* offsets;
    uintV* edges;
    std::vector<std::tuple<uintV, uintV>> batch;
    const bool small = argc < 2;
    if (small)
    {
        // the 6-vertex graph of tests/sampler.cpp:24-35, 2 walks per vertex of length 8, batch {(0,3),(4,0)}
        const uintV adj[] = {1, 2, 0, 2, 3, 0, 1, 3, 4, 5, 1, 2, 5, 2, 5, 2, 3, 4};
        const uintE off[] = {0, 2, 5, 10, 13, 15};
        n = 6;
        m = 18;
        offsets = pbbs::new_array_no_init<uintE>(n);
        edges = pbbs::new_array_no_init<uintV>(m);
        for (size_t i = 0; i < n; i++) offsets[i] = off[i];
        for (size_t i = 0; i < m; i++) edges[i] = adj[i];
        config::walks_per_vertex = 2;
        config::walk_length = 8;
        batch = {std::make_tuple(4u, 0u), std::make_tuple(0u, 3u)};
    }
    else
    {
        // an AdjacencyGraph file (the reference's own reader), 10 x 80, the reference's generator for the batch
        std::tie(n, m, offsets, edges) = read_unweighted_graph(argv[1], true, false);
        auto generated = utility::generate_batch_of_edges(argc > 2 ? std::atol(argv[2]) : 100, n, 0, false, false);
        batch.assign(generated.first, generated.first + generated.second);
        pbbs::free_array(generated.first);
    }
    config::deterministic_mode = true;
    const size_t walks = n * config::walks_per_vertex;

    dygrl::WharfMH reference(n, m, offsets, edges);
    reference.generate_initial_random_walks();
    if (small) for (size_t w = 0; w < walks; w++) std::cout << "reference " << w << ": " << reference.walk(w) << std::endl;

    wmh_engine* engine = b200_attach(reference);
    if (!engine) return 3;
    if (wmh_generate_initial_random_walks(engine) != WMH_OK) { std::cerr << "b200: " << wmh_last_error() << std::endl; return 4; }
    size_t bad = compare("initial", reference, engine, walks, small);
    bad += batch_on_both("inserted", false, reference, engine, batch, walks, small);
    bad += batch_on_both("deleted", true, reference, engine, batch, walks, small);
    wmh_destroy(engine);
    std::cout << "total differing: " << bad << std::endl;
    return bad ? 5 : 0;
}
