// The snapshot views and the walks.txt hand-off of the C++ re-creation (include/wharfmh_b200.hpp), driven the way the
// reference's tests and drivers use the originals: flatten_vertex_tree / map_vertices (graph/wharfmh.h:140,219, read as in
// tests/wharfmh.cpp:65-99,114-139) and walk(id) + endl per affected walk (experiments/src/vertex-classification.cpp:167-178).
// Same 6-vertex graph as tests/cpp/shim_driver.cpp (tests/sampler.cpp:24-35); argv[1] = scratch directory.
#include "wharfmh_b200.hpp"

#include <cstdio>
#include <fstream>
#include <string>

int main(int argc, char** argv)
{
    if (argc < 2) return 2;
    const uintV adj[] = {1, 2, 0, 2, 3, 0, 1, 3, 4, 5, 1, 2, 5, 2, 5, 2, 3, 4};
    const uintE off[] = {0, 2, 5, 10, 13, 15};
    const long n = 6, m = 18;
    uintE* offsets = pbbs::new_array_no_init<uintE>(n);
    uintV* edges = pbbs::new_array_no_init<uintV>(m);
    for (long i = 0; i < n; i++) offsets[i] = off[i];
    for (long i = 0; i < m; i++) edges[i] = adj[i];
    config::walks_per_vertex = 2;
    config::walk_length = 8;
    config::random_walk_model = types::DEEPWALK;
    config::deterministic_mode = true;

    dygrl::WharfMH wharf(n, m, offsets, edges);
    {
        auto tree = wharf.flatten_vertex_tree();                // before any walk: every index empty (tests/wharfmh.cpp:76-77)
        size_t entries = 0, roots = 0;
        for (size_t v = 0; v < tree.size(); v++) { entries += tree[v].inverted_index.size(); roots += tree[v].inverted_index.root != nullptr; }
        std::printf("fresh tree %zu vertices %zu index entries %zu roots degree(2) %zu samplers %zu\n", tree.size(), entries, roots,
                    tree[2].compressed_edges.degree(), tree[2].sampler_manager->size());
    }
    wharf.generate_initial_random_walks();
    {
        auto tree = wharf.flatten_vertex_tree();                // wharfmh.h:140, read the way tests/wharfmh.cpp:65-99 reads it
        size_t entries = 0, mapped = 0;
        for (size_t v = 0; v < tree.size(); v++) entries += tree[v].inverted_index.size();
        wharf.map_vertices([&](const auto& entry, size_t i) { mapped += entry.first == i && entry.second.compressed_edges.degree() == tree[i].compressed_edges.size(); });
        uintV* row = tree[2].compressed_edges.get_edges(2);
        std::printf("vertex tree %zu vertices %zu index entries, vertex 2: degree %zu first %u last %u index %zu root %d samplers %zu mapped %zu\n", tree.size(), entries,
                    tree[2].compressed_edges.degree(), row[0], row[4], tree[2].inverted_index.size(), tree[2].inverted_index.root != nullptr,
                    tree[2].sampler_manager->size(), mapped);
        pbbs::free_array(row);
    }

    auto* batch = pbbs::new_array_no_init<std::tuple<uintV, uintV>>(2);
    batch[0] = std::make_tuple(4u, 0u);
    batch[1] = std::make_tuple(0u, 3u);
    auto affected = wharf.insert_edges_batch(2, batch, false, true, 8);
    pbbs::free_array(batch);
    {
        const std::string walks_path = std::string(argv[1]) + "/walks.txt";
        const size_t bytes = wharf.write_walks(walks_path.c_str(), affected);                // vertex-classification.cpp:167-178
        std::ifstream f(walks_path);
        std::string line;
        size_t lines = 0;
        bool same = true;
        while (std::getline(f, line)) same = same && line == wharf.walk(affected[lines++]);
        std::printf("walks.txt %zu bytes %zu lines same %d\n", bytes, lines, (int) same);
    }
    wharf.destroy_index();
    {
        auto tree = wharf.flatten_vertex_tree();                // after destroy_index: empty again (tests/wharfmh.cpp:134-139)
        size_t entries = 0;
        for (size_t v = 0; v < tree.size(); v++) entries += tree[v].inverted_index.size();
        std::printf("after destroy_index %zu vertices %zu index entries degree(0) %zu\n", tree.size(), entries, tree[0].compressed_edges.degree());
    }
    wharf.destroy();
    std::printf("destroyed %zu vertices tree %zu\n", wharf.number_of_vertices(), wharf.flatten_vertex_tree().size());   // tests/wharfmh.cpp:108-117
    return 0;
}
