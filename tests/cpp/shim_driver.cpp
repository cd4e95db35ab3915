// Drives the C++ re-creation of dygrl::WharfMH (include/wharfmh_b200.hpp) the way the reference's own tests drive the
// original (tests/sampler.cpp:24-47, tests/wharfmh.cpp): the 6-vertex graph of tests/sampler.cpp, 2 walks per
// vertex of length 8, deterministic mode; prints every walk, inserts {(0,3),(4,0)}, prints, deletes them, prints.
// tests/test_cpp_shim.py compares the output with the corpora of the unmodified reference (tests/golden/golden.json).
#include "wharfmh_b200.hpp"

#include <cstdio>

int main()
{
    // rows: 0:{1,2} 1:{0,2,3} 2:{0,1,3,4,5} 3:{1,2,5} 4:{2,5} 5:{2,3,4}
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

    dygrl::WharfMH wharf(n, m, offsets, edges);            // takes ownership (free_memory = true), wharfmh.h:58
    wharf.generate_initial_random_walks();
    auto dump = [&](const char* tag) {
        for (types::WalkID w = 0; w < (types::WalkID) n * config::walks_per_vertex; w++) std::printf("%s %llu: %s\n", tag, (unsigned long long) w, wharf.walk(w).c_str());
    };
    dump("initial");
    std::printf("vertex_at_walk(2,3) = %u\n", (unsigned) wharf.vertex_at_walk(2, 3));

    auto* batch = pbbs::new_array_no_init<std::tuple<uintV, uintV>>(2);
    batch[0] = std::make_tuple(4u, 0u);                      // unsorted on purpose: the call sorts it in place (wharfmh.h:452)
    batch[1] = std::make_tuple(0u, 3u);
    auto affected = wharf.insert_edges_batch(2, batch, false, true, 8);
    std::printf("insert affected %zu first (%u,%u) edges %zu\n", affected.size(), std::get<0>(batch[0]), std::get<1>(batch[0]), wharf.number_of_edges());
    dump("inserted");
    affected = wharf.delete_edges_batch(2, batch, true, true, 8);
    std::printf("delete affected %zu edges %zu\n", affected.size(), wharf.number_of_edges());
    dump("deleted");
    pbbs::free_array(batch);

    auto flat = wharf.flatten_graph();
    std::printf("degree(2) = %zu\n", (size_t) flat[2].degree);
    wharf.destroy();
    return 0;
}
