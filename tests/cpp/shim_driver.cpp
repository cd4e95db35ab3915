// Drives the C++ re-creation of dygrl::WharfMH (include/wharfmh_b200.hpp) the way the reference's own tests drive the
// original (tests/sampler.cpp:24-47, tests/wharfmh.cpp): the 6-vertex graph of tests/sampler.cpp, 2 walks per
// vertex of length 8, deterministic mode; prints every walk, inserts {(0,3),(4,0)}, prints, deletes them, prints.
// tests/test_cpp_shim.py compares the output with the corpora of the unmodified reference (tests/golden/golden.json).
#include "wharfmh_b200.hpp"

#include <cstdio>
#include <fstream>
#include <string>

// Host-only part (runs without a GPU): the text formats either side of the path.  argv[1] = scratch directory.
static void text_formats(const std::string& dir)
{
    const std::string adj_path = dir + "/toy.adj", stream_path = dir + "/toy.stream";
    { std::ofstream f(adj_path); f << "AdjacencyGraph\n3\n4\n0\n2\n3\n1\n2\n0\n0\n"; }
    auto [gn, gm, goff, gadj] = read_unweighted_graph(adj_path.c_str(), true);               // IO.h:67
    std::printf("adjacency n %zu m %zu offsets %llu %llu %llu edges %u %u %u %u\n", gn, gm, (unsigned long long) goff[0], (unsigned long long) goff[1],
                (unsigned long long) goff[2], gadj[0], gadj[1], gadj[2], gadj[3]);
    pbbs::free_array(goff);
    pbbs::free_array(gadj);
    { std::ofstream f(stream_path); f << "# five lines\n0 1\n1 2\n2 3\n3 4\n4 5\n"; }
    std::vector<std::pair<std::tuple<uintV, uintV>*, uintV>> stream;
    create_edge_stream(stream_path.c_str(), 2, stream);                                       // vertex-classification.cpp:5 with -eps 2
    std::printf("stream partitions %zu sizes", stream.size());
    for (auto& part : stream) std::printf(" %u", part.second);
    std::printf(" last (%u,%u) (%u,%u)\n", std::get<0>(stream.back().first[0]), std::get<1>(stream.back().first[0]),
                std::get<0>(stream.back().first[1]), std::get<1>(stream.back().first[1]));
    for (auto& part : stream) delete[] part.first;
    try { read_unweighted_graph(stream_path.c_str(), true); }
    catch (const std::exception& e) { std::printf("bad file: %s\n", e.what()); }
    std::fflush(stdout);
}

int main(int argc, char** argv_main)
{
    if (argc > 1) text_formats(argv_main[1]);
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
