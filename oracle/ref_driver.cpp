// TEST INFRASTRUCTURE ONLY — never linked into, imported by, or executed from the product path.
//
// Command-line driver around the UNMODIFIED reference (djordjijeK/DynamicGraphRepresentationLearning).
// It #includes the reference's headers where they lie (/root/reference, via -I flags in
// oracle/Makefile) and is compiled into oracle/_ref/wharf_ref.  No reference source is copied here.
//
// Purpose: (1) pin the plain-C restatement (oracle/wharf_oracle.c) and the golden fixtures under
// tests/golden/ against the real reference, (2) serve as `cpu_baseline.kind == "reference"` in bench.py.
//
// Binary formats (little endian):
//   graph file : u64 n, u64 m, u64 offsets[n], u32 edges[m]       (reference ctor layout: n offsets, wharfmh.h:58)
//   batch file : u64 count, then count x (u32 src, u32 dst)
//   walks file : u64 n_walks, u64 walk_length, u32 vertex[n_walks][walk_length]; UINT32_MAX = "no vertex"
//   ids   file : u64 count, u64 id[count]   (affected walk ids, sorted ascending)
//
// Commands:
//   wharf_ref rng <seed> <k>                       -> k lrand() outputs of utility::Random(seed)
//   wharf_ref drand <seed> <k>                     -> k drand() outputs (printed as %.17g)
//   wharf_ref pair32 <x> <y> | unpair32 <z> | triplet32 a b c | untriplet32 z | pair64/unpair64/...
//   wharf_ref genbatch <edges> <vertices> <seed> <self_loops> <directed> <out.batch>
//   wharf_ref sampler <model> <p> <q> <iters> <cur> <prev> <seed>   (tests/sampler.cpp fixture, prints counts)
//   wharf_ref run <graph> <out_prefix> [key=value ...] [ins:<batch>|del:<batch> ...]
//        keys: wpv, len, model(deepwalk|node2vec), p, q, det(0|1), init(random|burnin|weight),
//              sorted(0|1), dedup(0|1), nn, seq(0|1), dump(0|1), trials, rngseed
//        writes <out_prefix>.walks.<k>.bin after init (k=0) and after every batch op (k=1..) when dump=1,
//        <out_prefix>.affected.<k>.bin for each op, and prints one JSON line per phase with timings
//        (init_times = wall seconds of each of the `trials` generate_initial_random_walks calls).

#include <wharfmh.h>

#include <chrono>
#include <cinttypes>
#include <cstdio>
#include <cstring>
#include <map>
#include <string>
#include <vector>

namespace
{
    using Edge = std::tuple<uintV, uintV>;

    double now_s()
    {
        return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
    }

    void die(const std::string& msg)
    {
        std::fprintf(stderr, "wharf_ref: %s\n", msg.c_str());
        std::exit(2);
    }

    void read_graph(const char* path, size_t& n, size_t& m, uintE*& offsets, uintV*& edges)
    {
        FILE* f = std::fopen(path, "rb");
        if (!f) die(std::string("cannot open graph ") + path);
        uint64_t hdr[2];
        if (std::fread(hdr, 8, 2, f) != 2) die("short graph header");
        n = hdr[0]; m = hdr[1];
        offsets = pbbs::new_array_no_init<uintE>(n);
        edges   = pbbs::new_array_no_init<uintV>(m ? m : 1);
        static_assert(sizeof(uintE) == 8, "build with -DEDGELONG");
        static_assert(sizeof(uintV) == 4, "uintV must be 32-bit");
        if (std::fread(offsets, 8, n, f) != n) die("short offsets");
        if (m && std::fread(edges, 4, m, f) != m) die("short edges");
        std::fclose(f);
    }

    std::pair<Edge*, size_t> read_batch(const char* path)
    {
        FILE* f = std::fopen(path, "rb");
        if (!f) die(std::string("cannot open batch ") + path);
        uint64_t cnt;
        if (std::fread(&cnt, 8, 1, f) != 1) die("short batch header");
        std::vector<uint32_t> raw(2 * cnt);
        if (cnt && std::fread(raw.data(), 4, 2 * cnt, f) != 2 * cnt) die("short batch");
        std::fclose(f);
        Edge* e = pbbs::new_array<Edge>(cnt ? cnt : 1);
        for (size_t i = 0; i < cnt; i++) e[i] = std::make_tuple(raw[2 * i], raw[2 * i + 1]);
        return {e, cnt};
    }

    void write_batch(const char* path, Edge* e, size_t cnt)
    {
        FILE* f = std::fopen(path, "wb");
        if (!f) die(std::string("cannot write ") + path);
        uint64_t c = cnt;
        std::fwrite(&c, 8, 1, f);
        std::vector<uint32_t> raw(2 * cnt);
        for (size_t i = 0; i < cnt; i++) { raw[2 * i] = std::get<0>(e[i]); raw[2 * i + 1] = std::get<1>(e[i]); }
        if (cnt) std::fwrite(raw.data(), 4, 2 * cnt, f);
        std::fclose(f);
    }

    // Dumps the corpus through the reference's own retrieval path (walk() semantics, wharfmh.h:365-394):
    // follow find_next until the sentinel; pad the row with UINT32_MAX.  Also self-checks:
    // every non-isolated start must retrieve exactly L vertices.
    size_t dump_walks(dygrl::WharfMH& w, const std::string& path, bool write)
    {
        const size_t n = w.number_of_vertices();
        const size_t L = config::walk_length;
        const size_t W = n * config::walks_per_vertex;
        std::vector<uint32_t> out(W * L, UINT32_MAX);
        auto flat = w.flatten_vertex_tree();
        std::atomic<size_t> bad(0);
        parallel_for(0, W, [&](size_t walk_id)
        {
            types::Vertex cur = walk_id % n;
            size_t pos = 0;
            while (cur != std::numeric_limits<uint32_t>::max() - 1 && pos < L)
            {
                out[walk_id * L + pos] = cur;
                auto r = flat[cur].inverted_index.find((uintV)(walk_id * L + pos));
                if (!r.valid) { bad++; break; }
                cur = r.value;
                pos++;
            }
        });
        if (bad.load()) std::fprintf(stderr, "wharf_ref: SELF-CHECK FAILED: %zu walks hit a missing index entry\n", bad.load());
        if (write)
        {
            FILE* f = std::fopen(path.c_str(), "wb");
            if (!f) die("cannot write " + path);
            uint64_t hdr[2] = {W, L};
            std::fwrite(hdr, 8, 2, f);
            std::fwrite(out.data(), 4, out.size(), f);
            std::fclose(f);
        }
        return bad.load();
    }

    void write_ids(const std::string& path, pbbs::sequence<types::WalkID>& ids)
    {
        std::vector<uint64_t> v(ids.size());
        for (size_t i = 0; i < ids.size(); i++) v[i] = ids[i];
        std::sort(v.begin(), v.end());
        FILE* f = std::fopen(path.c_str(), "wb");
        if (!f) die("cannot write " + path);
        uint64_t c = v.size();
        std::fwrite(&c, 8, 1, f);
        if (c) std::fwrite(v.data(), 8, c, f);
        std::fclose(f);
    }

    int cmd_run(int argc, char** argv)
    {
        if (argc < 4) die("run <graph> <out_prefix> ...");
        std::string prefix = argv[3];
        std::map<std::string, std::string> kv = {
            {"wpv", "10"}, {"len", "80"}, {"model", "deepwalk"}, {"p", "0.5"}, {"q", "2.0"}, {"det", "1"},
            {"init", "weight"}, {"sorted", "0"}, {"dedup", "1"}, {"nn", ""}, {"seq", "0"}, {"dump", "1"},
            {"trials", "1"}, {"rngseed", ""}};
        std::vector<std::pair<char, std::string>> ops;
        for (int i = 4; i < argc; i++)
        {
            std::string a = argv[i];
            if (a.rfind("ins:", 0) == 0) { ops.push_back({'i', a.substr(4)}); continue; }
            if (a.rfind("del:", 0) == 0) { ops.push_back({'d', a.substr(4)}); continue; }
            auto eq = a.find('=');
            if (eq == std::string::npos) die("bad arg " + a);
            kv[a.substr(0, eq)] = a.substr(eq + 1);
        }

        config::walks_per_vertex = (uint8_t) std::stoi(kv["wpv"]);
        config::walk_length      = (uint8_t) std::stoi(kv["len"]);
        config::random_walk_model = kv["model"] == "node2vec" ? types::RandomWalkModelType::NODE2VEC
                                                             : types::RandomWalkModelType::DEEPWALK;
        config::paramP = std::stof(kv["p"]);
        config::paramQ = std::stof(kv["q"]);
        config::deterministic_mode = kv["det"] != "0";
        config::sampler_init_strategy = kv["init"] == "random" ? types::SamplerInitStartegy::RANDOM
                                      : kv["init"] == "burnin" ? types::SamplerInitStartegy::BURNIN
                                                               : types::SamplerInitStartegy::WEIGHT;
        if (!kv["rngseed"].empty()) config::random = utility::Random(std::stoull(kv["rngseed"]));
        const bool sorted = kv["sorted"] != "0", dedup = kv["dedup"] != "0", seq = kv["seq"] != "0";
        const bool dump = kv["dump"] != "0";
        const int trials = std::stoi(kv["trials"]);

        size_t n, m; uintE* offsets; uintV* edges;
        read_graph(argv[2], n, m, offsets, edges);
        double t0 = now_s();
        dygrl::WharfMH engine(n, m, offsets, edges, true);
        double t_build = now_s() - t0;

        // initial generation (trials > 1 re-generates after destroy_index, throughput-latency.cpp:176-187)
        double t_init = 0;
        std::string init_times = "[";
        for (int t = 0; t < trials; t++)
        {
            if (t) engine.destroy_index();
            t0 = now_s();
            engine.generate_initial_random_walks();
            double dt = now_s() - t0;
            t_init = (t == 0) ? dt : std::min(t_init, dt);
            init_times += (t ? "," : "") + std::to_string(dt);
        }
        init_times += "]";
        size_t nv = engine.number_of_vertices();
        size_t nonisolated = 0;
        {
            auto flat = engine.flatten_graph();
            for (size_t v = 0; v < nv; v++) nonisolated += flat[v].degree > 0;
        }
        size_t bad = dump_walks(engine, prefix + ".walks.0.bin", dump);
        std::printf("{\"phase\":\"init\",\"threads\":%d,\"n\":%zu,\"m\":%zu,\"nonisolated\":%zu,\"build_s\":%.6f,"
                    "\"init_s\":%.6f,\"init_times\":%s,\"steps\":%zu,\"selfcheck_bad\":%zu}\n",
                    num_workers(), nv, (size_t) engine.number_of_edges(), nonisolated, t_build, t_init, init_times.c_str(),
                    nonisolated * (size_t) config::walks_per_vertex * (size_t) config::walk_length, bad);
        std::fflush(stdout);

        int k = 1;
        for (auto& op : ops)
        {
            auto batch = read_batch(op.second.c_str());
            size_t nn = kv["nn"].empty() ? std::numeric_limits<size_t>::max() : std::stoull(kv["nn"]);
            graph_update_time_on_insert.reset(); walk_update_time_on_insert.reset();
            graph_update_time_on_delete.reset(); walk_update_time_on_delete.reset();
            t0 = now_s();
            auto affected = (op.first == 'i')
                ? engine.insert_edges_batch(batch.second, batch.first, sorted, dedup, nn, true, seq)
                : engine.delete_edges_batch(batch.second, batch.first, sorted, dedup, nn, true, seq);
            double dt = now_s() - t0;
            double tg = op.first == 'i' ? graph_update_time_on_insert.get_total() : graph_update_time_on_delete.get_total();
            double tw = op.first == 'i' ? walk_update_time_on_insert.get_total() : walk_update_time_on_delete.get_total();
            write_ids(prefix + ".affected." + std::to_string(k) + ".bin", affected);
            bad = dump_walks(engine, prefix + ".walks." + std::to_string(k) + ".bin", dump);
            std::printf("{\"phase\":\"%s\",\"batch\":%zu,\"affected\":%zu,\"m\":%zu,\"total_s\":%.6f,\"graph_s\":%.6f,"
                        "\"walk_s\":%.6f,\"selfcheck_bad\":%zu}\n",
                        op.first == 'i' ? "insert" : "delete", batch.second, affected.size(),
                        (size_t) engine.number_of_edges(), dt, tg, tw, bad);
            std::fflush(stdout);
            pbbs::free_array(batch.first);
            k++;
        }
        return 0;
    }

    // tests/sampler.cpp:43-127 restated as a command: fixed state, `iters` calls of sample(), counts per vertex.
    int cmd_sampler(int argc, char** argv)
    {
        if (argc < 9) die("sampler <model> <p> <q> <iters> <cur> <prev> <seed>");
        std::string model_name = argv[2];
        float p = std::stof(argv[3]), q = std::stof(argv[4]);
        int iters = std::stoi(argv[5]);
        uintV cur = std::stoul(argv[6]), prev = std::stoul(argv[7]);
        config::random = utility::Random(std::stoull(argv[8]));

        const long nv = 6, ne = 18;
        uintE* offsets = pbbs::new_array_no_init<uintE>(nv);
        uintV* edges   = pbbs::new_array_no_init<uintV>(ne);
        const uintE off[6]  = {0, 2, 5, 10, 13, 15};
        const uintV adj[18] = {1, 2, 0, 2, 3, 0, 1, 3, 4, 5, 1, 2, 5, 2, 5, 2, 3, 4};
        for (int i = 0; i < nv; i++) offsets[i] = off[i];
        for (int i = 0; i < ne; i++) edges[i] = adj[i];

        dygrl::WharfMH engine(nv, ne, offsets, edges);
        auto* graph = new dygrl::FlatGraph(engine.flatten_graph());
        dygrl::RandomWalkModel* model = model_name == "node2vec"
            ? (dygrl::RandomWalkModel*) new dygrl::Node2Vec(graph, p, q)
            : (dygrl::RandomWalkModel*) new dygrl::DeepWalk(graph);
        types::State state(cur, model_name == "node2vec" ? prev : cur);
        dygrl::MetropolisHastingsSampler sampler(state, model);
        std::map<uintV, long> counts;
        for (int i = 0; i < iters; i++) counts[sampler.sample(state, model).first]++;
        std::printf("{\"init\":%u,\"counts\":{", sampler.previously_sampled_vertex());
        bool first = true;
        for (auto& c : counts) { std::printf("%s\"%u\":%ld", first ? "" : ",", c.first, c.second); first = false; }
        std::printf("}}\n");
        std::fflush(stdout);
        std::_Exit(0);   // the reference model dtor deletes a snapshot it does not own (deepwalk.h:28-31)
    }
}

int main(int argc, char** argv)
{
    if (argc < 2) die("missing command");
    std::string cmd = argv[1];

    if (cmd == "rng" || cmd == "drand")
    {
        utility::Random r(std::stoull(argv[2]));
        int k = std::stoi(argv[3]);
        for (int i = 0; i < k; i++)
        {
            if (cmd == "rng") std::printf("%" PRIu64 "\n", r.lrand());
            else              std::printf("%.17g\n", r.drand());
        }
        return 0;
    }
    if (cmd == "pair32")      { std::printf("%u\n", pairings::Szudzik<uint32_t>::pair({(uint32_t) std::stoul(argv[2]), (uint32_t) std::stoul(argv[3])})); return 0; }
    if (cmd == "unpair32")    { auto r = pairings::Szudzik<uint32_t>::unpair((uint32_t) std::stoul(argv[2])); std::printf("%u %u\n", r.first, r.second); return 0; }
    if (cmd == "triplet32")   { std::printf("%u\n", pairings::Szudzik<uint32_t>::pair_triplet({(uint32_t) std::stoul(argv[2]), (uint32_t) std::stoul(argv[3]), (uint32_t) std::stoul(argv[4])})); return 0; }
    if (cmd == "untriplet32") { auto r = pairings::Szudzik<uint32_t>::unpair_triplet((uint32_t) std::stoul(argv[2])); std::printf("%u %u %u\n", std::get<0>(r), std::get<1>(r), std::get<2>(r)); return 0; }
    if (cmd == "pair64")      { std::printf("%zu\n", pairings::Szudzik<size_t>::pair({std::stoull(argv[2]), std::stoull(argv[3])})); return 0; }
    if (cmd == "unpair64")    { auto r = pairings::Szudzik<size_t>::unpair(std::stoull(argv[2])); std::printf("%zu %zu\n", r.first, r.second); return 0; }
    if (cmd == "triplet64")   { std::printf("%zu\n", pairings::Szudzik<size_t>::pair_triplet({std::stoull(argv[2]), std::stoull(argv[3]), std::stoull(argv[4])})); return 0; }
    if (cmd == "untriplet64") { auto r = pairings::Szudzik<size_t>::unpair_triplet(std::stoull(argv[2])); std::printf("%zu %zu %zu\n", std::get<0>(r), std::get<1>(r), std::get<2>(r)); return 0; }

    if (cmd == "genbatch")
    {
        if (argc < 8) die("genbatch <edges> <vertices> <seed> <self_loops> <directed> <out>");
        auto b = utility::generate_batch_of_edges(std::stoull(argv[2]), std::stoull(argv[3]), std::stoull(argv[4]),
                                                  std::stoi(argv[5]) != 0, std::stoi(argv[6]) != 0);
        write_batch(argv[7], b.first, b.second);
        std::printf("{\"generated\":%zu}\n", b.second);
        return 0;
    }
    if (cmd == "sampler") return cmd_sampler(argc, argv);
    if (cmd == "run")     return cmd_run(argc, argv);
    die("unknown command " + cmd);
    return 2;
}
