// Text formats either side of the path: what the reference's drivers read before they construct
// a WharfMH and what they write after a batch.  Host code only - nothing here launches a kernel or needs a device;
// the device work of ingest (symmetrise, sort, de-duplicate) is the engine's own batch insert, and the rows that are
// written come from wmh_walks_gather.
//
//   Ligra "AdjacencyGraph" reader     libs/compressed_trees/common/IO.h:67-106 (read_unweighted_graph)
//   SNAP edge list                    experiments/bin/SNAPtoAdj input; vertex-classification.cpp:5-38 reads the same
//                                     "u v" lines for its edge stream
//   walks.txt                         vertex-classification.cpp:167-178: WharfMH::walk(id) + newline per affected walk
#include "engine.hpp"

#include <cerrno>
#include <cstdio>
#include <cstring>
#include <algorithm>
#include <string>
#include <thread>
#include <vector>

using namespace wharf;

namespace
{
inline bool is_space(char c) { return c == ' ' || c == '\t' || c == '\n' || c == '\r' || c == '\f' || c == '\v'; }

// Unsigned decimal at [p, end); advances p.  false when no digit is there or the value does not fit 64 bits.
inline bool parse_u64(const char*& p, const char* end, u64& value)
{
    if (p == end || *p < '0' || *p > '9') return false;
    u64 v = 0;
    while (p != end && *p >= '0' && *p <= '9')
    {
        const u64 d = static_cast<u64>(*p - '0');
        if (v > (~0ull - d) / 10) return false;
        v = v * 10 + d;
        ++p;
    }
    value = v;
    return true;
}

// Decimal digits of v followed by `tail`, written backwards from a small buffer.
inline char* put_u64(char* out, u64 v, char tail)
{
    char tmp[20];
    int k = 0;
    do
    {
        tmp[k++] = static_cast<char>('0' + v % 10);
        v /= 10;
    } while (v);
    while (k) *out++ = tmp[--k];
    *out++ = tail;
    return out;
}

inline u64 digits_of_u32(u32 v)
{
    return v < 100000u ? (v < 100u ? (v < 10u ? 1 : 2) : (v < 10000u ? (v < 1000u ? 3 : 4) : 5))
                       : (v < 10000000u ? (v < 1000000u ? 6 : 7) : (v < 1000000000u ? (v < 100000000u ? 8 : 9) : 10));
}

// Decimal digits of a 32-bit value followed by `tail`: the digit count is known first, so the digits go straight to
// their places, two at a time.
inline char* put_u32(char* out, u32 v, char tail)
{
    static const char kPairs[] = "0001020304050607080910111213141516171819202122232425262728293031323334353637383940414243444546474849"
                                 "5051525354555657585960616263646566676869707172737475767778798081828384858687888990919293949596979899";
    const u64 d = digits_of_u32(v);
    char* q = out + d;
    while (v >= 100)
    {
        const u32 r = v % 100;
        v /= 100;
        q -= 2;
        q[0] = kPairs[2 * r];
        q[1] = kPairs[2 * r + 1];
    }
    if (v >= 10)
    {
        q -= 2;
        q[0] = kPairs[2 * v];
        q[1] = kPairs[2 * v + 1];
    }
    else *--q = static_cast<char>('0' + v);
    out[d] = tail;
    return out + d + 1;
}

inline u64 digits_of(u64 v)
{
    u64 k = 1;
    while (v >= 10)
    {
        v /= 10;
        ++k;
    }
    return k;
}
}  // namespace

extern "C"
{

// libs/compressed_trees/common/IO.h:86-104.  Tokens: "AdjacencyGraph", n, m, n offsets, m neighbours.
int wmh_parse_adjacency_graph(const char* text, uint64_t bytes, uint64_t* n_out, uint64_t* m_out, uint64_t* offsets, uint32_t* edges)
{
    if (!text || !n_out || !m_out) return report_failure(WMH_ERR_INVALID, "wmh_parse_adjacency_graph: null argument");
    const char* p = text;
    const char* end = text + bytes;
    auto skip = [&] { while (p != end && is_space(*p)) ++p; };
    skip();
    static const char kHeader[] = "AdjacencyGraph";
    const size_t hl = sizeof(kHeader) - 1;
    if (static_cast<size_t>(end - p) < hl || std::memcmp(p, kHeader, hl) != 0 || (p + hl != end && !is_space(p[hl])))
        return report_failure(WMH_ERR_INVALID, "not an AdjacencyGraph file (IO.h:87)");
    p += hl;
    u64 n = 0, m = 0;
    skip();
    if (!parse_u64(p, end, n)) return report_failure(WMH_ERR_INVALID, "AdjacencyGraph: vertex count missing");
    skip();
    if (!parse_u64(p, end, m)) return report_failure(WMH_ERR_INVALID, "AdjacencyGraph: edge count missing");
    // every further token takes at least two bytes: a header that promises more than the text can hold is refused
    // before the caller sizes its arrays by it
    if (n > bytes || m > bytes || 2 * (n + m) > static_cast<u64>(end - p) + 1)
        return report_failure(WMH_ERR_INVALID, "Bad input file: fewer than n + m + 3 tokens (IO.h:94)");
    *n_out = n;
    *m_out = m;
    if (!offsets && !edges) return WMH_OK;                      // sizes only
    if (!offsets || (!edges && m)) return report_failure(WMH_ERR_INVALID, "wmh_parse_adjacency_graph: offsets and edges must both be given");
    u64 previous = 0;
    for (u64 i = 0; i < n; i++)
    {
        u64 v = 0;
        skip();
        if (!parse_u64(p, end, v)) return report_failure(WMH_ERR_INVALID, "Bad input file: fewer than n + m + 3 tokens (IO.h:94)");
        if (v < previous || v > m) return report_failure(WMH_ERR_INVALID, "AdjacencyGraph: offsets must be non-decreasing and <= m (offset " + std::to_string(i) + ")");
        offsets[i] = previous = v;
    }
    for (u64 i = 0; i < m; i++)
    {
        u64 v = 0;
        skip();
        if (!parse_u64(p, end, v)) return report_failure(WMH_ERR_INVALID, "Bad input file: fewer than n + m + 3 tokens (IO.h:94)");
        if (v >= n) return report_failure(WMH_ERR_INVALID, "AdjacencyGraph: neighbour id >= n (edge " + std::to_string(i) + ")");
        edges[i] = static_cast<u32>(v);
    }
    skip();
    if (p != end) return report_failure(WMH_ERR_INVALID, "Bad input file: more than n + m + 3 tokens (IO.h:94)");
    return WMH_OK;
}

uint64_t wmh_format_adjacency_graph(uint64_t n, uint64_t m, const uint64_t* offsets, const uint32_t* edges, char* out, uint64_t capacity)
{
    if ((n && !offsets) || (m && !edges)) return 0;
    u64 need = 15 + digits_of(n) + 1 + digits_of(m) + 1;
    for (u64 i = 0; i < n; i++) need += digits_of(offsets[i]) + 1;
    for (u64 i = 0; i < m; i++) need += digits_of(edges[i]) + 1;
    if (!out || capacity < need) return need;
    char* q = out;
    std::memcpy(q, "AdjacencyGraph\n", 15);
    q += 15;
    q = put_u64(q, n, '\n');
    q = put_u64(q, m, '\n');
    for (u64 i = 0; i < n; i++) q = put_u64(q, offsets[i], '\n');
    for (u64 i = 0; i < m; i++) q = put_u64(q, edges[i], '\n');
    return need;
}

// One "u v" pair per line; blank lines and lines whose first character is '#' or '%' are skipped (SNAP headers).
// *count receives the number of pairs in the text even when `pairs` is null or too small (call twice);
// *vertices = largest id + 1.
int wmh_parse_edge_list(const char* text, uint64_t bytes, uint32_t* pairs, uint64_t capacity, uint64_t* count, uint64_t* vertices)
{
    if (!text || !count) return report_failure(WMH_ERR_INVALID, "wmh_parse_edge_list: null argument");
    const char* p = text;
    const char* end = text + bytes;
    u64 k = 0, top = 0, line = 0;
    bool any = false;
    while (p != end)
    {
        ++line;
        while (p != end && *p != '\n' && is_space(*p)) ++p;
        if (p == end) break;
        if (*p == '\n')
        {
            ++p;
            continue;
        }
        if (*p == '#' || *p == '%')
        {
            while (p != end && *p != '\n') ++p;
            if (p != end) ++p;
            continue;
        }
        u64 a = 0, b = 0;
        bool ok = parse_u64(p, end, a);
        while (ok && p != end && *p != '\n' && is_space(*p)) ++p;
        ok = ok && parse_u64(p, end, b);
        while (ok && p != end && *p != '\n' && is_space(*p)) ++p;
        if (!ok || (p != end && *p != '\n')) return report_failure(WMH_ERR_INVALID, "edge list: line " + std::to_string(line) + " is not \"u v\"");
        if (a >= 0xFFFFFFFEull || b >= 0xFFFFFFFEull) return report_failure(WMH_ERR_DOMAIN, "edge list: line " + std::to_string(line) + ": vertex id does not fit 32 bits");
        if (pairs && k < capacity)
        {
            pairs[2 * k] = static_cast<u32>(a);
            pairs[2 * k + 1] = static_cast<u32>(b);
        }
        top = a > top ? a : top;
        top = b > top ? b : top;
        any = true;
        ++k;
        if (p != end) ++p;                                       // the newline that ended this pair
    }
    *count = k;
    if (vertices) *vertices = any ? top + 1 : 0;
    if (pairs && k > capacity) return report_failure(WMH_ERR_RANGE, "wmh_parse_edge_list: buffer holds " + std::to_string(capacity) + " pairs, the text has " + std::to_string(k));
    return WMH_OK;
}

// The text WharfMH::walk returns (graph/wharfmh.h:365-394: every vertex followed by one space), one walk per line.
// Large corpora are formatted by several host threads: the walks are cut into contiguous segments, every segment is
// sized, then written at its offset.
uint64_t wmh_format_walks(const uint32_t* rows, uint64_t count, uint32_t walk_length, char* out, uint64_t capacity)
{
    if (count && !rows) return 0;
    const u64 kWalksPerThread = 8192;
    unsigned hw = std::thread::hardware_concurrency();
    if (hw == 0) hw = 1;
    const u64 parts = std::max<u64>(1, std::min<u64>({(u64) hw, 16, count / kWalksPerThread}));
    std::vector<u64> begin(parts + 1), bytes(parts + 1, 0);
    for (u64 s = 0; s <= parts; s++) begin[s] = count / parts * s + std::min<u64>(s, count % parts);
    auto size_of = [&](u64 s) {
        u64 need = 0;
        for (u64 w = begin[s]; w < begin[s + 1]; w++)
        {
            const u32* row = rows + w * walk_length;
            for (u32 p = 0; p < walk_length && row[p] != WMH_NO_VERTEX; p++) need += digits_of_u32(row[p]) + 1;
            need += 1;
        }
        bytes[s + 1] = need;
    };
    auto write_of = [&](u64 s) {
        char* q = out + bytes[s];
        for (u64 w = begin[s]; w < begin[s + 1]; w++)
        {
            const u32* row = rows + w * walk_length;
            for (u32 p = 0; p < walk_length && row[p] != WMH_NO_VERTEX; p++) q = put_u32(q, row[p], ' ');
            *q++ = '\n';
        }
    };
    auto run = [&](auto&& job) {
        if (parts == 1) { job(0); return; }
        std::vector<std::thread> pool;
        for (u64 s = 1; s < parts; s++) pool.emplace_back([&job, s] { job(s); });
        job(0);
        for (auto& t : pool) t.join();
    };
    run(size_of);
    for (u64 s = 0; s < parts; s++) bytes[s + 1] += bytes[s];      // -> offsets
    const u64 need = bytes[parts];
    if (!out || capacity < need) return need;
    run(write_of);
    return need;
}

// vertex-classification.cpp:167-178: the walks a batch changed, as text, for the downstream trainer.
int wmh_write_walks(wmh_engine* e, const uint64_t* walk_ids, uint64_t count, const char* path, int append, uint64_t* bytes_written)
{
    if (!e || !path || (count && !walk_ids)) return report_failure(WMH_ERR_INVALID, "wmh_write_walks: null argument");
    wmh_config cfg;
    if (wmh_get_config(e, &cfg) != WMH_OK) return WMH_ERR_INVALID;
    const u32 L = cfg.walk_length;                               // fixed while a corpus exists (wmh_set_config)
    if (count && !wmh_device_walks(e)) return report_failure(WMH_ERR_STATE, "wmh_write_walks: walks have not been generated");
    std::FILE* f = std::fopen(path, append ? "ab" : "wb");
    if (!f) return report_failure(WMH_ERR_INVALID, std::string("wmh_write_walks: cannot open ") + path + ": " + std::strerror(errno));
    const u64 kChunk = 1u << 16;
    std::vector<u32> local, rows;
    std::vector<char> text;
    u64 total = 0;
    int rc = WMH_OK;
    for (u64 first = 0; first < count && rc == WMH_OK; first += kChunk)
    {
        const u64 k = count - first < kChunk ? count - first : kChunk;
        local.resize(k);
        for (u64 i = 0; i < k; i++)
        {
            const int64_t l = wmh_global_to_local_walk(e, walk_ids[first + i]);
            if (l < 0)
            {
                rc = report_failure(WMH_ERR_RANGE, "wmh_write_walks: walk " + std::to_string(walk_ids[first + i]) + " is not held by this engine (rank)");
                break;
            }
            local[i] = static_cast<u32>(l);
        }
        if (rc != WMH_OK) break;
        rows.resize(k * L);
        rc = wmh_walks_gather(e, local.data(), k, rows.data());
        if (rc != WMH_OK) break;
        const u64 need = wmh_format_walks(rows.data(), k, L, nullptr, 0);
        text.resize(need);
        wmh_format_walks(rows.data(), k, L, text.data(), need);
        if (std::fwrite(text.data(), 1, need, f) != need) rc = report_failure(WMH_ERR_INVALID, std::string("wmh_write_walks: short write to ") + path);
        total += need;
    }
    if (std::fclose(f) != 0 && rc == WMH_OK) rc = report_failure(WMH_ERR_INVALID, std::string("wmh_write_walks: close failed for ") + path);
    if (bytes_written) *bytes_written = total;
    return rc;
}

}  // extern "C"
