// CPU-only unit checks of host pieces that otherwise need a GPU run to be reached: the device-ready lists sidecar of the index
// cache (index_cache.hpp) and the streaming CSV writer (csv_writer.hpp) driven by a scripted progress source.
// Built and run by tests/test_host_units.py; prints "OK" and exits 0 when every check holds.
#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "../parfastaai_b200/host/csv_writer.hpp"
#include "../parfastaai_b200/host/index_cache.hpp"

using namespace pfaai_host;

#define CHECK(c) do { if (!(c)) { std::fprintf(stderr, "FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); return 1; } } while (0)

static std::string slurp(const std::string &p) { std::ifstream f(p, std::ios::binary); std::stringstream s; s << f.rdbuf(); return s.str(); }

int main(int argc, char **argv) {
    if (argc < 2) return 2;
    const std::string dir = argv[1];
    // ---- lists sidecar: round trip, wrong key / shape, damage, truncation ----
    {
        const int P = 3, G = 5;
        std::vector<int64_t> so(P * G + 1);
        std::vector<uint32_t> v;
        for (int i = 0; i < P * G; i++) { so[i] = (int64_t)v.size(); for (int k = 0; k <= i % 4; k++) v.push_back(1000u * i + k); }
        so[P * G] = (int64_t)v.size();
        IndexCacheKey key; key.size = 12345; key.mtime_ns = 987654321;
        const std::string path = dir + "/x.lists";
        CHECK(write_lists_cache(path, key, P, G, v.size(), so.data(), v.data()));
        CHECK(lists_cache_header_ok(path, key));
        std::vector<int64_t> so2; std::vector<uint32_t> v2;
        CHECK(read_lists_cache(path, key, P, G, v.size(), so2, v2));
        CHECK(so2 == so && v2 == v);
        IndexCacheKey other = key; other.mtime_ns++;
        CHECK(!lists_cache_header_ok(path, other));
        CHECK(!read_lists_cache(path, other, P, G, v.size(), so2, v2));          // another version of the database
        CHECK(!read_lists_cache(path, key, P, G + 1, v.size(), so2, v2));        // another shape
        CHECK(!read_lists_cache(path, key, P, G, v.size() + 1, so2, v2));        // another |F|
        std::string blob = slurp(path);
        { std::string b = blob; b[60] ^= 0x55; std::ofstream(path, std::ios::binary) << b; }
        CHECK(!read_lists_cache(path, key, P, G, v.size(), so2, v2));            // damaged offsets: checksum
        { std::ofstream(path, std::ios::binary) << blob.substr(0, blob.size() - 3); }
        CHECK(!read_lists_cache(path, key, P, G, v.size(), so2, v2));            // truncated
        { std::ofstream(path, std::ios::binary) << blob + "x"; }
        CHECK(!read_lists_cache(path, key, P, G, v.size(), so2, v2));            // trailing bytes
        CHECK(so2 == so && v2 == v);                                             // failed reads leave the outputs alone
    }
    // ---- streaming writer == plain writer, whatever the progress schedule ----
    {
        const size_t n = 37;
        std::vector<std::string> names;
        std::vector<double> M(n * n);
        for (size_t i = 0; i < n; i++) names.push_back("g" + std::to_string(i));
        for (size_t i = 0; i < n; i++) for (size_t j = 0; j < n; j++) M[i * n + j] = i == j ? 0.0 : 1.0 / (double)(1 + i + j);
        const std::string a = dir + "/plain.csv", b = dir + "/stream.csv";
        CHECK(write_aji_csv(a, ";", names, names, M.data()));
        const int64_t schedule[] = {0, 0, 5, 5, 5, 12, 30, 30};                 // rows_final per poll, then "finished"
        size_t at = 0, overlapped = 0;
        auto poll = [&](int64_t &rows, int &fin) {
            if (at < sizeof(schedule) / sizeof(schedule[0])) { rows = schedule[at++]; fin = 0; }
            else { rows = (int64_t)n; fin = 1; }
            return true;
        };
        CHECK(write_aji_csv_streaming(b, ";", names, names, M.data(), poll, &overlapped));
        CHECK(slurp(a) == slurp(b) && overlapped == 30);
        // a compute that fails (poll returns false) must not report success, one that finishes without all rows neither
        auto bad = [&](int64_t &, int &) { return false; };
        CHECK(!write_aji_csv_streaming(b, ";", names, names, M.data(), bad, nullptr));
        auto shortfall = [&](int64_t &rows, int &fin) { rows = 3; fin = 1; return true; };
        CHECK(!write_aji_csv_streaming(b, ";", names, names, M.data(), shortfall, nullptr));
    }
    std::puts("OK");
    return 0;
}
