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
        { std::string b = blob; b[b.size() - 5] ^= 0x01; std::ofstream(path, std::ios::binary) << b; }
        CHECK(!read_lists_cache(path, key, P, G, v.size(), so2, v2));            // one bit of a tetramer flipped (the segment stays ascending): checksum of vals
        { std::string b = blob; b[56 + 8 * (P * G + 1) + 1] ^= 0x40; std::ofstream(path, std::ios::binary) << b; }
        CHECK(!read_lists_cache(path, key, P, G, v.size(), so2, v2));            // first posting damaged
        { std::string b = blob; b[7] = '1'; std::ofstream(path, std::ios::binary) << b; }
        CHECK(!lists_cache_header_ok(path, key) && !read_lists_cache(path, key, P, G, v.size(), so2, v2));   // the older PFLST001 layout is not trusted
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
    // ---- query-DB mode on the cached main index (CachedQTDataBase) == the ATTACH path, up to the tetramers only one side has ----
    if (argc >= 4) {
        const std::string mainDb = argv[2], qryDb = argv[3], cache = dir + "/main.pfidx";
        auto genome_major = [](const HostIndex &x, std::vector<int64_t> &so, std::vector<uint32_t> &sv) {
            const int P = x.nProteins, G = x.nGenomes;
            so.assign((size_t)G * P + 1, 0);
            for (const pfaai_pg_t &e : x.F) so[(size_t)e.genome * P + e.protein + 1]++;
            for (size_t i = 1; i < so.size(); i++) so[i] += so[i - 1];
            sv.resize(x.F.size());
            std::vector<int64_t> cur(so.begin(), so.end() - 1);
            size_t k = 0;
            for (uint32_t t = 0; t < (uint32_t)NTETRAMERS; t++)
                for (int32_t c = 0; c < x.Lc[t]; c++, k++) sv[(size_t)cur[(size_t)x.F[k].genome * P + x.F[k].protein]++] = t;
        };
        {   // what a first single-DB run leaves behind: .pfidx (names, Lc, F, T) + .lists (device-ready lists)
            SQLiteSCPDataBase m(mainDb);
            CHECK(m.getDBErrorCode() == SQLITE_OK);
            HostIndex ix;
            CHECK(m.loadIndex(ix) == SQLITE_OK);
            IndexCacheKey key;
            CHECK(index_cache_key(mainDb, key));
            CHECK(write_index_cache(cache, key, m.getMeta(), ix));
            std::vector<int64_t> so; std::vector<uint32_t> sv;
            genome_major(ix, so, sv);
            CHECK(write_lists_cache(cache + ".lists", key, ix.nProteins, ix.nGenomes, sv.size(), so.data(), sv.data()));
        }
        QTSQLiteSCPDataBase ref(mainDb, qryDb);
        CachedQTDataBase inc(mainDb, qryDb, cache);
        CHECK(ref.validate() == PFAAI_OK && inc.validate() == PFAAI_OK && inc.fromCache());
        CHECK(ref.getMeta().proteinSet == inc.getMeta().proteinSet);            // same SCPs in the same (summation) order
        CHECK(ref.getMeta().genomeSet == inc.getMeta().genomeSet && ref.getMeta().qyGenomeSet == inc.getMeta().qyGenomeSet);
        HostIndex a, b;
        CHECK(ref.loadIndex(a) == SQLITE_OK && inc.loadIndex(b) == SQLITE_OK);
        CHECK(a.nProteins == b.nProteins && a.nGenomes == b.nGenomes && a.nTargets == b.nTargets && a.nQueries == b.nQueries);
        CHECK(a.T == b.T && b.haveLists() && b.F.empty());
        // per SCP: restrict the incremental lists to the tetramers that occur on BOTH sides -> the inner-join index of the reference
        std::vector<int64_t> so; std::vector<uint32_t> sv;
        genome_major(a, so, sv);
        const int P = a.nProteins, G = a.nGenomes, nT = a.nTargets;
        size_t at = 0;
        for (int p = 0; p < P; p++) {
            std::vector<char> inT(NTETRAMERS, 0), inQ(NTETRAMERS, 0);
            for (int g = 0; g < G; g++)
                for (int64_t k = b.segOff[(size_t)g * P + p]; k < b.segOff[(size_t)g * P + p + 1]; k++) (g < nT ? inT : inQ)[b.segVals[(size_t)k]] = 1;
            for (int g = 0; g < G; g++) {
                std::vector<uint32_t> kept;
                for (int64_t k = b.segOff[(size_t)g * P + p]; k < b.segOff[(size_t)g * P + p + 1]; k++)
                    if (inT[b.segVals[(size_t)k]] && inQ[b.segVals[(size_t)k]]) kept.push_back(b.segVals[(size_t)k]);
                const std::vector<uint32_t> want(sv.begin() + so[(size_t)g * P + p], sv.begin() + so[(size_t)g * P + p + 1]);
                CHECK(kept == want);
                at += kept.size();
            }
        }
        CHECK(at == a.F.size());
        // a main database that changed after the cache was written: the cache is ignored, the ATTACH path answers
        IndexCacheKey stale; CHECK(index_cache_key(mainDb, stale)); stale.size += 1;
        { HostIndex ix; DBMetaData meta; SQLiteSCPDataBase m(mainDb); CHECK(m.loadIndex(ix) == SQLITE_OK); CHECK(write_index_cache(cache, stale, m.getMeta(), ix)); }
        CachedQTDataBase fallback(mainDb, qryDb, cache);
        CHECK(!fallback.fromCache() && fallback.validate() == PFAAI_OK && fallback.getMeta().proteinSet == ref.getMeta().proteinSet);
    }
    std::puts("OK");
    return 0;
}
