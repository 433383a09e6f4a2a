// par_fastaai_main.cpp -- `par_fastaai.x`, the reference's command line on the B200 AJI path (C++17 host).
//
//   par_fastaai.x [-r query.db] [-s sep] [-q query_list.txt] path_to_input_db path_to_output_file
//
// Same options, modes, stdout banners, exit codes and CSV bytes as the reference's src/main.cpp:56-356
// (all-vs-all :177-202, query subset :234-266, query DB vs main DB :302-335, printOutput :133-175);
// exit codes: 0 ok, 1 DB open failure, 3 bad subset / overlapping query DB / construct failure,
// 105 missing input file, 106 missing required argument, 109 unknown argument (CLI11's codes).
// Everything between the SQLite load and the CSV writer runs on the GPU behind include/pfaai_gpu.h.
// Extra switches of this build: --device N, --gpus N (row-block sharding over N GPUs), --no-ref-compat (query-DB mode: corrected T lookup instead of the
// reference's label quirk), --dump-index PREFIX (host only: write Lc/F/T as raw little-endian arrays and exit;
// used by the CPU tests of the loader), --format-doubles FILE (test hook of the CSV cell formatter).
#include <sys/stat.h>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <memory>
#include <string>
#include <unordered_set>
#include <vector>

#include "csv_writer.hpp"
#include "fastaai_reader.hpp"
#include "index_cache.hpp"
#include "pfaai_driver.hpp"

using namespace pfaai_host;

namespace {

struct AppParams {
    std::string pathToDatabase, pathToQryDatabase, pathToQrySubsetFile, pathToOutputFile;
    std::string outFieldSeparator = ",";
    std::vector<std::string> qryGenomeSet;
    std::string dumpIndexPrefix, formatDoublesFile, indexCachePath;
    int device = 0, nGpus = 1;
    bool refCompat = true, sameDevice = false, noStreamCsv = false;

    static bool is_file(const std::string &p) {
        struct stat st;
        return ::stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode);
    }
    static void usage(std::ostream &os, const char *argv0) {
        os << "Usage: " << argv0 << " [OPTIONS] path_to_input_db path_to_output_file\n\n"
           << "Positionals:\n"
           << "  path_to_input_db TEXT:FILE REQUIRED   Path to the Input Database\n"
           << "  path_to_output_file TEXT REQUIRED     Path to output csv file.\n\n"
           << "Options:\n"
           << "  -h,--help                   Print this help message and exit\n"
           << "  -r,--query_db TEXT:FILE     Path to the Query Database [Optional (default: Same as the Input DB)]\n"
           << "  -s,--separator TEXT [,]     Field Separator in the output file [Optional (default: ,)].\n"
           << "  -q,--query_subset TEXT:FILE Path to Query List (Should be subset of genomoes in the input DB.)\n"
           << "  --device INT [0]            first CUDA device ordinal\n"
           << "  --gpus INT [1]              number of GPUs (devices device..device+gpus-1); output row blocks are sharded over them\n"
           << "  --no-ref-compat             query-DB mode: corrected T lookup (default reproduces the reference)\n"
           << "  --no-stream-csv             write the CSV after the computation (default, one GPU all-vs-all: rows are written while later row blocks are computed)\n"
           << "  --index-cache TEXT          device-ready index cache of the input DB (read if valid, else written by the single-DB modes; with -r the\n"
           << "                              cached main DB is reused and only the query DB is read)\n"
           << "  --dump-index TEXT           host only: write PREFIX.{lc,f,t}.bin and exit\n";
    }
    // returns -1 to continue, otherwise the process exit code
    int parse(int argc, char *argv[]) {
        std::vector<std::string> pos;
        auto fail = [&](int code, const std::string &msg) {
            std::cerr << msg << "\nRun with --help for more information.\n";
            return code;
        };
        for (int i = 1; i < argc; i++) {
            std::string a = argv[i], val;
            bool has_val = false;
            if (a.rfind("--", 0) == 0) {
                const size_t eq = a.find('=');
                if (eq != std::string::npos) { val = a.substr(eq + 1); a = a.substr(0, eq); has_val = true; }
            } else if (a.size() > 2 && a[0] == '-' && a[1] != '-') {
                val = a.substr(2); a = a.substr(0, 2); has_val = true;
            }
            auto need = [&](std::string &dst) -> bool {
                if (has_val) { dst = val; return true; }
                if (i + 1 >= argc) return false;
                dst = argv[++i];
                return true;
            };
            if (a == "-h" || a == "--help") { usage(std::cout, argv[0]); return 0; }
            else if (a == "-r" || a == "--query_db") {
                if (!need(pathToQryDatabase)) return fail(114, a + ": 1 required TEXT:FILE missing");
                if (!is_file(pathToQryDatabase)) return fail(105, "--query_db: File does not exist: " + pathToQryDatabase);
            } else if (a == "-s" || a == "--separator") {
                if (!need(outFieldSeparator)) return fail(114, a + ": 1 required TEXT missing");
            } else if (a == "-q" || a == "--query_subset") {
                if (!need(pathToQrySubsetFile)) return fail(114, a + ": 1 required TEXT:FILE missing");
                if (!is_file(pathToQrySubsetFile)) return fail(105, "--query_subset: File does not exist: " + pathToQrySubsetFile);
            } else if (a == "--device") {
                std::string d;
                if (!need(d)) return fail(114, "--device: 1 required INT missing");
                device = std::atoi(d.c_str());
            } else if (a == "--gpus") {
                std::string d;
                if (!need(d)) return fail(114, "--gpus: 1 required INT missing");
                nGpus = std::max(1, std::atoi(d.c_str()));
            } else if (a == "--same-device") sameDevice = true;   // test hook: all shards on one GPU
            else if (a == "--no-ref-compat") refCompat = false;
            else if (a == "--no-stream-csv") noStreamCsv = true;
            else if (a == "--index-cache") {
                if (!need(indexCachePath)) return fail(114, "--index-cache: 1 required TEXT missing");
            } else if (a == "--dump-index") {
                if (!need(dumpIndexPrefix)) return fail(114, "--dump-index: 1 required TEXT missing");
            } else if (a == "--format-doubles") {
                if (!need(formatDoublesFile)) return fail(114, "--format-doubles: 1 required TEXT missing");
            } else if (a.size() > 1 && a[0] == '-') return fail(109, "The following argument was not expected: " + std::string(argv[i]));
            else pos.push_back(argv[i]);
        }
        if (!formatDoublesFile.empty()) return format_doubles();
        if (pos.empty()) return fail(106, "path_to_input_db is required");
        pathToDatabase = pos[0];
        if (!is_file(pathToDatabase)) return fail(105, "path_to_input_db: File does not exist: " + pathToDatabase);
        if (pos.size() < 2) return fail(106, "path_to_output_file is required");
        pathToOutputFile = pos[1];
        if (pos.size() > 2) return fail(109, "The following argument was not expected: " + pos[2]);
        return -1;
    }

    // test hook: print every float64 of a raw little-endian file with the CSV cell formatter, one per line
    int format_doubles() const {
        std::ifstream in(formatDoublesFile, std::ios::binary);
        if (!in) return 105;
        double v;
        std::string out;
        while (in.read(reinterpret_cast<char *>(&v), sizeof(v))) { append_double(out, v); out += '\n'; }
        std::fwrite(out.data(), 1, out.size(), stdout);
        return 0;
    }

    // the argument box of main.cpp:89-113 (box-drawing characters are 3 bytes in UTF-8)
    void print() const {
        const std::string a[5] = {" Input Database  : " + pathToDatabase + " ", " Query Database  : " + pathToQryDatabase + " ",
                                  " Query Subset    : " + pathToQrySubsetFile + " ", " Output File     : " + pathToOutputFile + " ",
                                  " Field Separator : " + outFieldSeparator + " "};
        size_t w = 0;
        for (auto &s : a) w = std::max(w, s.size());
        std::string bar;
        for (size_t i = 0; i < w; i++) bar += "─";
        std::cout << " ┌" << bar << "┐\n";
        for (auto &s : a) std::cout << " │" << s << std::string(w - s.size(), ' ') << "│\n";
        std::cout << " └" << bar << "┘\n";
        std::cout.flush();
    }

    void load_query_genomes() {
        std::ifstream in(pathToQrySubsetFile);
        std::string s;
        while (in >> s) qryGenomeSet.push_back(s);
    }
};

int dump_index(const HostIndex &ix, const std::string &prefix) {
    auto put = [&](const std::string &suffix, const void *p, size_t bytes) {
        FILE *f = std::fopen((prefix + suffix).c_str(), "wb");
        if (!f) return false;
        const bool ok = bytes == 0 || std::fwrite(p, 1, bytes, f) == bytes;
        return (std::fclose(f) == 0) && ok;
    };
    const int32_t dims[4] = {ix.nProteins, ix.nGenomes, ix.nTargets, ix.nQueries};
    bool ok = put(".dims.bin", dims, sizeof(dims)) && put(".lc.bin", ix.Lc.data(), ix.Lc.size() * 4) &&
              put(".f.bin", ix.F.data(), ix.F.size() * sizeof(pfaai_pg_t)) && put(".t.bin", ix.T.data(), ix.T.size() * 4);
    return ok ? PFAAI_OK : PFAAI_ERR_CONSTRUCT;
}

void printOutput(const DataStructBase &pfdata, const ParFAAIImpl &impl, const AppParams &args) {
    std::printf("Writing output with %d query genomes and %d target genomes. \n", pfdata.qrySetSize(), pfdata.tgtSetSize());
    std::fflush(stdout);
    PhaseTimer tm;
    if (impl.streamed()) {
        // the rows were formatted and written by the writer thread while the GPU computed (pfaai_driver.hpp, csv_writer.hpp)
        if (!impl.streamOk()) std::cerr << "Error in writing " << args.pathToOutputFile << std::endl;
        std::printf("CSV output          : %10.3f ms; (streamed: %zu of %d rows were written while the tile kernel was still running)\n",
                    impl.streamTailMs(), impl.streamOverlapRows(), pfdata.qrySetSize());
        return;
    }
    if (!write_aji_csv(args.pathToOutputFile, args.outFieldSeparator, pfdata.refQuerySet(), pfdata.refTargetSet(), impl.matrixData()))
        std::cerr << "Error in writing " << args.pathToOutputFile << std::endl;
    std::printf("CSV output          : %10.3f ms;\n", tm.lap_ms());
}

template <typename DataT> int construct_run_print(DataT &data, const AppParams &args) {
    data.setDevices(args.device, args.nGpus, args.sameDevice);
    const pfaai_status rc = data.construct();    // (the reference tests the wrong variable here; we return the code)
    if (rc != PFAAI_OK) return rc;
    ParFAAIImpl impl(data);
    if (!args.pathToOutputFile.empty() && !args.noStreamCsv) impl.setStreamingOutput(args.pathToOutputFile, args.outFieldSeparator);
    const int rrc = impl.run();
    if (rrc != PFAAI_OK) return rrc;
#ifndef NDEBUG
    if (impl.haveLists()) impl.print_aji();
#endif
    if (!args.pathToOutputFile.empty()) printOutput(data, impl, args);
    return PFAAI_OK;
}

int missing_banner(const std::vector<std::string> &missing) {
    std::cout << "--------------------ERROR-----------------------------" << std::endl
              << " In the query subset file, the following genomes are " << std::endl
              << " missing from the database : " << std::endl;
    for (size_t i = 0; i < missing.size(); i++) std::cout << "    " << missing[i] << (i + 1 < missing.size() ? "\n" : "");
    std::cout << std::endl << std::endl
              << " Please remove them and before running Fast AAI. " << std::endl
              << "------------------------------------------------------" << std::endl;
    return PFAAI_ERR_CONSTRUCT;
}

int overlap_banner(const std::vector<std::string> &common) {
    std::cout << "---------------------ERROR------------------------------" << std::endl
              << "  Query database should have no intersecting genes with " << std::endl
              << "  the main data base." << std::endl << std::endl
              << "  In the query data base, the following genomes are " << std::endl
              << "  overlapping with the target database:" << std::endl;
    for (size_t i = 0; i < common.size(); i++) std::cout << "    " << common[i] << (i + 1 < common.size() ? "\n" : "");
    std::cout << std::endl << std::endl
              << "  Please remove them before running Fast AAI." << std::endl
              << "--------------------------------------------------------" << std::endl;
    return PFAAI_ERR_CONSTRUCT;
}

// the database object of the single-DB modes: plain SQLite, or SQLite behind the index cache (--index-cache)
std::unique_ptr<SQLiteSCPDataBase> open_main_db(const AppParams &args) {
    if (args.indexCachePath.empty()) return std::unique_ptr<SQLiteSCPDataBase>(new SQLiteSCPDataBase(args.pathToDatabase));
    return std::unique_ptr<SQLiteSCPDataBase>(new CachedSCPDataBase(args.pathToDatabase, args.indexCachePath));
}

int parallel_fastaai(const AppParams &args) {                       // main.cpp:177-202
    std::unique_ptr<SQLiteSCPDataBase> dbp = open_main_db(args);
    SQLiteSCPDataBase &db = *dbp;
    if (db.validate() != PFAAI_OK) return PFAAI_ERR_SQLITE_DB;
    if (!args.dumpIndexPrefix.empty()) {
        HostIndex ix;
        if (db.loadIndex(ix) != SQLITE_OK) return PFAAI_ERR_SQLITE_DB;
        return dump_index(ix, args.dumpIndexPrefix);
    }
    ParFAAIData data(db, db.getMeta());
    return construct_run_print(data, args);
}

int parallel_subset_fastaai(const AppParams &args) {                // main.cpp:234-266
    std::unique_ptr<SQLiteSCPDataBase> dbp = open_main_db(args);
    SQLiteSCPDataBase &db = *dbp;
    if (db.validate() != PFAAI_OK) return PFAAI_ERR_SQLITE_DB;
    const DBMetaData &meta = db.getMeta();
    std::unordered_set<std::string> have(meta.genomeSet.begin(), meta.genomeSet.end());
    std::vector<std::string> missing;
    for (auto &g : args.qryGenomeSet)
        if (!have.count(g)) missing.push_back(g);
    if (!missing.empty()) return missing_banner(missing);
    ParFAAIQSubData data(db, meta, args.qryGenomeSet);
    return construct_run_print(data, args);
}

int parallel_qry2tgt_fastaai(const AppParams &args) {               // main.cpp:302-335
    // with --index-cache: the main database comes from its cache, only the query database is read (index_cache.hpp)
    std::unique_ptr<QTSQLiteSCPDataBase> dbp(args.indexCachePath.empty() || !args.dumpIndexPrefix.empty()
                                                 ? new QTSQLiteSCPDataBase(args.pathToDatabase, args.pathToQryDatabase)
                                                 : new CachedQTDataBase(args.pathToDatabase, args.pathToQryDatabase, args.indexCachePath));
    QTSQLiteSCPDataBase &db = *dbp;
    if (db.validate() != PFAAI_OK) return PFAAI_ERR_SQLITE_DB;
    const DBMetaData &meta = db.getMeta();
    std::unordered_set<std::string> have(meta.genomeSet.begin(), meta.genomeSet.end());
    std::vector<std::string> common;
    for (auto &g : meta.qyGenomeSet)
        if (have.count(g)) common.push_back(g);
    if (!common.empty()) return overlap_banner(common);
    if (!args.dumpIndexPrefix.empty()) {
        HostIndex ix;
        if (db.loadIndex(ix) != SQLITE_OK) return PFAAI_ERR_SQLITE_DB;
        return dump_index(ix, args.dumpIndexPrefix);
    }
    ParFAAIQryTgtData data(db, meta, meta.proteinSet, args.refCompat);
    return construct_run_print(data, args);
}

}  // namespace

int main(int argc, char *argv[]) {
    AppParams args;
    const int prc = args.parse(argc, argv);
    if (prc >= 0) return prc;
    args.print();
    if (args.pathToQryDatabase.empty() || args.pathToQryDatabase == args.pathToDatabase) {
        if (args.pathToQrySubsetFile.empty()) return parallel_fastaai(args);
        args.load_query_genomes();
#ifndef NDEBUG
        std::cout << "Query Set: [" << args.qryGenomeSet.size() << ", (";
        for (size_t i = 0; i < args.qryGenomeSet.size(); i++) std::cout << (i ? ", " : "") << args.qryGenomeSet[i];
        std::cout << ")]" << std::endl;
#endif
        return parallel_subset_fastaai(args);
    }
    return parallel_qry2tgt_fastaai(args);
}
