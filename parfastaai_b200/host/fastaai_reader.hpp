// fastaai_reader.hpp -- host-side FastAAI SQLite ingest for the B200 AJI path (C++17, header-only).
//
// "SQLite/tetramer loading stays on the host" (BASELINE north_star).  The two classes keep the names and
// the call surface of the reference's storage adapters (paths relative to the reference repo):
//     SQLiteSCPDataBase     include/pfaai/scp_db.hpp:59-263     single DB (all-vs-all, query-subset)
//     QTSQLiteSCPDataBase   include/pfaai/scp_db.hpp:268-590    main DB + attached query DB (-r)
//     SQLiteHelper          include/pfaai/db_helper.hpp:34-218  metadata queries
// and deliver the very arrays the reference's DataBaseInterface virtuals deliver piecewise
// (tetramerOccCounts -> Lc, proteinSetGPPairs -> F, proteinTetramerCounts -> T; interface.hpp:153-196)
// -- but in ONE sequential pass per `<SCP>_tetras` table (one read-only connection per OpenMP thread)
// instead of the reference's per-thread P-way UNION ALL range scans.
//
// Output layout (what pfaai_gpu_upload_index consumes):
//     Lc [160000]  postings per tetramer
//     F  [sum Lc]  (protein, genome) ordered by (tetramer, protein, genome-as-stored-in-the-blob);
//                  query-DB mode: inner join on the tetramer per SCP, target genomes first, then
//                  query genomes renumbered nT + id   (scp_db.hpp:450-528)
//     T  [P x G]   row-major, T(p,g) = length(tetramers blob of (p,g)) / 4, unrestricted (scp_db.hpp:219-262, 531-589)
// Protein index = first-appearance order of `SELECT DISTINCT scp_acc` (same SQL text as the reference so
// that the query planner returns the same order; tests pin it).
#pragma once
#include <omp.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include "../../include/pfaai_gpu.h"
#include "sqlite3_abi.h"

namespace pfaai_host {

constexpr int NTETRAMERS = PFAAI_NTETRAMERS;

// == DBMetaData (interface.hpp:145-149)
struct DBMetaData {
    std::vector<std::string> proteinSet;
    std::vector<std::string> genomeSet;
    std::vector<std::string> qyGenomeSet;
};

// == DatabaseNames (scp_db.hpp:37-55): only the names the path reads
struct DatabaseNames {
    const std::string TMTAB_SUFFIX = "_tetras";
    const std::string TMTAB_COLUMN_TT = "tetramer";
    const std::string TMTAB_COLUMN_GM = "genomes";
    const std::string GNMTAB_SUFFIX = "_genomes";
    const std::string GNMTAB_COLUMN_GID = "genome_id";
    const std::string GNMTAB_COLUMN_TMS = "tetramers";
    const std::string GMTTAB = "genome_metadata";
    const std::string GMTTAB_COLUMN_GNAME = "genome_name";
    const std::string SCPDTAB = "scp_data";
    const std::string SCPDTAB_COLUMN_ACC = "scp_acc";
};

// The device-ready host arrays of one run.
struct HostIndex {
    int nProteins = 0;
    int nGenomes = 0;          // total genome ids (query-DB mode: nT + nQ)
    int nTargets = 0, nQueries = 0;
    std::vector<int32_t> Lc;   // [160000]
    std::vector<pfaai_pg_t> F; // [sum Lc]
    std::vector<int32_t> T;    // [P x nGenomes]
    // device-ready form (index cache, pfaai_gpu_import_lists): genome-major lists; when present F may be empty
    std::vector<int64_t> segOff;    // [P*nGenomes + 1]
    std::vector<uint32_t> segVals;  // [sum Lc]
    bool haveLists() const { return !segOff.empty(); }
};

namespace detail {

inline int query_strings(sqlite3 *db, const std::string &sql, std::vector<std::string> &out) {
    sqlite3_stmt *st = nullptr;
    int rc = sqlite3_prepare_v2(db, sql.c_str(), -1, &st, nullptr);
    if (rc != SQLITE_OK) {
        std::cerr << "Error in preparing sql statement " << sql << std::endl;
        std::cerr << "The error was: " << sqlite3_errmsg(db);
        return rc;
    }
    out.clear();
    while (sqlite3_step(st) == SQLITE_ROW) {
        const unsigned char *txt = sqlite3_column_text(st, 0);
        out.emplace_back(txt ? reinterpret_cast<const char *>(txt) : "");
    }
    sqlite3_finalize(st);
    return SQLITE_OK;
}

// one `<SCP>_tetras` table: rows in ascending tetramer order
struct TetraTable {
    std::vector<int32_t> tetramer;   // [rows]
    std::vector<int64_t> start;      // [rows + 1] offsets into genomes
    std::vector<int32_t> genomes;    // concatenated blobs (little-endian int32 genome ids)
};

inline int read_tetras(sqlite3 *db, const std::string &qualified_table, const DatabaseNames &n, TetraTable &t) {
    const std::string sql = "SELECT " + n.TMTAB_COLUMN_TT + ", " + n.TMTAB_COLUMN_GM + " FROM " + qualified_table +
                            " ORDER BY " + n.TMTAB_COLUMN_TT;
    sqlite3_stmt *st = nullptr;
    int rc = sqlite3_prepare_v2(db, sql.c_str(), -1, &st, nullptr);
    if (rc != SQLITE_OK) {
        std::cerr << "Error in preparing sql statement " << sql << std::endl;
        std::cerr << "The error was: " << sqlite3_errmsg(db) << std::endl;
        return rc;
    }
    t.tetramer.clear(); t.start.assign(1, 0); t.genomes.clear();
    while ((rc = sqlite3_step(st)) == SQLITE_ROW) {
        const int tet = sqlite3_column_int(st, 0);
        const void *blob = sqlite3_column_blob(st, 1);
        const int n_ids = sqlite3_column_bytes(st, 1) / 4;
        if (tet < 0 || tet >= NTETRAMERS) continue;   // outside the 20^4 universe: never indexed by the reference
        const size_t old = t.genomes.size();
        t.genomes.resize(old + n_ids);
        if (n_ids) std::memcpy(t.genomes.data() + old, blob, sizeof(int32_t) * (size_t)n_ids);
        t.tetramer.push_back(tet);
        t.start.push_back((int64_t)t.genomes.size());
    }
    sqlite3_finalize(st);
    return rc == SQLITE_DONE ? SQLITE_OK : rc;
}

// one `<SCP>_genomes` table -> T(p, id_offset + genome_id) = bytes / 4
inline int read_tcounts(sqlite3 *db, const std::string &qualified_table, const DatabaseNames &n, int32_t *Trow, int id_offset,
                        int n_cols) {
    const std::string sql = "SELECT " + n.GNMTAB_COLUMN_GID + ", length(" + n.GNMTAB_COLUMN_TMS + ") FROM " + qualified_table;
    sqlite3_stmt *st = nullptr;
    int rc = sqlite3_prepare_v2(db, sql.c_str(), -1, &st, nullptr);
    if (rc != SQLITE_OK) {
        std::cerr << "Error in preparing sql statement " << sql << std::endl;
        std::cerr << "The error was: " << sqlite3_errmsg(db) << std::endl;
        return rc;
    }
    while ((rc = sqlite3_step(st)) == SQLITE_ROW) {
        const int gid = sqlite3_column_int(st, 0) + id_offset;
        if (gid >= 0 && gid < n_cols) Trow[gid] = sqlite3_column_int(st, 1) / 4;
    }
    sqlite3_finalize(st);
    return rc == SQLITE_DONE ? SQLITE_OK : rc;
}

// Lc and F from the per-protein posting tables (already restricted / renumbered by the caller).
inline void assemble(const std::vector<TetraTable> &tabs, HostIndex &ix) {
    const int P = (int)tabs.size();
    ix.Lc.assign(NTETRAMERS, 0);
    for (int p = 0; p < P; p++) {
        const TetraTable &t = tabs[p];
        for (size_t r = 0; r < t.tetramer.size(); r++) ix.Lc[t.tetramer[r]] += (int32_t)(t.start[r + 1] - t.start[r]);
    }
    std::vector<int64_t> Lp(NTETRAMERS + 1, 0);
    for (int t = 0; t < NTETRAMERS; t++) Lp[t + 1] = Lp[t] + ix.Lc[t];
    ix.F.resize((size_t)Lp[NTETRAMERS]);
    // every thread owns a contiguous tetramer range and walks the proteins in index order inside it:
    // F comes out ordered (tetramer, protein, blob order) without atomics
    const int nthreads = std::max(1, omp_get_max_threads());
#pragma omp parallel for schedule(static, 1)
    for (int th = 0; th < nthreads; th++) {
        const int t_lo = (int)((int64_t)NTETRAMERS * th / nthreads), t_hi = (int)((int64_t)NTETRAMERS * (th + 1) / nthreads);
        std::vector<int64_t> cur(Lp.begin() + t_lo, Lp.begin() + t_hi);
        for (int p = 0; p < P; p++) {
            const TetraTable &t = tabs[p];
            size_t r = std::lower_bound(t.tetramer.begin(), t.tetramer.end(), t_lo) - t.tetramer.begin();
            for (; r < t.tetramer.size() && t.tetramer[r] < t_hi; r++) {
                int64_t &c = cur[t.tetramer[r] - t_lo];
                for (int64_t k = t.start[r]; k < t.start[r + 1]; k++) {
                    ix.F[(size_t)c].protein = p;
                    ix.F[(size_t)c].genome = t.genomes[(size_t)k];
                    c++;
                }
            }
        }
    }
}

}  // namespace detail

// ---------------------------------------------------------------------------------------------------
// Single database: query set == target set (all-vs-all and query-subset modes)
// ---------------------------------------------------------------------------------------------------
class SQLiteSCPDataBase {
  protected:
    std::string m_pathToDb;
    sqlite3 *m_sqltDbPtr = nullptr;
    int m_dbErrorCode = SQLITE_OK;
    DatabaseNames m_dbNames;
    DBMetaData m_dbMeta;

    static sqlite3 *open_ro(const std::string &path, int *rc) {
        sqlite3 *db = nullptr;
        *rc = sqlite3_open_v2(path.c_str(), &db, SQLITE_OPEN_READONLY | SQLITE_OPEN_NOMUTEX, nullptr);
        return db;
    }

    struct DeferOpen {};
    SQLiteSCPDataBase(const std::string &dbPath, DeferOpen) : m_pathToDb(dbPath) {}   // a subclass decides whether SQLite is needed

    void openAndReadNames() {
        m_sqltDbPtr = open_ro(m_pathToDb, &m_dbErrorCode);
        if (m_dbErrorCode != SQLITE_OK) return;
        // same statements as db_helper.hpp:169-218 / 59-107 (protein order = first appearance)
        m_dbErrorCode = detail::query_strings(
            m_sqltDbPtr, "SELECT DISTINCT " + m_dbNames.SCPDTAB_COLUMN_ACC + " FROM " + m_dbNames.SCPDTAB, m_dbMeta.proteinSet);
        if (m_dbErrorCode == SQLITE_OK)
            m_dbErrorCode = detail::query_strings(
                m_sqltDbPtr, "SELECT " + m_dbNames.GMTTAB_COLUMN_GNAME + " FROM " + m_dbNames.GMTTAB, m_dbMeta.genomeSet);
    }

  public:
    explicit SQLiteSCPDataBase(const std::string &dbPath) : m_pathToDb(dbPath) { openAndReadNames(); }
    SQLiteSCPDataBase(const SQLiteSCPDataBase &) = delete;
    SQLiteSCPDataBase &operator=(const SQLiteSCPDataBase &) = delete;
    virtual ~SQLiteSCPDataBase() { if (m_sqltDbPtr) closeDB(); }

    int getDBErrorCode() const { return m_dbErrorCode; }
    std::string getDBPath() const { return m_pathToDb; }
    const DBMetaData &getMeta() const { return m_dbMeta; }
    // device-ready lists cache (index_cache.hpp): the driver announces whether it can take the genome-major lists instead of F
    // (single-context or replicated builds) and hands the lists of a fresh device build back for the next run
    virtual void preferLists(bool) const {}
    virtual bool wantsLists() const { return false; }
    virtual bool storeLists(const HostIndex &, const int64_t *, const uint32_t *) const { return false; }
    virtual void dropLists() const {}     // the lists just served were refused by the device: forget them, serve F from now on

    int closeDB() {
        if (m_sqltDbPtr == nullptr) {
            std::cerr << "Database already closed : " << m_pathToDb << std::endl;
            return -1;
        }
        sqlite3_close(m_sqltDbPtr);
        m_sqltDbPtr = nullptr;
        m_dbErrorCode = SQLITE_OK;
        return SQLITE_OK;
    }

    // == validate() (scp_db.hpp:111-118)
    virtual pfaai_status validate() const {
        if (m_dbErrorCode != SQLITE_OK || m_sqltDbPtr == nullptr) {
            std::cerr << "Error in opening " << getDBPath() << std::endl;
            if (m_dbErrorCode != SQLITE_OK) std::cerr << "DB ERROR: " << sqlite3_errstr(m_dbErrorCode) << std::endl;
            return PFAAI_ERR_SQLITE_DB;
        }
        return PFAAI_OK;
    }

    // Lc, F, T of the whole database (B.6: subset mode indexes the whole DB as well).  Returns an SQLite rc.
    virtual int loadIndex(HostIndex &ix) const {
        const int P = (int)m_dbMeta.proteinSet.size(), G = (int)m_dbMeta.genomeSet.size();
        ix.nProteins = P; ix.nGenomes = G; ix.nTargets = G; ix.nQueries = G;
        ix.T.assign((size_t)P * G, 0);
        std::vector<detail::TetraTable> tabs(P);
        int rc_all = SQLITE_OK;
#pragma omp parallel
        {
            int rc = SQLITE_OK;
            sqlite3 *db = open_ro(m_pathToDb, &rc);   // one connection per thread: no shared-handle serialisation
#pragma omp for schedule(dynamic, 1)
            for (int p = 0; p < P; p++) {
                if (rc != SQLITE_OK) continue;
                const std::string &acc = m_dbMeta.proteinSet[p];
                rc = detail::read_tetras(db, "`" + acc + m_dbNames.TMTAB_SUFFIX + "`", m_dbNames, tabs[p]);
                if (rc == SQLITE_OK)
                    rc = detail::read_tcounts(db, "`" + acc + m_dbNames.GNMTAB_SUFFIX + "`", m_dbNames, ix.T.data() + (size_t)p * G, 0, G);
            }
            if (db) sqlite3_close(db);
            if (rc != SQLITE_OK) {
#pragma omp critical
                rc_all = rc;
            }
        }
        if (rc_all != SQLITE_OK) return rc_all;
        detail::assemble(tabs, ix);
        return SQLITE_OK;
    }
};

// ---------------------------------------------------------------------------------------------------
// Main (target) database + attached query database (-r)
// ---------------------------------------------------------------------------------------------------
class QTSQLiteSCPDataBase {
  protected:
    std::string m_pathToTgtDb, m_pathToQryDb;
    sqlite3 *m_sqltDbPtr = nullptr;
    int m_dbErrorCode = SQLITE_OK;
    DatabaseNames m_dbNames;
    DBMetaData m_dbMeta;

    sqlite3 *open_pair(int *rc, bool verbose) const {
        sqlite3 *db = nullptr;
        *rc = sqlite3_open_v2(m_pathToTgtDb.c_str(), &db, SQLITE_OPEN_READONLY | SQLITE_OPEN_NOMUTEX, nullptr);
        if (*rc != SQLITE_OK) return db;
        std::string quoted;
        for (char ch : m_pathToQryDb) { quoted += ch; if (ch == '\'') quoted += '\''; }
        const std::string attach = "ATTACH DATABASE '" + quoted + "' as QueryDB ;";     // scp_db.hpp:343-346
        *rc = sqlite3_exec(db, attach.c_str(), nullptr, nullptr, nullptr);
        if (*rc != SQLITE_OK && verbose) {
            std::cerr << "Error in attaching query database : " << m_pathToQryDb << std::endl;
            std::cerr << "SQL Error: " << sqlite3_errmsg(db);
        }
        return db;
    }

    struct DeferOpen {};
    QTSQLiteSCPDataBase(const std::string &tgtDBPath, const std::string &qryDBPath, DeferOpen)   // a subclass decides whether the pair is opened
        : m_pathToTgtDb(tgtDBPath), m_pathToQryDb(qryDBPath) {}
    // SCPs present in BOTH databases, same statement as db_helper.hpp:139-146.  The ORDER of its rows is whatever SQLite's
    // planner makes of the two-table join, and that order is the fp64 summation order of S: anything that wants the
    // reference's bytes takes the order from this statement, not from either database alone.
    std::string commonScpStatement() const {
        const std::string acc = m_dbNames.SCPDTAB_COLUMN_ACC;
        return "SELECT DISTINCT target_table." + acc + " \n  FROM `main`." + m_dbNames.SCPDTAB + " as target_table, `QueryDB`." +
               m_dbNames.SCPDTAB + " as query_table \n  WHERE target_table." + acc + " = query_table." + acc + ";";
    }
    void openPairAndReadNames() {
        m_dbMeta = DBMetaData();
        m_sqltDbPtr = open_pair(&m_dbErrorCode, true);
        if (m_dbErrorCode != SQLITE_OK) return;
        m_dbErrorCode = detail::query_strings(m_sqltDbPtr, commonScpStatement(), m_dbMeta.proteinSet);
        if (m_dbErrorCode == SQLITE_OK)
            m_dbErrorCode = detail::query_strings(
                m_sqltDbPtr, "SELECT " + m_dbNames.GMTTAB_COLUMN_GNAME + " FROM `main`." + m_dbNames.GMTTAB, m_dbMeta.genomeSet);
        if (m_dbErrorCode == SQLITE_OK)
            m_dbErrorCode = detail::query_strings(
                m_sqltDbPtr, "SELECT " + m_dbNames.GMTTAB_COLUMN_GNAME + " FROM `QueryDB`." + m_dbNames.GMTTAB, m_dbMeta.qyGenomeSet);
    }

  public:
    QTSQLiteSCPDataBase(const std::string &tgtDBPath, const std::string &qryDBPath)
        : m_pathToTgtDb(tgtDBPath), m_pathToQryDb(qryDBPath) {
        openPairAndReadNames();
    }
    QTSQLiteSCPDataBase(const QTSQLiteSCPDataBase &) = delete;
    QTSQLiteSCPDataBase &operator=(const QTSQLiteSCPDataBase &) = delete;
    virtual ~QTSQLiteSCPDataBase() { if (m_sqltDbPtr) closeDB(); }

    int getDBErrorCode() const { return m_dbErrorCode; }
    std::string getDBPath() const { return m_pathToTgtDb; }
    const DBMetaData &getMeta() const { return m_dbMeta; }

    int closeDB() {
        if (m_sqltDbPtr == nullptr) {
            std::cerr << "Database already closed : " << m_pathToQryDb << std::endl;
            return m_dbErrorCode;
        }
        sqlite3_close(m_sqltDbPtr);   // closing the connection detaches QueryDB
        m_sqltDbPtr = nullptr;
        m_dbErrorCode = SQLITE_OK;
        return SQLITE_OK;
    }

    virtual pfaai_status validate() const {
        if (m_dbErrorCode != SQLITE_OK || m_sqltDbPtr == nullptr) {
            std::cerr << "Error in opening " << getDBPath() << std::endl;
            if (m_dbErrorCode != SQLITE_OK) std::cerr << "DB ERROR: " << sqlite3_errstr(m_dbErrorCode) << std::endl;
            return PFAAI_ERR_SQLITE_DB;
        }
        return PFAAI_OK;
    }

    // targets keep ids 0..nT-1, queries become nT + id; per SCP only tetramers present in BOTH databases
    // are indexed (inner join, scp_db.hpp:408-414,459-466), T is unrestricted.
    // the index as the reference's own route delivers it (ATTACH + joins), even when a subclass can serve it from a cache
    virtual int loadIndexUncached(HostIndex &ix) const { return loadIndex(ix); }
    virtual int loadIndex(HostIndex &ix) const {
        const int P = (int)m_dbMeta.proteinSet.size();
        const int nT = (int)m_dbMeta.genomeSet.size(), nQ = (int)m_dbMeta.qyGenomeSet.size(), G = nT + nQ;
        ix.nProteins = P; ix.nGenomes = G; ix.nTargets = nT; ix.nQueries = nQ;
        ix.T.assign((size_t)P * G, 0);
        std::vector<detail::TetraTable> tabs(P);
        int rc_all = SQLITE_OK;
#pragma omp parallel
        {
            int rc = SQLITE_OK;
            sqlite3 *db = open_pair(&rc, false);
            detail::TetraTable tt, qt;
#pragma omp for schedule(dynamic, 1)
            for (int p = 0; p < P; p++) {
                if (rc != SQLITE_OK) continue;
                const std::string &acc = m_dbMeta.proteinSet[p];
                rc = detail::read_tetras(db, "main.`" + acc + m_dbNames.TMTAB_SUFFIX + "`", m_dbNames, tt);
                if (rc == SQLITE_OK) rc = detail::read_tetras(db, "QueryDB.`" + acc + m_dbNames.TMTAB_SUFFIX + "`", m_dbNames, qt);
                if (rc != SQLITE_OK) continue;
                detail::TetraTable &out = tabs[p];
                out.start.assign(1, 0);
                size_t a = 0, b = 0;
                while (a < tt.tetramer.size() && b < qt.tetramer.size()) {   // sorted merge = inner join on the tetramer
                    if (tt.tetramer[a] < qt.tetramer[b]) a++;
                    else if (tt.tetramer[a] > qt.tetramer[b]) b++;
                    else {
                        out.tetramer.push_back(tt.tetramer[a]);
                        out.genomes.insert(out.genomes.end(), tt.genomes.begin() + tt.start[a], tt.genomes.begin() + tt.start[a + 1]);
                        for (int64_t k = qt.start[b]; k < qt.start[b + 1]; k++) out.genomes.push_back(qt.genomes[(size_t)k] + nT);
                        out.start.push_back((int64_t)out.genomes.size());
                        a++; b++;
                    }
                }
                int32_t *Trow = ix.T.data() + (size_t)p * G;
                rc = detail::read_tcounts(db, "main.`" + acc + m_dbNames.GNMTAB_SUFFIX + "`", m_dbNames, Trow, 0, nT);
                if (rc == SQLITE_OK)
                    rc = detail::read_tcounts(db, "QueryDB.`" + acc + m_dbNames.GNMTAB_SUFFIX + "`", m_dbNames, Trow, nT, G);
            }
            if (db) sqlite3_close(db);
            if (rc != SQLITE_OK) {
#pragma omp critical
                rc_all = rc;
            }
        }
        if (rc_all != SQLITE_OK) return rc_all;
        detail::assemble(tabs, ix);
        return SQLITE_OK;
    }
};

}  // namespace pfaai_host
