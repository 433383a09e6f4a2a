/*
 * pfaai_oracle.c -- CPU restatement of the ParFastAAI AJI hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under oracle/ is part of the product: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it, and only
 * as the checker.  The product path (parfastaai_b200/csrc, the C-ABI in include/pfaai_gpu.h) never
 * links or calls it and has no CPU fallback.
 *
 * Parity: PINNED.  tests/test_oracle.py checks every function here against the reference's own
 * golden vectors (tests/golden/<set>.npz, re-packed from /root/reference/data/<set>_<x>.bin): Lc, Lp, F, T,
 * sorted E, JAC (S,N) and AJI for the xantho / subset1 / subset2 / query-subset / query-DB fixtures,
 * and against oracle/_ref (the reference itself, compiled from /root/reference by oracle/Makefile)
 * on synthetic databases.
 *
 * Each function cites the reference file:line (relative to /root/reference) whose behaviour it
 * restates.  Plain C99, scalar; pfo_direct_aji() alone uses OpenMP (it is the "strong CPU baseline"
 * and the large-N spot checker, and is an independent, definition-level formulation).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define PFO_NTETRA 160000 /* 20^4, include/pfaai/interface.hpp:233 */

typedef struct { int32_t protein, genome; } pfo_pg_t;               /* DPair<int,int>, utils.hpp:204-225 */
typedef struct { int32_t protein, gA, gB; } pfo_e_t;                 /* ETriple<int>, interface.hpp:92-121 */
typedef struct { int32_t gA, gB; double S; int32_t N; int32_t pad; } pfo_jac_t; /* JACTuple, interface.hpp:61-75 */

enum { PFO_ALL_VS_ALL = 0, PFO_QUERY_SUBSET = 1, PFO_QUERY_DB = 2 };
enum { PFO_OK = 0, PFO_ERR_ARG = 3 };

/* ------------------------------------------------------------------------------------------------
 * Mode semantics: the three DataStructInterface implementations (include/pfaai/ds_impl.hpp).
 *   all-vs-all   ParFAAIData        :39-151   ids 0..G-1, pairs a<b
 *   query subset ParFAAIQSubData    :159-337  qry_ids = ids of the query genomes in query-FILE order
 *   query DB     ParFAAIQryTgtData  :344-490  targets 0..nT-1, queries nT..nT+nQ-1 (G = nT+nQ)
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
    int mode, G, nQ, nT;
    uint8_t *is_qry;      /* m_qryIndicator */
    int32_t *index_map;   /* m_genomeIndexMap */
    int32_t *qry_lookup;  /* m_qryLookup  (subset) */
    int32_t *tgt_lookup;  /* m_tgtLookup  (subset) */
} pfo_mode_t;

static void mode_free(pfo_mode_t *m) {
    free(m->is_qry); free(m->index_map); free(m->qry_lookup); free(m->tgt_lookup);
    memset(m, 0, sizeof(*m));
}

static int mode_init(pfo_mode_t *m, int mode, int G, int nQ, const int32_t *qry_ids) {
    memset(m, 0, sizeof(*m));
    m->mode = mode; m->G = G;
    if (G <= 0) return PFO_ERR_ARG;
    m->is_qry = (uint8_t *)calloc((size_t)G, 1);
    m->index_map = (int32_t *)calloc((size_t)G, sizeof(int32_t));
    if (mode == PFO_ALL_VS_ALL) {                      /* ds_impl.hpp:76-90 */
        m->nQ = G; m->nT = G;
        for (int g = 0; g < G; g++) { m->is_qry[g] = 1; m->index_map[g] = g; }
    } else if (mode == PFO_QUERY_SUBSET) {             /* ds_impl.hpp:188-230 */
        if (nQ < 0 || nQ > G || (nQ > 0 && !qry_ids)) { mode_free(m); return PFO_ERR_ARG; }
        m->nQ = nQ; m->nT = G - nQ;
        m->qry_lookup = (int32_t *)calloc((size_t)(nQ > 0 ? nQ : 1), sizeof(int32_t));
        m->tgt_lookup = (int32_t *)calloc((size_t)(m->nT > 0 ? m->nT : 1), sizeof(int32_t));
        for (int q = 0; q < nQ; q++) {
            int g = qry_ids[q];
            if (g < 0 || g >= G || m->is_qry[g]) { mode_free(m); return PFO_ERR_ARG; }
            m->is_qry[g] = 1; m->qry_lookup[q] = g; m->index_map[g] = q;
        }
        for (int g = 0, j = 0; g < G; g++)
            if (!m->is_qry[g]) { m->index_map[g] = j; m->tgt_lookup[j] = g; j++; }
    } else if (mode == PFO_QUERY_DB) {                 /* ds_impl.hpp:372-392 */
        if (nQ <= 0 || nQ >= G) { mode_free(m); return PFO_ERR_ARG; }
        m->nQ = nQ; m->nT = G - nQ;
        for (int g = 0; g < m->nT; g++) { m->is_qry[g] = 0; m->index_map[g] = g; }
        for (int q = 0; q < nQ; q++) { m->is_qry[m->nT + q] = 1; m->index_map[m->nT + q] = q; }
    } else { mode_free(m); return PFO_ERR_ARG; }
    return PFO_OK;
}

static inline int64_t mode_npairs(const pfo_mode_t *m) {
    switch (m->mode) {
    case PFO_ALL_VS_ALL:   return (int64_t)m->G * (m->G - 1) / 2;                              /* :79-81 */
    case PFO_QUERY_SUBSET: return (int64_t)m->nQ * m->nT + (int64_t)m->nQ * (m->nQ - 1) / 2;   /* :244-249 */
    default:               return (int64_t)m->nQ * m->nT;                                      /* :403 */
    }
}

static inline int mode_valid_pair(const pfo_mode_t *m, int a, int b) {
    switch (m->mode) {
    case PFO_ALL_VS_ALL:   return a < b;                                                       /* :89-91 */
    case PFO_QUERY_SUBSET: return (m->is_qry[a] && m->is_qry[b] && a < b) ||
                                  (m->is_qry[a] && !m->is_qry[b] && a != b);                   /* :270-273 */
    default:               return m->is_qry[a] && !m->is_qry[b];                               /* :421-423 */
    }
}

static inline int64_t mode_pair_index(const pfo_mode_t *m, int a, int b) {
    if (m->mode == PFO_ALL_VS_ALL)                                                             /* :83-86 */
        return (int64_t)m->G * a + b - (int64_t)(a + 2) * (a + 1) / 2;
    if (m->mode == PFO_QUERY_SUBSET) {                                                         /* :251-263 */
        int64_t gia = m->index_map[a], gib = m->index_map[b];
        if (m->is_qry[a] && !m->is_qry[b]) return gia * m->nT + gib;
        return (int64_t)m->nQ * m->nT + ((int64_t)m->nQ * gia + gib - (gia + 2) * (gia + 1) / 2);
    }
    return (int64_t)m->index_map[a] * m->nT + m->index_map[b];                                 /* :411-413 */
}

int64_t pfo_n_pairs(int mode, int G, int nQ) {
    pfo_mode_t m; memset(&m, 0, sizeof(m));
    m.mode = mode; m.G = G;
    if (mode == PFO_ALL_VS_ALL) { m.nQ = G; m.nT = G; } else { m.nQ = nQ; m.nT = G - nQ; }
    return mode_npairs(&m);
}

/* initJAC of the three classes: pair labels in JAC order, S = 0, N = 0.
 * all: ds_impl.hpp:99-114; subset: :278-305; query DB: :426-438.  `qt_label_quirk` != 0 reproduces the
 * reference's query-DB labels (q, nQ + t) literally (they are what computeEBlockJAC uses for its T
 * lookups, algorithm_impl.hpp:225-226,252-254); 0 gives the intended ids (nT + q, t). */
int pfo_init_jac(int mode, int G, int nQ, const int32_t *qry_ids, int qt_label_quirk, pfo_jac_t *jac) {
    pfo_mode_t m; int rc = mode_init(&m, mode, G, nQ, qry_ids);
    if (rc) return rc;
    int64_t np = mode_npairs(&m), i = 0;
    if (mode == PFO_ALL_VS_ALL) {
        for (int a = 0; a < G; a++) for (int b = a + 1; b < G; b++, i++) { jac[i].gA = a; jac[i].gB = b; }
    } else if (mode == PFO_QUERY_SUBSET) {
        for (int q = 0; q < m.nQ; q++) for (int t = 0; t < m.nT; t++, i++) {
            jac[i].gA = m.qry_lookup[q]; jac[i].gB = m.tgt_lookup[t];
        }
        for (int a = 0; a < m.nQ; a++) for (int b = a + 1; b < m.nQ; b++, i++) {
            jac[i].gA = m.qry_lookup[a]; jac[i].gB = m.qry_lookup[b];
        }
    } else {
        for (int q = 0; q < m.nQ; q++) for (int t = 0; t < m.nT; t++, i++) {
            if (qt_label_quirk) { jac[i].gA = q; jac[i].gB = m.nQ + t; }
            else                { jac[i].gA = m.nT + q; jac[i].gB = t; }
        }
    }
    for (int64_t k = 0; k < np; k++) { jac[k].S = 0.0; jac[k].N = 0; jac[k].pad = 0; }
    mode_free(&m);
    return (i == np) ? PFO_OK : PFO_ERR_ARG;
}

/* ------------------------------------------------------------------------------------------------
 * Phase 1/2: Lc, Lp, F, T from unsorted postings (protein, tetramer, genome).
 *   Lc  DataStructHelper::constructLc       ds_helper.hpp:82-109  (Lc[t] = #postings of tetramer t)
 *   Lp  DataStructHelper::parallelPrefixSum ds_helper.hpp:112-122 (exclusive scan)
 *   F   DataStructHelper::constructF        ds_helper.hpp:126-162 (order: tetramer, protein, genome --
 *       the SQL "ORDER BY tetramer, source_table" of scp_db.hpp:182 plus ascending genome blobs)
 *   T   DataStructHelper::constructT        ds_helper.hpp:46-79   (T(p,g) = #tetramers of SCP p in genome g;
 *       equals the (p,g) histogram of F in the single-DB modes)
 * ---------------------------------------------------------------------------------------------- */
void pfo_prefix_sum(const int32_t *Lc, int32_t *Lp, int n) {
    int32_t run = 0;
    for (int i = 0; i < n; i++) { Lp[i] = run; run += Lc[i]; }
}

static int cmp_pg(const void *x, const void *y) {
    const pfo_pg_t *a = (const pfo_pg_t *)x, *b = (const pfo_pg_t *)y;
    if (a->protein != b->protein) return a->protein < b->protein ? -1 : 1;
    if (a->genome != b->genome) return a->genome < b->genome ? -1 : 1;
    return 0;
}

int pfo_build_index(int64_t n, const int32_t *protein, const int32_t *tetramer, const int32_t *genome,
                    int P, int G, int32_t *Lc, int32_t *Lp, pfo_pg_t *F, int32_t *T) {
    memset(Lc, 0, sizeof(int32_t) * PFO_NTETRA);
    for (int64_t i = 0; i < n; i++) {
        if (tetramer[i] < 0 || tetramer[i] >= PFO_NTETRA || protein[i] < 0 || protein[i] >= P ||
            genome[i] < 0 || genome[i] >= G) return PFO_ERR_ARG;
        Lc[tetramer[i]]++;
    }
    pfo_prefix_sum(Lc, Lp, PFO_NTETRA);
    int32_t *cur = (int32_t *)malloc(sizeof(int32_t) * PFO_NTETRA);
    memcpy(cur, Lp, sizeof(int32_t) * PFO_NTETRA);
    for (int64_t i = 0; i < n; i++) {
        pfo_pg_t e; e.protein = protein[i]; e.genome = genome[i];
        F[cur[tetramer[i]]++] = e;
    }
    free(cur);
    for (int t = 0; t < PFO_NTETRA; t++)
        if (Lc[t] > 1) qsort(F + Lp[t], (size_t)Lc[t], sizeof(pfo_pg_t), cmp_pg);
    if (T) {
        memset(T, 0, sizeof(int32_t) * (size_t)P * G);
        for (int64_t i = 0; i < n; i++) T[(size_t)F[i].protein * G + F[i].genome]++;
    }
    return PFO_OK;
}

/* ------------------------------------------------------------------------------------------------
 * Phase 2b: the E enumeration, restated as dense counting.
 *   constructTetramerTuples ds_helper.hpp:270-357: for every tetramer block of F, every maximal run of
 *   equal protein p inside it, every ordered (i, j) of the run with isQryGenome(g_i) and
 *   isValidPair(g_i, g_j) emits the triple (p, g_i, g_j).  After the sort (psort.hpp:27-53,
 *   ETriple::operator< interface.hpp:103-111) the run length of (gA, gB, p) in E is
 *   counts[pairIndex(gA,gB)][p] below; |E| = sum of counts (countTetramerTuples, ds_helper.hpp:206-265).
 * counts is [n_pairs x P] row-major int32, zeroed here.  Returns |E| through e_size.
 * ---------------------------------------------------------------------------------------------- */
int pfo_counts(int mode, int G, int nQ, const int32_t *qry_ids, int P, const int32_t *Lc,
               const pfo_pg_t *F, int64_t nF, int32_t *counts, int64_t *e_size) {
    pfo_mode_t m; int rc = mode_init(&m, mode, G, nQ, qry_ids);
    if (rc) return rc;
    int64_t np = mode_npairs(&m), total = 0, pos = 0;
    memset(counts, 0, sizeof(int32_t) * (size_t)np * P);
    for (int t = 0; t < PFO_NTETRA; t++) {
        int64_t end = pos + Lc[t];
        if (end > nF) { mode_free(&m); return PFO_ERR_ARG; }
        int64_t left = pos;
        while (left < end) {
            int p = F[left].protein;
            int64_t right = left;
            while (right < end && F[right].protein == p) right++;
            for (int64_t i = left; i < right; i++) {
                int a = F[i].genome;
                if (!m.is_qry[a]) continue;                        /* ds_helper.hpp:314,342 */
                for (int64_t j = left; j < right; j++) {
                    int b = F[j].genome;
                    if (!mode_valid_pair(&m, a, b)) continue;      /* ds_helper.hpp:319,347 */
                    int64_t idx = mode_pair_index(&m, a, b);
                    /* the reference's subset-mode formula assumes the query file lists genomes in
                     * ascending id order (ds_impl.hpp:251-263); anything else indexes out of range */
                    if (idx < 0 || idx >= np) { mode_free(&m); return PFO_ERR_ARG; }
                    counts[idx * P + p]++;
                    total++;
                }
            }
            left = right;
        }
        pos = end;
    }
    if (e_size) *e_size = total;
    mode_free(&m);
    return (pos == nF) ? PFO_OK : PFO_ERR_ARG;
}

/* The sorted E array itself, rebuilt from the counts: triples ordered by (gA, gB, protein)
 * (interface.hpp:103-111), each (pair, p) repeated counts times.  `E` must hold e_size triples. */
static int cmp_lab(const void *x, const void *y) {
    const int64_t *a = (const int64_t *)x, *b = (const int64_t *)y;
    return (a[0] < b[0]) ? -1 : (a[0] > b[0]);
}
int pfo_sorted_e(int mode, int G, int nQ, const int32_t *qry_ids, int P, const int32_t *counts, pfo_e_t *E) {
    pfo_mode_t m; int rc = mode_init(&m, mode, G, nQ, qry_ids);
    if (rc) return rc;
    int64_t np = mode_npairs(&m);
    pfo_jac_t *lab = (pfo_jac_t *)malloc(sizeof(pfo_jac_t) * (size_t)(np > 0 ? np : 1));
    pfo_init_jac(mode, G, nQ, qry_ids, 0, lab);     /* true genome ids, as stored in the triples */
    int64_t *key = (int64_t *)malloc(sizeof(int64_t) * 2 * (size_t)(np > 0 ? np : 1));
    for (int64_t i = 0; i < np; i++) { key[2 * i] = (int64_t)lab[i].gA * G + lab[i].gB; key[2 * i + 1] = i; }
    qsort(key, (size_t)np, 2 * sizeof(int64_t), cmp_lab);
    int64_t o = 0;
    for (int64_t k = 0; k < np; k++) {
        int64_t i = key[2 * k + 1];
        for (int p = 0; p < P; p++)
            for (int c = counts[i * P + p]; c > 0; c--) { E[o].protein = p; E[o].gA = lab[i].gA; E[o].gB = lab[i].gB; o++; }
    }
    free(key); free(lab); mode_free(&m);
    return PFO_OK;
}

/* ------------------------------------------------------------------------------------------------
 * Phase 3/4: JAC and AJI.
 *   computeEBlockJAC algorithm_impl.hpp:222-277: for each pair, for each protein run (ascending p) of
 *   length c >= 1: J = (double)c / (double)(T(p, gA) + T(p, gB) - c); S += J; N += 1, where (gA, gB) are
 *   the LABELS from initJAC.  computeAJI :309-322: AJI = S / N.
 *   Deviation (documented in DESIGN.md): a pair with no shared tetramer gets S = 0, N = 0,
 *   AJI = 0.0 instead of the reference's E[0]-derived garbage.
 * jac carries the labels on input (pfo_init_jac); T is [P x t_cols] row-major.
 * ---------------------------------------------------------------------------------------------- */
int pfo_jac_aji(int64_t n_pairs, int P, const int32_t *counts, const int32_t *T, int t_cols,
                pfo_jac_t *jac, double *aji) {
    for (int64_t i = 0; i < n_pairs; i++) {
        int a = jac[i].gA, b = jac[i].gB;
        if (a < 0 || a >= t_cols || b < 0 || b >= t_cols) return PFO_ERR_ARG;
        double S = 0.0; int N = 0;
        for (int p = 0; p < P; p++) {
            int c = counts[i * P + p];
            if (c <= 0) continue;
            double J = (double)c / (double)(T[(size_t)p * t_cols + a] + T[(size_t)p * t_cols + b] - c);
            S += J; N += 1;
        }
        jac[i].S = S; jac[i].N = N;
        if (aji) aji[i] = (N > 0) ? S / (double)N : 0.0;
    }
    return PFO_OK;
}

/* printOutput's matrix fill, src/main.cpp:143-154: M[mapQueryId(gA)][mapTargetId(gB)] = AJI, mirrored when
 * gB is also a query genome (modes 0 and 1); untouched cells (diagonal / self) stay 0.
 * `jac` must carry TRUE genome ids (pfo_init_jac with quirk = 0 gives the same pair order). M is [nQ x nCols]. */
int pfo_fill_matrix(int mode, int G, int nQ, const int32_t *qry_ids, const pfo_jac_t *jac, const double *aji,
                    double *M) {
    pfo_mode_t m; int rc = mode_init(&m, mode, G, nQ, qry_ids);
    if (rc) return rc;
    int64_t np = mode_npairs(&m);
    int rows = m.nQ, cols = (mode == PFO_QUERY_DB) ? m.nT : G;
    memset(M, 0, sizeof(double) * (size_t)rows * cols);
    for (int64_t i = 0; i < np; i++) {
        int a = jac[i].gA, b = jac[i].gB;
        int mq_a = m.index_map[a];                                       /* mapQueryId */
        int mt_b = (mode == PFO_QUERY_DB) ? m.index_map[b] : b;          /* mapTargetId */
        M[(size_t)mq_a * cols + mt_b] = aji[i];
        if (mode != PFO_QUERY_DB && m.is_qry[b]) M[(size_t)m.index_map[b] * cols + a] = aji[i];
    }
    mode_free(&m);
    return PFO_OK;
}

/* ------------------------------------------------------------------------------------------------
 * Independent, definition-level formulation (doc/pfaai_algorithm.tex:42-55 with the reference's N
 * semantics): for a rectangle rows x cols of genomes,
 *     c(p,a,b) = |Q(p,a) ^ Q(p,b)|  by merging the two sorted tetramer lists,
 *     S = sum_{p asc, c>0} c / (T(p,ta) + T(p,tb) - c),  N = #{p : c>0},  AJI = S/N (0 if N == 0).
 * Input is the genome-major index: list (g,p) = vals[off[g*P+p] .. off[g*P+p+1]) ascending.
 * row_t / col_t give the T column used for each row / col (identity except the query-DB label quirk).
 * Cells with rows[i] == cols[j] are written as 0 (self).  OpenMP over rows: this is the "strong CPU
 * baseline (ours)" of BASELINE.md section 3 and the spot checker for N the reference cannot reach.
 * Any of S, N, aji, counts may be NULL.  counts is [n_rows x n_cols x P].
 * ---------------------------------------------------------------------------------------------- */
int pfo_direct_aji(int P, int G, const int64_t *off, const int32_t *vals, const int32_t *T, int t_cols,
                   int n_rows, const int32_t *rows, const int32_t *row_t,
                   int n_cols, const int32_t *cols, const int32_t *col_t,
                   double *S_out, int32_t *N_out, double *aji, int32_t *counts) {
    for (int i = 0; i < n_rows; i++) if (rows[i] < 0 || rows[i] >= G) return PFO_ERR_ARG;
    for (int j = 0; j < n_cols; j++) if (cols[j] < 0 || cols[j] >= G) return PFO_ERR_ARG;
#pragma omp parallel for schedule(dynamic, 1)
    for (int i = 0; i < n_rows; i++) {
        int a = rows[i], ta = row_t ? row_t[i] : a;
        for (int j = 0; j < n_cols; j++) {
            int b = cols[j], tb = col_t ? col_t[j] : b;
            size_t cell = (size_t)i * n_cols + j;
            double S = 0.0; int N = 0;
            for (int p = 0; p < P; p++) {
                int c = 0;
                if (a != b) {
                    const int32_t *x = vals + off[(size_t)a * P + p], *xe = vals + off[(size_t)a * P + p + 1];
                    const int32_t *y = vals + off[(size_t)b * P + p], *ye = vals + off[(size_t)b * P + p + 1];
                    while (x < xe && y < ye) {
                        int32_t u = *x, v = *y;
                        c += (u == v); x += (u <= v); y += (v <= u);
                    }
                }
                if (counts) counts[cell * P + p] = c;
                if (c > 0) {
                    S += (double)c / (double)(T[(size_t)p * t_cols + ta] + T[(size_t)p * t_cols + tb] - c);
                    N += 1;
                }
            }
            if (S_out) S_out[cell] = S;
            if (N_out) N_out[cell] = N;
            if (aji) aji[cell] = (N > 0) ? S / (double)N : 0.0;
        }
    }
    return PFO_OK;
}

/* Genome-major index from F + Lc (what the device index build must reproduce): off is [G*P+1] int64,
 * vals is [nF] int32 tetramer ids; list (g,p) ascending in tetramer because F is tetramer-major. */
int pfo_genome_major(int P, int G, const int32_t *Lc, const pfo_pg_t *F, int64_t nF, int64_t *off, int32_t *vals) {
    size_t nseg = (size_t)G * P;
    memset(off, 0, sizeof(int64_t) * (nseg + 1));
    for (int64_t i = 0; i < nF; i++) {
        if (F[i].protein < 0 || F[i].protein >= P || F[i].genome < 0 || F[i].genome >= G) return PFO_ERR_ARG;
        off[(size_t)F[i].genome * P + F[i].protein + 1]++;
    }
    for (size_t s = 0; s < nseg; s++) off[s + 1] += off[s];
    int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (nseg > 0 ? nseg : 1));
    memcpy(cur, off, sizeof(int64_t) * nseg);
    int64_t pos = 0;
    for (int t = 0; t < PFO_NTETRA; t++)
        for (int k = 0; k < Lc[t]; k++, pos++) {
            if (pos >= nF) { free(cur); return PFO_ERR_ARG; }
            vals[cur[(size_t)F[pos].genome * P + F[pos].protein]++] = t;
        }
    free(cur);
    return (pos == nF) ? PFO_OK : PFO_ERR_ARG;
}

/* Checker for the tile kernel's reciprocal-table division (parfastaai_b200/csrc/aji_tile.cuh, jaccard_exact):
 * counts the (c, den), 0 <= c <= den <= den_max, for which  q' = fma(fma(-q, den, c), r, q)  with r = RN(1/den),
 * q = RN(c*r)  differs from the IEEE quotient the reference computes (algorithm_impl.hpp:250-254).  Must be 0. */
#include <math.h>
int64_t pfo_check_table_division(int den_max) {
    int64_t bad = 0;
#pragma omp parallel for schedule(dynamic, 64) reduction(+ : bad)
    for (int den = 1; den <= den_max; den++) {
        const double dd = (double)den, r = 1.0 / dd;
        for (int c = 0; c <= den; c++) {
            const double cd = (double)c;
            volatile double q = cd * r;            /* volatile: no contraction of the product into the fma below */
            const double rem = fma(-q, dd, cd);
            const double q2 = fma(rem, r, q);
            if (q2 != cd / dd) bad++;
        }
    }
    return bad;
}
