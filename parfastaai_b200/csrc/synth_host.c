/*
 * synth_host.c -- host-side TOOLING: synthetic FastAAI-like databases (SURVEY 8(d) recipe) and the
 * conversion of genome-major tetramer lists into the reference's arrays (Lc, F).  Used by bench.py and
 * the tests to make inputs; it is not on the GPU path and computes no AJI.
 *
 * Recipe (BASELINE configs 4/5: N genomes x P SCPs, ~300-aa proteins):
 *   per SCP p : length L = clip(trunc(Normal(300, 40)), 80, 600); ancestral sequence uniform over 20 aa
 *   clades    : K = max(1, N/16); genome -> clade uniform; clade divergence d_k ~ U(0.05, 0.35),
 *               genome divergence e_g ~ U(0, 0.03)   ("dense" family: K = 4, d_k ~ U(0, 0.05), e_g ~ U(0, 0.01))
 *   sequences : clade = ancestor with each site resampled (uniform aa) w.p. d_k; genome = clade with each
 *               site resampled w.p. e_g
 *   tetramers : unique of the L-3 overlapping words, id = a0*8000 + a1*400 + a2*20 + a3 in [0, 160000)
 * Every random draw is a counter-based hash of (seed, stream, indices), so the output is independent of
 * the thread count and any (genome, SCP) list can be regenerated on its own.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define NTETRA 160000
typedef struct { int32_t protein, genome; } pg_t;

static inline uint64_t mix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
static inline uint64_t h4(uint64_t seed, uint64_t stream, uint64_t a, uint64_t b, uint64_t c) {
    uint64_t x = mix64(seed ^ (stream * 0xD6E8FEB86659FD93ull));
    x = mix64(x ^ a); x = mix64(x ^ (b + 0x632BE59BD9B4E019ull)); x = mix64(x ^ (c + 0xA0761D6478BD642Full));
    return x;
}
static inline double u01(uint64_t h) { return (double)(h >> 11) * (1.0 / 9007199254740992.0); }

enum { S_LEN = 1, S_ANC, S_CLADE_OF, S_DK, S_EG, S_CMUT, S_CAA, S_GMUT, S_GAA };

int pfs_protein_length(uint64_t seed, int p) {
    /* Box-Muller */
    double u1 = u01(h4(seed, S_LEN, (uint64_t)p, 0, 0)), u2 = u01(h4(seed, S_LEN, (uint64_t)p, 1, 0));
    if (u1 < 1e-300) u1 = 1e-300;
    double z = sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
    int L = (int)(300.0 + 40.0 * z);
    if (L < 80) L = 80;
    if (L > 600) L = 600;
    return L;
}

static int cmp_i32(const void *a, const void *b) {
    int32_t x = *(const int32_t *)a, y = *(const int32_t *)b;
    return (x > y) - (x < y);
}

/* sorted unique tetramers of (genome g, SCP p) into out (capacity >= L-3); returns the count */
static int gen_list(uint64_t seed, int N, int dense, int p, int g, int32_t *out) {
    const int L = pfs_protein_length(seed, p);
    const int K = dense ? 4 : (N / 16 > 1 ? N / 16 : 1);
    const int k = (int)(h4(seed, S_CLADE_OF, (uint64_t)g, 0, 0) % (uint64_t)K);
    const double dk = dense ? 0.05 * u01(h4(seed, S_DK, (uint64_t)k, (uint64_t)p, 0))
                            : 0.05 + 0.30 * u01(h4(seed, S_DK, (uint64_t)k, (uint64_t)p, 0));
    const double eg = (dense ? 0.01 : 0.03) * u01(h4(seed, S_EG, (uint64_t)g, (uint64_t)p, 0));
    uint8_t aa[600];
    for (int s = 0; s < L; s++) {
        int a = (int)(h4(seed, S_ANC, (uint64_t)p, (uint64_t)s, 0) % 20);
        if (u01(h4(seed, S_CMUT, (uint64_t)p, (uint64_t)k, (uint64_t)s)) < dk)
            a = (int)(h4(seed, S_CAA, (uint64_t)p, (uint64_t)k, (uint64_t)s) % 20);
        if (u01(h4(seed, S_GMUT, (uint64_t)p, (uint64_t)g, (uint64_t)s)) < eg)
            a = (int)(h4(seed, S_GAA, (uint64_t)p, (uint64_t)g, (uint64_t)s) % 20);
        aa[s] = (uint8_t)a;
    }
    const int n = L - 3;
    for (int s = 0; s < n; s++) out[s] = aa[s] * 8000 + aa[s + 1] * 400 + aa[s + 2] * 20 + aa[s + 3];
    qsort(out, (size_t)n, sizeof(int32_t), cmp_i32);
    int m = 0;
    for (int s = 0; s < n; s++) if (s == 0 || out[s] != out[s - 1]) out[m++] = out[s];
    return m;
}

/* Pass 1: seg_off[g*P+p+1] = list length (then prefix-summed in place); returns |F| */
int64_t pfs_count(int N, int P, uint64_t seed, int dense, int64_t *seg_off /*[N*P+1]*/) {
    seg_off[0] = 0;
#pragma omp parallel for schedule(dynamic, 4)
    for (int g = 0; g < N; g++) {
        int32_t buf[600];
        for (int p = 0; p < P; p++) seg_off[(size_t)g * P + p + 1] = gen_list(seed, N, dense, p, g, buf);
    }
    for (size_t s = 0; s < (size_t)N * P; s++) seg_off[s + 1] += seg_off[s];
    return seg_off[(size_t)N * P];
}

/* Pass 2: the lists themselves, genome-major (segment g*P+p), ascending tetramers */
void pfs_fill(int N, int P, uint64_t seed, int dense, const int64_t *seg_off, int32_t *vals) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int g = 0; g < N; g++) {
        int32_t buf[600];
        for (int p = 0; p < P; p++) {
            int m = gen_list(seed, N, dense, p, g, buf);
            memcpy(vals + seg_off[(size_t)g * P + p], buf, sizeof(int32_t) * (size_t)m);
        }
    }
}

/* The same two passes for a SUBSET of the genomes of the N-genome database (ids[k] in [0, N)): list (k, p) of the result is list
 * (ids[k], p) of the full database.  Lets a checker rebuild a few rows / columns of a large synthetic workload on its own. */
int64_t pfs_count_subset(int N, int P, uint64_t seed, int dense, const int32_t *ids, int n, int64_t *seg_off /*[n*P+1]*/) {
    seg_off[0] = 0;
#pragma omp parallel for schedule(dynamic, 4)
    for (int k = 0; k < n; k++) {
        int32_t buf[600];
        for (int p = 0; p < P; p++) seg_off[(size_t)k * P + p + 1] = gen_list(seed, N, dense, p, ids[k], buf);
    }
    for (size_t s = 0; s < (size_t)n * P; s++) seg_off[s + 1] += seg_off[s];
    return seg_off[(size_t)n * P];
}
void pfs_fill_subset(int N, int P, uint64_t seed, int dense, const int32_t *ids, int n, const int64_t *seg_off, int32_t *vals) {
#pragma omp parallel for schedule(dynamic, 4)
    for (int k = 0; k < n; k++) {
        int32_t buf[600];
        for (int p = 0; p < P; p++) {
            int m = gen_list(seed, N, dense, p, ids[k], buf);
            memcpy(vals + seg_off[(size_t)k * P + p], buf, sizeof(int32_t) * (size_t)m);
        }
    }
}

/* genome-major lists -> the reference's arrays: Lc[160000] and F in (tetramer, protein, genome) order
 * (what DataBaseInterface::tetramerOccCounts / proteinSetGPPairs deliver, scp_db.hpp:121-216). */
void pfs_to_reference(int N, int P, const int64_t *seg_off, const int32_t *vals, int32_t *Lc, pg_t *F) {
    const int64_t nF = seg_off[(size_t)N * P];
    memset(Lc, 0, sizeof(int32_t) * NTETRA);
    for (int64_t i = 0; i < nF; i++) Lc[vals[i]]++;
    int64_t *Lp = (int64_t *)malloc(sizeof(int64_t) * (NTETRA + 1));
    Lp[0] = 0;
    for (int t = 0; t < NTETRA; t++) Lp[t + 1] = Lp[t] + Lc[t];
    /* each thread owns a tetramer range and walks (p outer, g inner) so that its cursors are private */
    const int NR = 256;
#pragma omp parallel for schedule(dynamic, 1)
    for (int rgn = 0; rgn < NR; rgn++) {
        /* ranges balanced by posting count */
        const int64_t lo_cnt = nF * rgn / NR, hi_cnt = nF * (rgn + 1) / NR;
        int t0 = 0, t1 = 0;
        { int a = 0, b = NTETRA; while (a < b) { int m = (a + b) >> 1; if (Lp[m] < lo_cnt) a = m + 1; else b = m; } t0 = a; }
        { int a = 0, b = NTETRA; while (a < b) { int m = (a + b) >> 1; if (Lp[m] < hi_cnt) a = m + 1; else b = m; } t1 = a; }
        if (rgn == NR - 1) t1 = NTETRA;
        if (t0 >= t1) continue;
        int64_t *cur = (int64_t *)malloc(sizeof(int64_t) * (size_t)(t1 - t0));
        for (int t = t0; t < t1; t++) cur[t - t0] = Lp[t];
        for (int p = 0; p < P; p++)
            for (int g = 0; g < N; g++) {
                const int32_t *x = vals + seg_off[(size_t)g * P + p], *xe = vals + seg_off[(size_t)g * P + p + 1];
                /* first element >= t0 */
                const int32_t *a = x, *b = xe;
                while (a < b) { const int32_t *m = a + (b - a) / 2; if (*m < t0) a = m + 1; else b = m; }
                for (; a < xe && *a < t1; a++) { pg_t e; e.protein = p; e.genome = g; F[cur[*a - t0]++] = e; }
            }
        free(cur);
    }
    free(Lp);
}
