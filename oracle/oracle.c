/* oracle.c -- TEST INFRASTRUCTURE ONLY.  NOT part of the product.
 *
 * A plain-C, sequential restatement of the BFQzip hot path:
 *   Part A  reads -> eBWT / permuted qualities / LCP        (gsufsort semantics)
 *   Part B  positional clusters, quality smoothing, noise reduction (bfq_ext semantics)
 *   Part C  inverse LF walk -> reads in input order           (BCRdecode semantics)
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this file's shared object.  The CUDA product (bfqzip_b200/) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below against the
 * unmodified reference binaries built by oracle/Makefile into oracle/_ref/ (run in this
 * container) and against the golden vectors those binaries produced (tests/golden/).
 *
 * Every function cites the reference file:line it restates (paths under /root/reference).
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>

typedef struct {
    int32_t K;          /* bfq_ext -k, default 16   (bfq_ext.cpp:69,1026) */
    int32_t m;          /* bfq_ext -m, default 2    (bfq_ext.cpp:73,1027) */
    int32_t qthr;       /* bfq_ext -t, default 20   (bfq_ext.cpp:81,1030) */
    int32_t max_len;    /* bfq_ext -l               (bfq_ext.cpp:986)      */
    int32_t M;          /* compile-time mode 0..3   (bfq_ext.cpp:643-654)  */
    int32_t B;          /* Illumina binning 0/1     (parameters.h:41-45)   */
    float   fthr;       /* bfq_ext -f, default 40.0 (bfq_ext.cpp:85,1031)  */
    uint8_t eos;        /* bfq_ext -s, default '#'  (bfq_ext.cpp:96,1005)  */
    uint8_t default_q;  /* bfq_ext -v, default '>'  (bfq_ext.cpp:77,1028)  */
} oracle_params;

typedef struct {
    uint64_t clusters;      /* tot_clust            bfq_ext.cpp:1114 */
    uint64_t discarded;     /* num_clust_discarded  bfq_ext.cpp:1115 */
    uint64_t ambiguous;     /* num_clust_amb_discarded :1116 */
    uint64_t processed;     /* num_clust_mod        :1117 */
    uint64_t bases_inside;  /* :1118 */
    uint64_t qs_smoothed;   /* :1121 */
    uint64_t modified;      /* :1126 */
    uint64_t nreads;        /* numSeq, bfq_ext.cpp:153-154 */
} oracle_stats;

/* ------------------------------------------------------------------------------------
 * Part A.  external/gsufsort/lib/utils.c:193-249 (cat_all), external/gsacak.c:113-133
 * (separators ordered by text position), main.c:848-867,891-895 (bwt / bwt.qs / lcp).
 *
 * Suffix (d,o), o = 0..len_d.  Order: lexicographic up to the end-marker, end-marker
 * smaller than every base, ties (identical strings) by read index d.
 * ---------------------------------------------------------------------------------- */
typedef struct { uint32_t d; uint32_t o; } sfx_t;

static const uint8_t  *g_bases;
static const uint64_t *g_off;

static int sfx_cmp(const void *a, const void *b)
{
    const sfx_t *x = (const sfx_t *)a, *y = (const sfx_t *)b;
    const uint8_t *s = g_bases + g_off[x->d] + x->o;
    const uint8_t *t = g_bases + g_off[y->d] + y->o;
    uint64_t ls = g_off[x->d + 1] - g_off[x->d] - x->o;
    uint64_t lt = g_off[y->d + 1] - g_off[y->d] - y->o;
    uint64_t l = ls < lt ? ls : lt;
    int c = memcmp(s, t, l);
    if (c) return c;
    if (ls != lt) return ls < lt ? -1 : 1;       /* end-marker < every base        */
    if (x->d != y->d) return x->d < y->d ? -1 : 1; /* equal strings: by read index  */
    return 0;
}

/* bases/quals: concatenated reads without separators; off[nreads+1]; outputs N bytes each,
 * N = sum(len_d + 1).  Returns N.  lcp is written as the low byte (main.c:891-895). */
uint64_t oracle_build_ebwt(const uint8_t *bases, const uint8_t *quals, const uint64_t *off,
                           uint64_t nreads, uint8_t eos,
                           uint8_t *bwt, uint8_t *qs, uint8_t *lcp)
{
    uint64_t N = off[nreads] + nreads, k = 0;
    sfx_t *sa = (sfx_t *)malloc((N ? N : 1) * sizeof(sfx_t));
    for (uint64_t d = 0; d < nreads; d++)
        for (uint64_t o = 0; o <= off[d + 1] - off[d]; o++) {
            sa[k].d = (uint32_t)d; sa[k].o = (uint32_t)o; k++;
        }
    g_bases = bases; g_off = off;
    qsort(sa, N, sizeof(sfx_t), sfx_cmp);
    for (uint64_t i = 0; i < N; i++) {
        uint64_t p = off[sa[i].d] + sa[i].o;
        if (sa[i].o == 0) { bwt[i] = eos; qs[i] = eos; }      /* main.c:853,863 */
        else { bwt[i] = bases[p - 1]; qs[i] = quals[p - 1]; }
        if (i == 0) { lcp[i] = 0; continue; }
        uint64_t q = off[sa[i - 1].d] + sa[i - 1].o;
        uint64_t lp = off[sa[i].d + 1] - p, lq = off[sa[i - 1].d + 1] - q;
        uint64_t l = lp < lq ? lp : lq, h = 0;
        while (h < l && bases[p + h] == bases[q + h]) h++;
        lcp[i] = (uint8_t)h;
    }
    free(sa);
    return N;
}

/* ------------------------------------------------------------------------------------
 * Part B.  src_ext_mem/bfq_ext.cpp.
 * ---------------------------------------------------------------------------------- */

/* decode.cpp:72-84 */
static uint8_t illumina_8_level_binning(uint8_t q)
{
    if (q >= 40) q = 40;
    else if (q >= 35) q = 37;
    else if (q >= 30) q = 33;
    else if (q >= 25) q = 27;
    else if (q >= 20) q = 22;
    else if (q >= 10) q = 15;
    else if (q >= 2) q = 6;
    return (uint8_t)(q + 33);
}
uint8_t oracle_binning(uint8_t ascii_q) { return illumina_8_level_binning((uint8_t)(ascii_q - 33)); }

/* bfq_ext.cpp:502-518 */
static uint8_t max_qs(const uint8_t *b, const uint8_t *q, uint64_t len, uint8_t eos)
{
    uint8_t mx = 0;
    for (uint64_t j = 0; j < len; j++) if (b[j] != eos && q[j] > mx) mx = q[j];
    return mx;
}
/* bfq_ext.cpp:522-543: double, summed in row order */
static uint8_t mean_error(const uint8_t *b, const uint8_t *q, uint64_t len, uint8_t eos)
{
    double sum_err = 0.0; uint64_t num = 0;
    for (uint64_t j = 0; j < len; j++)
        if (b[j] != eos) { num++; sum_err = sum_err + pow(10, -((double)q[j] - 33.0) / 10.0); }
    double avg_err = sum_err / num;
    uint8_t v = (uint8_t)round(-10 * log10(avg_err));
    return (uint8_t)(v + 33);
}
/* bfq_ext.cpp:479-498: int sum, float division */
static uint8_t avg_qs(const uint8_t *b, const uint8_t *q, uint64_t len, uint8_t eos)
{
    int sum = 0; uint64_t num = 0;
    for (uint64_t j = 0; j < len; j++) if (b[j] != eos) { sum = sum + q[j]; num++; }
    if (sum == 0) return 0;
    return (uint8_t)round((float)sum / num);
}
/* exposed so tests can pin the floating-point helpers on hand-made clusters */
uint8_t oracle_newqs(int M, const uint8_t *b, const uint8_t *q, uint64_t len, uint8_t eos, uint8_t dflt)
{
    switch (M) {
    case 0: return max_qs(b, q, len, eos);
    case 1: return mean_error(b, q, len, eos);
    case 2: return dflt;
    default: return avg_qs(b, q, len, eos);
    }
}

typedef struct {
    const uint8_t *bwt; const uint8_t *qs_in; uint64_t n;
    uint8_t *qs; uint8_t *obwt;
    const uint64_t *LF;            /* LF[i] = C[bwt[i]] + rank_{bwt[i]}(i), bfq_ext.cpp:415-476 */
    const oracle_params *P; oracle_stats *S;
    int err;
} smooth_ctx;

/* bfq_ext.cpp:546-573 */
static void mod_bases_smooth_qs(smooth_ctx *c, uint64_t start, uint64_t len, uint8_t sym,
                                uint8_t newqs, const uint8_t *hiq)
{
    const uint8_t eos = c->P->eos;
    for (uint64_t j = start; j < start + len; j++) {
        uint8_t b = c->bwt[j];
        if (b == eos) continue;
        if (b != sym && !hiq[b]) { c->obwt[j] = sym; c->S->modified++; }
        else if (b == sym) { c->qs[j] = newqs; c->S->qs_smoothed++; }
        else if (newqs < c->qs[j]) { c->qs[j] = newqs; c->S->qs_smoothed++; }
    }
}

/* bfq_ext.cpp:581-824.  Returns 1 when the cluster counts towards tot_clust. */
static int process_cluster(smooth_ctx *c, uint64_t start, uint64_t len)
{
    const oracle_params *P = c->P; const uint8_t eos = P->eos;
    if ((int)len < P->m) return 0;                                    /* :584 */
    uint64_t freqs[256]; uint8_t hiq[256];
    memset(freqs, 0, sizeof freqs); memset(hiq, 0, sizeof hiq);
    uint64_t base_num = 0;
    const uint8_t *b = c->bwt + start, *q = c->qs_in + start;
    for (uint64_t j = 0; j < len; j++)                                 /* :612-627 */
        if (b[j] != eos) {
            freqs[b[j]]++; base_num++;
            if (q[j] >= P->qthr + 33) hiq[b[j]] = 1;
        }
    if (base_num == 0) return 0;                                       /* :633 */
    c->S->bases_inside += base_num;
    uint8_t newqs = oracle_newqs(P->M, b, q, len, eos, P->default_q);  /* :643-654 */

    uint8_t F[256]; int nF = 0;                                        /* :665-677, alphabet order */
    for (int s = 0; s < 256; s++)
        if (freqs[s] > 0) {
            uint8_t perc = (uint8_t)((100 * freqs[s]) / base_num);
            if (perc >= P->fthr) F[nF++] = (uint8_t)s;
        }
    if (nF >= 3) { c->err = 1; return 1; }                             /* assert :683 */

    if (nF == 0) c->S->discarded++;
    else if (nF == 1) {
        if (F[0] == 'N') c->S->discarded++;
        else { mod_bases_smooth_qs(c, start, len, F[0], newqs, hiq); c->S->processed++; }
    }
    else if (base_num < (unsigned)P->m) c->S->discarded++;             /* :697 */
    else if (F[0] == 'N') { mod_bases_smooth_qs(c, start, len, F[1], newqs, hiq); c->S->processed++; }
    else if (F[1] == 'N') { mod_bases_smooth_qs(c, start, len, F[0], newqs, hiq); c->S->processed++; }
    else {                                                             /* :713-802 */
        uint8_t prec0 = 0, prec1 = 0; int seen0[256] = {0}, seen1[256] = {0};
        for (uint64_t j = 0; j < len; j++) {
            uint8_t lf = c->bwt[c->LF[start + j]];
            if (b[j] == F[0] && lf != eos && lf != 'N') { seen0[lf] = 1; prec0 = lf; }
            if (b[j] == F[1] && lf != eos && lf != 'N') { seen1[lf] = 1; prec1 = lf; }
        }
        /* :744-747 sums the A,C,G,T slots only */
        int n0 = seen0['A'] + seen0['C'] + seen0['G'] + seen0['T'];
        int n1 = seen1['A'] + seen1['C'] + seen1['G'] + seen1['T'];
        if (n0 == 1 && n1 == 1 && prec0 != prec1) {
            c->S->processed++;
            for (uint64_t j = 0; j < len; j++) {
                uint64_t r = start + j;
                if (b[j] == eos) continue;
                if (b[j] != F[0] && b[j] != F[1] && !hiq[b[j]]) {
                    uint8_t lf = c->bwt[c->LF[r]];
                    if (lf == prec0) { c->obwt[r] = F[0]; c->S->modified++; }
                    else if (lf == prec1) { c->obwt[r] = F[1]; c->S->modified++; }
                }
                else if (b[j] == F[0] || b[j] == F[1]) { c->qs[r] = newqs; c->S->qs_smoothed++; }
                else if (newqs < c->qs[r]) { c->qs[r] = newqs; c->S->qs_smoothed++; }
            }
        } else c->S->ambiguous++;
    }
    return 1;
}

/* LF over the ORIGINAL bwt: pile starts as buildFreq (bfq_ext.cpp:132-187, symbols in byte
 * order), rank as findBlockToRead + in-block count (decode.cpp:353-366, 801-837, 898-903). */
static uint64_t *build_lf(const uint8_t *bwt, uint64_t n)
{
    uint64_t cnt[256] = {0}, C[256], run[256] = {0}, acc = 0;
    for (uint64_t i = 0; i < n; i++) cnt[bwt[i]]++;
    for (int s = 0; s < 256; s++) { C[s] = acc; acc += cnt[s]; }
    uint64_t *LF = (uint64_t *)malloc((n ? n : 1) * sizeof(uint64_t));
    for (uint64_t i = 0; i < n; i++) { LF[i] = C[bwt[i]] + run[bwt[i]]; run[bwt[i]]++; }
    return LF;
}

/* buildBitVectors (bfq_ext.cpp:350-412) + run (bfq_ext.cpp:837-948).
 * Outputs: qs_out = qualities after smoothing (no binning), obwt = output bases
 * (bwt with the noise-reduced rows replaced).  Returns 0, or 1 if the reference's
 * assert(FreqSymb.size() < 3) would have fired. */
int oracle_smooth(const uint8_t *bwt, const uint8_t *qs, const uint8_t *lcp, uint64_t n,
                  const oracle_params *P, uint8_t *qs_out, uint8_t *obwt, oracle_stats *S)
{
    memset(S, 0, sizeof *S);
    memcpy(qs_out, qs, n); memcpy(obwt, bwt, n);
    for (uint64_t i = 0; i < n; i++) if (bwt[i] == P->eos) S->nreads++;
    uint8_t *thr = (uint8_t *)calloc(n + 1, 1), *mn = (uint8_t *)calloc(n + 1, 1);
    {   /* the `decreasing` state machine, literally */
        int decreasing = 0; uint8_t prev = 0;
        for (uint64_t i = 0; i < n; i++) {
            if (lcp[i] >= P->K) thr[i] = 1;
            if (decreasing && prev <= lcp[i]) { decreasing = 0; mn[i - 1] = 1; }
            else if (prev > lcp[i]) decreasing = 1;
            prev = lcp[i];
        }
    }
    uint64_t *LF = build_lf(bwt, n);
    smooth_ctx c = { bwt, qs, n, qs_out, obwt, LF, P, S, 0 };
    int open = 0; uint64_t begin = 0, len = 0;
    for (uint64_t i = 1; i < n; i++) {
        if (thr[i] && !mn[i]) {
            if (open) len++;
            else { open = 1; begin = i - 1; len = 2; }
        } else {
            if (open) S->clusters += process_cluster(&c, begin, len);
            open = 0; len = 0;
        }
    }
    if (open) S->clusters += process_cluster(&c, begin, len);
    free(LF); free(thr); free(mn);
    return c.err;
}

/* ------------------------------------------------------------------------------------
 * Part C.  decode.cpp:499-966 (inverse LF walk, <= max_len rounds) and :968-1152
 * (strip padding, reverse, emit).  Read g starts at row g (posN = g+1, decode.cpp:578-582).
 * out_off[nreads+1]; out_bases/out_quals sized >= n - nreads.  B=1 bins every emitted
 * quality (decode.cpp:879-883).
 * ---------------------------------------------------------------------------------- */
uint64_t oracle_rebuild(const uint8_t *bwt, const uint8_t *obwt, const uint8_t *qs, uint64_t n,
                        const oracle_params *P, uint64_t nreads,
                        uint8_t *out_bases, uint8_t *out_quals, uint64_t *out_off)
{
    uint64_t *LF = build_lf(bwt, n);
    uint8_t *tb = (uint8_t *)malloc(P->max_len + 1), *tq = (uint8_t *)malloc(P->max_len + 1);
    uint64_t w = 0;
    for (uint64_t g = 0; g < nreads; g++) {
        uint64_t row = g; int k = 0;
        out_off[g] = w;
        while (k < P->max_len) {
            uint8_t c = bwt[row];
            if (c == P->eos) break;
            tb[k] = obwt[row];
            tq[k] = P->B ? illumina_8_level_binning((uint8_t)(qs[row] - 33)) : qs[row];
            k++;
            row = LF[row];
        }
        for (int j = k - 1; j >= 0; j--) { out_bases[w] = tb[j]; out_quals[w] = tq[j]; w++; }
    }
    out_off[nreads] = w;
    free(LF); free(tb); free(tq);
    return w;
}

/* whole path in one call (what bench.py's cpu_baseline "port" leg times) */
int oracle_pipeline(const uint8_t *bases, const uint8_t *quals, const uint64_t *off, uint64_t nreads,
                    const oracle_params *P, uint8_t *bwt, uint8_t *qs, uint8_t *lcp,
                    uint8_t *out_bases, uint8_t *out_quals, uint64_t *out_off, oracle_stats *S)
{
    uint64_t n = oracle_build_ebwt(bases, quals, off, nreads, P->eos, bwt, qs, lcp);
    uint8_t *sq = (uint8_t *)malloc(n ? n : 1), *ob = (uint8_t *)malloc(n ? n : 1);
    int err = oracle_smooth(bwt, qs, lcp, n, P, sq, ob, S);
    oracle_rebuild(bwt, ob, sq, n, P, nreads, out_bases, out_quals, out_off);
    free(sq); free(ob);
    return err;
}
