/*
 * oracle/pairhmm_oracle.c -- TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the reference's PairHMM forward algorithm (src/main/native/mgl_pairhmm/):
 *   Context.h:13-82,95-160       quality tables: ph2pr, matchToMatchProb through the "Jacobian log" table,
 *                                INITIAL_CONSTANT 2^120 (float) / 2^1020 (double)
 *   compute_prob_scalar.cc:48-83 per-read transition probabilities pMM, pGapM, pMX, pMY, pZZ and Distm
 *   compute_prob_scalar.cc:5-44  one cell: M, Y (same row, previous column), X (previous row, same column)
 *   pairhmm_common.h:131-169     the same recurrence written row by row (the reference's own debug restatement);
 *                                this file follows that order -- the operations per cell and the left-to-right
 *                                order of the final sum (compute_prob_scalar.cc:213,310) are the same
 *   ...MicrosoftPairHmm.cc:77-222  buffers, float first, double when the float result < MIN_ACCEPTED (1e-28f,
 *                                pairhmm_common.h:31), result = log10(sum) - log10(INITIAL_CONSTANT)
 *
 * Only tests/, __graft_entry__.smoke() and the CPU-baseline legs of the benches may load this.  Pinned by
 * tests/test_pairhmm_oracle.py against the reference's own fixture (pairhmm-testdata.txt, 104 cases at 1e-5,
 * MicrosoftPairHmmUnitTest.java:101) and against outputs of the reference itself (tests/golden/pairhmm_golden.json.gz).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#define MAX_QUAL 254                               /* Context.h:7 */
#define JAC_TOL 8.0                                /* Context.h:8 */
#define JAC_STEP 0.0001                            /* Context.h:9 */
#define JAC_INV_STEP (1.0 / JAC_STEP)
#define JAC_SIZE ((int)(JAC_TOL / JAC_STEP) + 1)
#define MM_SIZE (((MAX_QUAL + 1) * (MAX_QUAL + 2)) >> 1)
#define MIN_ACCEPTED 1e-28f                        /* pairhmm_common.h:31 */

/* ------------------------------------------------------------------ one precision, generated twice */
#define DEFINE_PRECISION(NUMBER, SFX, INIT_EXPR, LOG10_INIT_EXPR, PH2PR_EXPR)                                        \
    static NUMBER jac_##SFX[JAC_SIZE];                                                                                 \
    static NUMBER mm_##SFX[MM_SIZE];                                                                                   \
    static NUMBER ph2pr_##SFX[128];                                                                                    \
    static NUMBER initial_##SFX, log10_initial_##SFX;                                                                  \
    static int ready_##SFX = 0;                                                                                        \
                                                                                                                       \
    /* Context.h:60-62 */                                                                                              \
    static int fast_round_##SFX(NUMBER d) { return d > (NUMBER)0.0 ? (int)(d + (NUMBER)0.5) : (int)(d - (NUMBER)0.5); } \
                                                                                                                       \
    /* Context.h:63-79 */                                                                                              \
    static NUMBER approx_log10_sum_##SFX(NUMBER small, NUMBER big)                                                     \
    {                                                                                                                  \
        if (small > big) { NUMBER t = big; big = small; small = t; }                                                   \
        if (isinf(small) || isinf(big)) return big;                                                                    \
        const NUMBER diff = big - small;                                                                               \
        if (diff >= (NUMBER)JAC_TOL) return big;                                                                       \
        const int ind = fast_round_##SFX((NUMBER)(diff * (NUMBER)JAC_INV_STEP));                                       \
        return big + jac_##SFX[ind];                                                                                   \
    }                                                                                                                  \
                                                                                                                       \
    /* Context.h:41-58 (tables), :95-110 / :131-146 (the two constructors) */                                          \
    static void init_##SFX(void)                                                                                       \
    {                                                                                                                  \
        if (ready_##SFX) return;                                                                                       \
        for (int k = 0; k < JAC_SIZE; k++) jac_##SFX[k] = (NUMBER)log10(1.0 + pow(10.0, -((double)k) * JAC_STEP));    \
        const double inv_ln10 = 1.0 / log(10);                                                                         \
        for (int i = 0, offset = 0; i <= MAX_QUAL; offset += ++i)                                                      \
            for (int j = 0; j <= i; j++) {                                                                             \
                const double log10_sum = approx_log10_sum_##SFX(-0.1f * i, -0.1f * j);                                 \
                const double m2m_log10 = log1p(-fmin(1.0, pow(10, log10_sum))) * inv_ln10;                             \
                mm_##SFX[offset + j] = (NUMBER)pow(10, m2m_log10);                                                     \
            }                                                                                                          \
        for (int x = 0; x < 128; x++) ph2pr_##SFX[x] = PH2PR_EXPR;                                                     \
        initial_##SFX = INIT_EXPR;                                                                                     \
        log10_initial_##SFX = LOG10_INIT_EXPR;                                                                         \
        ready_##SFX = 1;                                                                                               \
    }                                                                                                                  \
                                                                                                                       \
    /* Context.h:118-128 / :154-164; qualities are masked to 0..127 by the caller, so the table always applies */     \
    static NUMBER mm_prob_##SFX(int ins_qual, int del_qual)                                                            \
    {                                                                                                                  \
        int lo = del_qual, hi = ins_qual;                                                                              \
        if (ins_qual <= del_qual) { lo = ins_qual; hi = del_qual; }                                                    \
        return mm_##SFX[((hi * (hi + 1)) >> 1) + lo];                                                                  \
    }                                                                                                                  \
                                                                                                                       \
    /* one (read, haplotype) pair: the sum over the last row of M + X (compute_prob_scalar.cc:213,310,                 \
       pairhmm_common.h:131-169); returns -1 on allocation failure through *ok */                                      \
    static NUMBER forward_##SFX(int rl, const uint8_t *rs, const uint8_t *q, const uint8_t *ins, const uint8_t *del,   \
                                const uint8_t *gcp, int hl, const uint8_t *hap, int *ok)                               \
    {                                                                                                                  \
        const size_t cols = (size_t)hl + 1;                                                                            \
        NUMBER *buf = (NUMBER *)malloc(sizeof(NUMBER) * cols * 6);                                                     \
        if (!buf) { *ok = 0; return (NUMBER)0; }                                                                       \
        NUMBER *Mp = buf, *Xp = buf + cols, *Yp = buf + 2 * cols, *Mc = buf + 3 * cols, *Xc = buf + 4 * cols,          \
               *Yc = buf + 5 * cols;                                                                                   \
        const NUMBER three_over = (NUMBER)1.0 / (NUMBER)3.0;                                                           \
        const NUMBER y_initial = initial_##SFX / (NUMBER)hl;              /* compute_prob_scalar.cc:104 */             \
        for (size_t c = 0; c < cols; c++) { Mp[c] = (NUMBER)0; Xp[c] = (NUMBER)0; Yp[c] = y_initial; }                 \
        for (int r = 1; r <= rl; r++) {                                                                                \
            const int qi = ins[r - 1] & 127, qd = del[r - 1] & 127, qc = gcp[r - 1] & 127;                             \
            const NUMBER pMM = mm_prob_##SFX(qi, qd);                      /* compute_prob_scalar.cc:73-81 */          \
            const NUMBER pGapM = (NUMBER)1.0 - ph2pr_##SFX[qc];                                                        \
            const NUMBER pMX = ph2pr_##SFX[qi], pMY = ph2pr_##SFX[qd], pZZ = ph2pr_##SFX[qc];                          \
            const NUMBER dq = ph2pr_##SFX[q[r - 1] & 127];                                                             \
            const uint8_t b = rs[r - 1];                                                                               \
            Mc[0] = (NUMBER)0; Xc[0] = (NUMBER)0; Yc[0] = (NUMBER)0;      /* compute_prob_scalar.cc:140-142 */         \
            for (size_t c = 1; c < cols; c++) {                                                                        \
                const uint8_t h = hap[c - 1];                                                                          \
                NUMBER distm = dq;                                         /* compute_prob_scalar.cc:25-34 */          \
                if (b == h || b == 'N' || h == 'N') distm = (NUMBER)1.0 - distm;                                       \
                else distm = distm * three_over;                                                                       \
                Mc[c] = distm * (Mp[c - 1] * pMM + (Xp[c - 1] + Yp[c - 1]) * pGapM);          /* :39 */                \
                Yc[c] = Mc[c - 1] * pMY + Yc[c - 1] * pZZ;                                      /* :41 */                \
                Xc[c] = Mp[c] * pMX + Xp[c] * pZZ;                                              /* :43 */                \
            }                                                                                                          \
            NUMBER *t;                                                                                                 \
            t = Mp; Mp = Mc; Mc = t; t = Xp; Xp = Xc; Xc = t; t = Yp; Yp = Yc; Yc = t;                                 \
        }                                                                                                              \
        NUMBER result = (NUMBER)0;                                                                                     \
        for (size_t c = 1; c < cols; c++) result += Mp[c] + Xp[c];                                                     \
        free(buf);                                                                                                     \
        return result;                                                                                                 \
    }

DEFINE_PRECISION(float, f, ldexpf(1.f, 120), log10f(ldexpf(1.f, 120)), powf(10.f, -((float)x) / 10.f))
DEFINE_PRECISION(double, d, ldexp(1.0, 1020), log10(ldexp(1.0, 1020)), pow(10.0, -((double)x) / 10.0))

/*
 * The reference's job (...MicrosoftPairHmm.cc:77-222): every read against every haplotype.  Buffers as the Java
 * class lays them out: lengths = { readCount, readLen[], hapCount, hapLen[] }; reads = per read bases | quals |
 * insGOP | delGOP | GCP; haps = bases back to back.  out[read * hapCount + hap] = log10 likelihood; escalated (may
 * be NULL) = 1 where the float result was < MIN_ACCEPTED and the pair was recomputed in double (.cc:176-203).
 * use_double: double throughout (.cc:165-171).  Returns 0, -1 bad arguments, -3 out of memory.
 */
int oracle_pairhmm(const int32_t *lengths, const uint8_t *reads, const uint8_t *haps, int use_double, double *out,
                   uint8_t *escalated)
{
    if (!lengths || !out) return -1;
    init_f();
    init_d();
    const int32_t R = lengths[0];
    if (R < 0) return -1;
    const int32_t *rl = lengths + 1;
    const int32_t H = lengths[1 + R];
    if (H < 0) return -1;
    const int32_t *hl = lengths + 2 + R;
    const uint8_t *rp = reads;
    for (int32_t r = 0; r < R; r++) {
        const int n = rl[r];
        if (n <= 0) return -1;
        const uint8_t *rs = rp, *q = rp + n, *ins = rp + 2 * (size_t)n, *del = rp + 3 * (size_t)n, *gcp = rp + 4 * (size_t)n;
        rp += 5 * (size_t)n;
        const uint8_t *hp = haps;
        for (int32_t h = 0; h < H; h++) {
            const int m = hl[h];
            if (m <= 0) return -1;
            int ok = 1, esc = 0;
            double result;
            if (!use_double) {
                const float rf = forward_f(n, rs, q, ins, del, gcp, m, hp, &ok);
                if (!ok) return -3;
                if ((double)rf < MIN_ACCEPTED) esc = 1;                                   /* .cc:182 */
                else result = log10((double)rf) - (double)log10_initial_f;                /* .cc:188 */
            }
            if (use_double || esc) {
                const double rd = forward_d(n, rs, q, ins, del, gcp, m, hp, &ok);
                if (!ok) return -3;
                result = log10(rd) - log10_initial_d;                                     /* .cc:206 */
            }
            out[(size_t)r * H + h] = result;
            if (escalated) escalated[(size_t)r * H + h] = (uint8_t)esc;
            hp += m;
        }
    }
    return 0;
}
