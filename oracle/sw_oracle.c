/*
 * oracle/sw_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of microsoft/mgl's Smith-Waterman affine-gap aligner, written
 * from the recurrence as explicit full matrices rather
 * than the reference's rolling arrays.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this file's
 * library; nothing under mgl_b200/ links or calls it.
 *
 * Parity status: PINNED.  The reference ships no golden vectors for this path
 * (SURVEY.md section 0, facts 1-2), so the restatement is pinned by executing the
 * reference's own C++ (oracle/_ref, built by oracle/Makefile from the sources
 * in place under /root/reference) -- see tests/test_oracle.py and the
 * committed outputs in tests/golden/.
 *
 * Reference lines restated (relative to src/main/native/mgl_sw/):
 *   boundary + seeds        sw.cpp:10-40,47-49
 *   cell recurrence         sw.cpp:54-93
 *   last-column argmax      sw.cpp:100-104
 *   last-row scan           sw.cpp:116-127
 *   traceback start cell    sw.cpp:155-170
 *   traceback walk          sw.cpp:173-214
 *   overhang post-process   sw.cpp:224-248
 *   text formatting         sw.cpp:251-252
 *   score sign normalising  com_microsoft_mgl_smithwaterman_MicrosoftSmithWaterman.cpp:52-55
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <stdio.h>

#define OS_SOFTCLIP 1   /* sw_common.h:22 */
#define OS_INDEL    2   /* sw_common.h:23 */
#define OS_LEAD_ID  4   /* sw_common.h:24 */
#define OS_IGNORE   8   /* sw_common.h:25 */
#define NEG_INF (-0x40000000) /* sw_common.h:33 */

/* CIGAR op codes used by the packed form: (len << 4) | op, BAM numbering. */
enum { OP_M = 0, OP_I = 1, OP_D = 2, OP_S = 4 };
static const char OP_CHR[] = "MIDNS";

typedef struct {
    int32_t offset;      /* alignment offset (return value of alignNative)     */
    int32_t n_ops;       /* number of CIGAR elements with length > 0            */
    int32_t score;       /* score of the traceback start cell (diagnostic)      */
    int32_t start_i;     /* traceback start row    (diagnostic)                 */
    int32_t start_j;     /* traceback start column (diagnostic)                 */
} oracle_result;

/*
 * Align one pair.  Scores are taken as the JNI passes them and normalised the
 * way the reference does.  ops_out receives up to ops_cap packed elements in
 * CIGAR (left to right) order.  Returns 0, or -1 on bad arguments, -2 when
 * ops_cap is too small (n_ops still reports the needed count), -3 on OOM.
 */
int oracle_sw_align(const uint8_t *ref, int tl, const uint8_t *alt, int ql,
                    int match, int mismatch, int gap_open, int gap_extend,
                    int strategy, uint32_t *ops_out, int ops_cap,
                    oracle_result *res)
{
    if (tl <= 0 || ql <= 0 || !ref || !alt || !res) return -1;
    if (strategy != OS_SOFTCLIP && strategy != OS_INDEL &&
        strategy != OS_LEAD_ID && strategy != OS_IGNORE) return -1;

    /* ...SmithWaterman.cpp:52-55 */
    const int w_match = match > 0 ? match : -match;
    const int w_mis   = mismatch < 0 ? mismatch : -mismatch;
    const int go      = gap_open > 0 ? gap_open : -gap_open;
    const int ge      = gap_extend > 0 ? gap_extend : -gap_extend;
    const int indel   = (strategy & (OS_INDEL | OS_LEAD_ID)) != 0;

    const size_t W = (size_t)ql + 2;          /* row stride, columns 0..ql+1 */
    const size_t cells = ((size_t)tl + 2) * W;
    int32_t *H  = (int32_t *)calloc(cells, sizeof(int32_t));
    int32_t *E  = (int32_t *)calloc(cells, sizeof(int32_t));  /* E[i][j]: vertical gap entering (i,j)   */
    int32_t *F  = (int32_t *)calloc(cells, sizeof(int32_t));  /* F[i][j]: horizontal gap entering (i,j) */
    int32_t *LE = (int32_t *)calloc(cells, sizeof(int32_t));  /* run length of E[i][j] */
    int32_t *LF = (int32_t *)calloc(cells, sizeof(int32_t));  /* run length of F[i][j] */
    int32_t *B  = (int32_t *)calloc(cells, sizeof(int32_t));  /* 0 diag, -k insertion run, +k deletion run */
    if (!H || !E || !F || !LE || !LF || !B) {
        free(H); free(E); free(F); free(LE); free(LF); free(B);
        return -3;
    }
#define AT(M, i, j) (M)[(size_t)(i) * W + (size_t)(j)]

    /* boundary: sw.cpp:10-40 (row 0) and :47-49 (column 0); H[0][0] stays 0 */
    for (int j = 1; j <= ql; j++) {
        AT(H, 0, j) = indel ? -go - (j - 1) * ge : 0;
        AT(E, 1, j) = AT(H, 0, j) - go;  AT(LE, 1, j) = 1;
    }
    for (int i = 1; i <= tl; i++) {
        AT(H, i, 0) = indel ? -go - (i - 1) * ge : 0;
        AT(F, i, 1) = AT(H, i, 0) - go;  AT(LF, i, 1) = 1;
    }

    int mqe = NEG_INF, mqe_t = -1;
    for (int i = 1; i <= tl; i++) {
        for (int j = 1; j <= ql; j++) {
            const int d = AT(H, i - 1, j - 1) + (ref[i - 1] == alt[j - 1] ? w_match : w_mis);
            const int e = AT(E, i, j), f = AT(F, i, j);
            int h;
            /* priority diag, then horizontal ("right"), then vertical: sw.cpp:59-71 */
            if (d >= e && d >= f)      { h = d; AT(B, i, j) = 0; }
            else if (f >= e)           { h = f; AT(B, i, j) = -AT(LF, i, j); }
            else                       { h = e; AT(B, i, j) =  AT(LE, i, j); }
            AT(H, i, j) = h;
            /* open only when strictly better than extending: sw.cpp:73-93 */
            if (h - go > e - ge) { AT(E, i + 1, j) = h - go; AT(LE, i + 1, j) = 1; }
            else                 { AT(E, i + 1, j) = e - ge; AT(LE, i + 1, j) = AT(LE, i, j) + 1; }
            if (h - go > f - ge) { AT(F, i, j + 1) = h - go; AT(LF, i, j + 1) = 1; }
            else                 { AT(F, i, j + 1) = f - ge; AT(LF, i, j + 1) = AT(LF, i, j) + 1; }
        }
        /* last column, ">=" with ascending i: sw.cpp:100-104 */
        if (AT(H, i, ql) >= mqe) { mqe = AT(H, i, ql); mqe_t = i; }
    }

    /* last row: sw.cpp:116-127 */
    int best = mqe, best_t = mqe_t, best_q = ql, seg = 0;
    for (int j = 1; j <= ql; j++) {
        const int s = AT(H, tl, j);
        if (s > best || (s == best && abs(tl - j) < abs(best_t - best_q))) {
            best = s; best_t = tl; best_q = j; seg = ql - j;
        }
    }

    /* traceback start: sw.cpp:155-170 */
    int I, J, run = 0;
    if (strategy == OS_INDEL)        { I = tl;     J = ql; }
    else if (strategy != OS_LEAD_ID) { I = best_t; J = best_q; run = seg; }
    else                             { I = mqe_t;  J = ql; }
    res->score = AT(H, I, J); res->start_i = I; res->start_j = J;

    /* elements are produced back to front; collect reversed, flip at the end */
    size_t cap = (size_t)tl + (size_t)ql + 8, n = 0;
    int32_t *el_len = (int32_t *)malloc(cap * sizeof(int32_t));
    uint8_t *el_op  = (uint8_t *)malloc(cap);
    if (!el_len || !el_op) {
        free(el_len); free(el_op);
        free(H); free(E); free(F); free(LE); free(LF); free(B);
        return -3;
    }
#define EMIT(op, len) do { el_op[n] = (uint8_t)(op); el_len[n] = (len); n++; } while (0)

    if (run > 0 && strategy == OS_SOFTCLIP) { EMIT(OP_S, run); run = 0; }   /* sw.cpp:173-176 */

    int state = OP_M;
    do {                                                                     /* sw.cpp:182-214 */
        const int b = AT(B, I, J);
        int next, step = 1;
        if (b > 0)      { next = OP_D; step = b;  I -= step; }
        else if (b < 0) { next = OP_I; step = -b; J -= step; }
        else            { next = OP_M; I--; J--; }
        if (next == state) run += step;
        else { EMIT(state, run); run = step; state = next; }
    } while (I > 0 && J > 0);

    int offset;
    if (strategy == OS_SOFTCLIP) {                                           /* sw.cpp:225-229 */
        EMIT(state, run);
        if (J > 0) EMIT(OP_S, J);
        offset = I;
    } else if (strategy == OS_IGNORE) {                                      /* sw.cpp:230-233 */
        EMIT(state, run + J);
        offset = I - J;
    } else {                                                                 /* sw.cpp:234-248 */
        EMIT(state, run);
        if (I > 0)      EMIT(OP_D, I);
        else if (J > 0) EMIT(OP_I, J);
        offset = 0;
    }

    /* keep only positive-length elements, in left-to-right order: sw.cpp:251-252 */
    int n_ops = 0, rc = 0;
    for (size_t k = n; k-- > 0;) {
        if (el_len[k] <= 0) continue;
        if (n_ops < ops_cap && ops_out) ops_out[n_ops] = ((uint32_t)el_len[k] << 4) | el_op[k];
        else rc = -2;
        n_ops++;
    }
    res->offset = offset;
    res->n_ops = n_ops;

    free(el_len); free(el_op);
    free(H); free(E); free(F); free(LE); free(LF); free(B);
    return rc;
#undef EMIT
#undef AT
}

/* Packed elements -> reference text form ("3M2D5M"); returns characters written (no NUL counted). */
int oracle_sw_format(const uint32_t *ops, int n_ops, char *out, int out_cap)
{
    int pos = 0;
    for (int k = 0; k < n_ops; k++) {
        char tmp[16];
        int len = snprintf(tmp, sizeof tmp, "%u%c", ops[k] >> 4, OP_CHR[ops[k] & 15]);
        if (pos + len >= out_cap) return -1;
        memcpy(out + pos, tmp, (size_t)len);
        pos += len;
    }
    if (pos < out_cap) out[pos] = 0;
    return pos;
}

/* Convenience: align and format in one call (what a caller of alignNative sees). */
int oracle_sw_align_text(const uint8_t *ref, int tl, const uint8_t *alt, int ql,
                         int match, int mismatch, int gap_open, int gap_extend,
                         int strategy, char *cigar, int cigar_cap, int32_t *offset)
{
    int cap = tl + ql + 8;
    uint32_t *ops = (uint32_t *)malloc((size_t)cap * sizeof(uint32_t));
    if (!ops) return -3;
    oracle_result r;
    int rc = oracle_sw_align(ref, tl, alt, ql, match, mismatch, gap_open, gap_extend,
                             strategy, ops, cap, &r);
    if (rc == 0) {
        *offset = r.offset;
        if (oracle_sw_format(ops, r.n_ops, cigar, cigar_cap) < 0) rc = -2;
    }
    free(ops);
    return rc;
}
