/* TEST HARNESS (not product, not oracle): a deterministic banded global aligner used by the iterative-refinement
 * test (BASELINE configs[4]) to re-align the reads to a new backbone between rounds.  Both
 * sides of the parity check (the CUDA path and the oracle) are fed the strings this produces, so only determinism
 * matters.  Unit-cost edit distance inside a band around the scaled diagonal; traceback prefers diagonal, then
 * deletion (target base against '-'), then insertion.
 *
 * int64 hal_align(q, n, t, m, band, qout, tout): aligns q[0..n) to t[0..m) end to end; writes the gapped strings
 * (capacity n+m+1 each) and returns their length, or -1 when the band cannot connect the corners. */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

int64_t hal_align(const char* q, int64_t n, const char* t, int64_t m, int64_t band, char* qout, char* tout) {
    const int64_t W = 2 * band + 1;
    int64_t i, k;
    if (n < 0 || m < 0 || band < 1) return -1;
    /* row i covers target columns j = c(i) - band .. c(i) + band with c(i) = i*m/n (the scaled diagonal) */
    uint8_t* tb = (uint8_t*)malloc((size_t)((n + 1) * W));
    int32_t* prev = (int32_t*)malloc(sizeof(int32_t) * (size_t)W);
    int32_t* cur = (int32_t*)malloc(sizeof(int32_t) * (size_t)W);
    const int32_t INF = 1 << 29;
    if (!tb || !prev || !cur) { free(tb); free(prev); free(cur); return -1; }
#define CEN(i_) ((n) ? ((i_) * m) / n : 0)
    for (k = 0; k < W; k++) {
        const int64_t j = CEN(0) - band + k;
        prev[k] = (j >= 0 && j <= m) ? (int32_t)j : INF;   /* leading deletions */
        tb[k] = 2;
    }
    for (i = 1; i <= n; i++) {
        const int64_t c0 = CEN(i - 1), c1 = CEN(i), sh = c1 - c0;   /* cur[k] sits at j = c1-band+k = prev index k+sh */
        for (k = 0; k < W; k++) {
            const int64_t j = c1 - band + k;
            int32_t best = INF;
            uint8_t dir = 0;
            if (j >= 0 && j <= m) {
                const int64_t kd = k + sh - 1, ku = k + sh;          /* (i-1, j-1) and (i-1, j) in prev */
                if (j >= 1 && kd >= 0 && kd < W && prev[kd] < INF) {
                    const int32_t v = prev[kd] + (q[i - 1] != t[j - 1]);
                    if (v < best) { best = v; dir = 1; }
                }
                if (j >= 1 && k >= 1 && cur[k - 1] < INF) {           /* (i, j-1): deletion */
                    const int32_t v = cur[k - 1] + 1;
                    if (v < best) { best = v; dir = 2; }
                }
                if (ku >= 0 && ku < W && prev[ku] < INF) {            /* (i-1, j): insertion */
                    const int32_t v = prev[ku] + 1;
                    if (v < best) { best = v; dir = 3; }
                }
            }
            cur[k] = best;
            tb[i * W + k] = dir;
        }
        { int32_t* x = prev; prev = cur; cur = x; }
    }
    {
        int64_t j = m, len = 0, a, b;
        i = n;
        k = j - (CEN(i) - band);
        if (k < 0 || k >= W || prev[k] >= INF) { free(tb); free(prev); free(cur); return -1; }
        while (i > 0 || j > 0) {
            const uint8_t dir = i == 0 ? 2 : tb[i * W + (j - (CEN(i) - band))];
            if (dir == 1) { qout[len] = q[i - 1]; tout[len] = t[j - 1]; i--; j--; }
            else if (dir == 2) { qout[len] = '-'; tout[len] = t[j - 1]; j--; }
            else if (dir == 3) { qout[len] = q[i - 1]; tout[len] = '-'; i--; }
            else { free(tb); free(prev); free(cur); return -1; }
            len++;
        }
        for (a = 0, b = len - 1; a < b; a++, b--) {
            char x = qout[a]; qout[a] = qout[b]; qout[b] = x;
            x = tout[a]; tout[a] = tout[b]; tout[b] = x;
        }
        qout[len] = tout[len] = 0;
        free(tb); free(prev); free(cur);
        return len;
    }
}

/* hal_lift: carries a read's alignment over from an OLD backbone to a NEW one through the alignment of the two
 * backbones (iterative refinement, BASELINE configs[4]: the consensus of a round is the next round's backbone).
 *   q, t, n        the read's gapped alignment to the old backbone, starting at old position s0 (0-based)
 *   map_new[p]     index of the new base aligned to old base p, or -1 when old base p has no partner
 *   next_new[p]    number of new bases in front of old position p (p = 0 .. old length)
 *   nb             the new backbone
 * Writes the gapped strings against the new backbone (capacity cap each), the 0-based new start and end, and
 * returns their length (-1 when cap is too small).  Deterministic and O(n): both sides of the parity check are fed
 * its output, and the bench excludes (and reports) its time. */
int64_t hal_lift(const char* q, const char* t, int64_t n, int64_t s0, const int32_t* map_new, const int32_t* next_new,
                 const char* nb, char* qo, char* to, int64_t cap, int64_t* new_start, int64_t* new_end) {
    int64_t p = s0, j = next_new[s0], len = 0, c;
    *new_start = j;
    for (c = 0; c < n; c++) {
        const char qc = q[c], tc = t[c];
        if (len + 2 >= cap) return -1;
        if (tc == '-') {
            if (qc != '-') { qo[len] = qc; to[len] = '-'; len++; }
            continue;
        }
        {
            const int32_t tgt = map_new[p];
            if (tgt >= 0) {
                while (j < tgt) { if (len + 2 >= cap) return -1; qo[len] = '-'; to[len] = nb[j]; len++; j++; }
                qo[len] = qc; to[len] = nb[j]; len++; j++;
            } else if (qc != '-') {
                qo[len] = qc; to[len] = '-'; len++;
            }
            p++;
        }
    }
    *new_end = j;
    return len;
}
