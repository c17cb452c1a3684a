/* Synthetic PacBio-like alignment generator for the parity tests and bench.py
 * (SURVEY.md section 8(d)).  Test/bench infrastructure, not part of the hot path.
 *
 * Error model = DAZZ_DB simulator's split (reference DAZZ_DB/simulator.c:56-59): of a total
 * error rate `err`, insertion .73333, deletion .20000, substitution .06667; bases uniform ACGT.
 * Alignment i of target t covers [s, s+len) with len = clamp(round(N(0.6 L, 0.2 L)), 1000, L),
 * s ~ U[0, L-len]; A = round(cov / 0.6) alignments so that mean depth = cov.  The truth
 * alignment is emitted as gapped strings in target orientation, substitutions left as mismatch
 * columns (normalizeGaps has to split them).  Everything is a pure function of
 * (seed, target index, alignment index): counter-based, so any shard of a multi-terabyte
 * config can be produced on the fly by any rank.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <stdio.h>

static inline uint64_t mix64(uint64_t z) {
    z += 0x9E3779B97F4A7C15ull;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
typedef struct { uint64_t s; } rng_t;
static inline void rng_seed(rng_t* r, uint64_t seed, uint64_t a, uint64_t b) {
    r->s = mix64(mix64(mix64(seed) ^ (a * 0xD1B54A32D192ED03ull)) ^ (b * 0x8CB92BA72F3D8DD7ull));
    if (!r->s) r->s = 1;
}
static inline uint64_t rng_next(rng_t* r) {      /* xorshift64* */
    uint64_t x = r->s;
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    r->s = x;
    return x * 0x2545F4914F6CDD1Dull;
}
static inline double rng_unif(rng_t* r) { return (double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline double rng_norm(rng_t* r) {
    double u1 = rng_unif(r), u2 = rng_unif(r);
    if (u1 < 1e-300) u1 = 1e-300;
    return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
}
static const char BASES[4] = {'A', 'C', 'G', 'T'};

typedef struct {
    uint64_t seed;
    uint32_t L;          /* backbone length */
    double cov;          /* nominal depth; A = round(cov/0.6) */
    double err;          /* total error rate, e.g. 0.15 */
} dgs_params;

int dgs_alns_per_target(const dgs_params* p) { int a = (int)floor(p->cov / 0.6 + 0.5); return a < 1 ? 1 : a; }

/* backbone of target t into out[L] */
void dgs_backbone(const dgs_params* p, uint64_t t, char* out) {
    rng_t r; uint32_t i;
    rng_seed(&r, p->seed, t, 0xFFFFFFFFull);
    for (i = 0; i < p->L; i += 32) {
        uint64_t x = rng_next(&r); uint32_t k, n = p->L - i < 32 ? p->L - i : 32;
        for (k = 0; k < n; k++) out[i + k] = BASES[(x >> (2 * k)) & 3];
    }
}

/* one alignment: writes q/t (capacity cap each), returns the number of columns (or the
 * needed capacity if cap is too small, in which case nothing useful was written). *start1 is
 * the 1-based target start (dagcon::Alignment::start convention), *tspan the target bases covered. */
static int64_t gen_aln(const dgs_params* p, uint64_t t, uint32_t a, const char* bb, char* q, char* tt,
                       int64_t cap, uint32_t* start1, uint32_t* tspan) {
    rng_t r; double e_ins = p->err * 0.73333, e_del = p->err * 0.2, e_sub = p->err * 0.06667;
    double dl; int64_t len, s, pos, n = 0;
    rng_seed(&r, p->seed, t, a);
    dl = floor(0.6 * p->L + 0.2 * p->L * rng_norm(&r) + 0.5);
    len = (int64_t)dl;
    if (len < 1000) len = 1000;
    if (len > (int64_t)p->L) len = p->L;
    s = (int64_t)(rng_unif(&r) * (double)(p->L - len + 1));
    if (s > (int64_t)p->L - len) s = (int64_t)p->L - len;
    *start1 = (uint32_t)s + 1; *tspan = (uint32_t)len;
    pos = s;
    while (pos < s + len) {
        double u = rng_unif(&r);
        char tb = bb[pos];
        if (u < e_ins) {
            if (n < cap) { q[n] = BASES[rng_next(&r) >> 62]; tt[n] = '-'; }
            n++;
        } else if (u < e_ins + e_del) {
            if (n < cap) { q[n] = '-'; tt[n] = tb; }
            n++; pos++;
        } else if (u < e_ins + e_del + e_sub) {
            char qb = BASES[rng_next(&r) >> 62];
            while (qb == tb) qb = BASES[rng_next(&r) >> 62];
            if (n < cap) { q[n] = qb; tt[n] = tb; }
            n++; pos++;
        } else {
            if (n < cap) { q[n] = tb; tt[n] = tb; }
            n++; pos++;
        }
    }
    return n;
}

/* A packed batch of targets [t0, t0+nt) in the product's dgc_batch layout (include/dagcon_b200.h):
 * per target tlen + aln_begin (CSR), per alignment start/len/q_off/t_off, one byte arena. */
typedef struct {
    uint32_t n_targets, n_alns;
    uint32_t* tgt_tlen; uint32_t* tgt_aln_begin;
    uint32_t* aln_start; uint32_t* aln_len; uint64_t* aln_q_off; uint64_t* aln_t_off;
    char* arena; uint64_t arena_bytes; uint64_t arena_cap;
    char* backbones;     /* nt * L, for tests */
} dgs_batch;

dgs_batch* dgs_batch_new(const dgs_params* p, uint64_t t0, uint32_t nt) {
    dgs_batch* b = (dgs_batch*)calloc(1, sizeof *b);
    uint32_t A = (uint32_t)dgs_alns_per_target(p), ti, a;
    int64_t cap1 = (int64_t)(p->L * 1.6) + 4096;
    char* bb;
    b->n_targets = nt; b->n_alns = nt * A;
    b->tgt_tlen = (uint32_t*)malloc(sizeof(uint32_t) * nt);
    b->tgt_aln_begin = (uint32_t*)malloc(sizeof(uint32_t) * ((size_t)nt + 1));
    b->aln_start = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)b->n_alns);
    b->aln_len = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)b->n_alns);
    b->aln_q_off = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)b->n_alns);
    b->aln_t_off = (uint64_t*)malloc(sizeof(uint64_t) * (size_t)b->n_alns);
    b->arena_cap = (uint64_t)nt * A * (uint64_t)(2 * 0.75 * p->L) + (uint64_t)(4 * cap1) + 4096;
    b->arena = (char*)malloc(b->arena_cap);
    b->backbones = (char*)malloc((size_t)nt * p->L + 1);
    for (ti = 0; ti < nt; ti++) {
        bb = b->backbones + (size_t)ti * p->L;
        dgs_backbone(p, t0 + ti, bb);
        b->tgt_tlen[ti] = p->L;
        b->tgt_aln_begin[ti] = ti * A;
        for (a = 0; a < A; a++) {
            uint32_t k = ti * A + a, st, span; int64_t n;
            if (b->arena_bytes + 2 * (uint64_t)cap1 + 64 > b->arena_cap) {
                b->arena_cap = b->arena_cap + b->arena_cap / 2 + 2 * (uint64_t)cap1 + 64;
                b->arena = (char*)realloc(b->arena, b->arena_cap);
            }
            n = gen_aln(p, t0 + ti, a, bb, b->arena + b->arena_bytes, b->arena + b->arena_bytes + cap1, cap1, &st, &span);
            if (n > cap1) { fprintf(stderr, "dgs: alignment exceeds capacity\n"); abort(); }
            /* compact t right behind q, then 16-byte align the next record */
            memmove(b->arena + b->arena_bytes + n, b->arena + b->arena_bytes + cap1, (size_t)n);
            b->aln_start[k] = st; b->aln_len[k] = (uint32_t)n;
            b->aln_q_off[k] = b->arena_bytes; b->aln_t_off[k] = b->arena_bytes + (uint64_t)n;
            b->arena_bytes += 2 * (uint64_t)n;
            b->arena_bytes = (b->arena_bytes + 15) & ~(uint64_t)15;
        }
    }
    b->tgt_aln_begin[nt] = nt * A;
    return b;
}
void dgs_batch_free(dgs_batch* b) {
    if (!b) return;
    free(b->tgt_tlen); free(b->tgt_aln_begin); free(b->aln_start); free(b->aln_len);
    free(b->aln_q_off); free(b->aln_t_off); free(b->arena); free(b->backbones); free(b);
}

static void revcomp_into(const char* s, int64_t n, char* out) {
    int64_t i;
    for (i = 0; i < n; i++) {
        char c = s[n - 1 - i];
        out[i] = c == 'A' ? 'T' : c == 'C' ? 'G' : c == 'G' ? 'C' : c == 'T' ? 'A' : c;
    }
}

/* Write targets [t0, t0+nt) as BLASR -m5 text (field layout: reference src/cpp/Alignment.cpp:44-80).
 * Odd alignments are written on the '-' strand, i.e. with both strings reverse-complemented, which
 * parseM5 undoes.  Returns bytes written or -1. */
int64_t dgs_write_m5(const dgs_params* p, uint64_t t0, uint32_t nt, const char* path) {
    FILE* f = fopen(path, "w");
    dgs_batch* b; uint32_t ti, k; int64_t total = 0; char *rq, *rt, *ms;
    size_t cap = (size_t)(p->L * 1.6) + 4096;
    if (!f) return -1;
    rq = (char*)malloc(cap); rt = (char*)malloc(cap); ms = (char*)malloc(cap);
    for (ti = 0; ti < nt; ti++) {
        b = dgs_batch_new(p, t0 + ti, 1);
        for (k = 0; k < b->n_alns; k++) {
            const char* q = b->arena + b->aln_q_off[k]; const char* t = b->arena + b->aln_t_off[k];
            int64_t n = b->aln_len[k], i, qb = 0, tb = 0;
            int minus = (k & 1);
            for (i = 0; i < n; i++) { qb += q[i] != '-'; tb += t[i] != '-'; }
            if (minus) { revcomp_into(q, n, rq); revcomp_into(t, n, rt); } else { memcpy(rq, q, (size_t)n); memcpy(rt, t, (size_t)n); }
            for (i = 0; i < n; i++) ms[i] = rq[i] == rt[i] ? '|' : '*';
            total += fprintf(f, "synthq/%llu_%u/0_%lld %lld 0 %lld + synth/%llu/0_%u %u %u %u %c -1000 0 0 0 0 254 ",
                             (unsigned long long)(t0 + ti), k, (long long)qb, (long long)qb, (long long)qb,
                             (unsigned long long)(t0 + ti), p->L, p->L, b->aln_start[k] - 1,
                             (unsigned)(b->aln_start[k] - 1 + tb), minus ? '-' : '+');
            total += (int64_t)fwrite(rq, 1, (size_t)n, f); fputc(' ', f);
            total += (int64_t)fwrite(ms, 1, (size_t)n, f); fputc(' ', f);
            total += (int64_t)fwrite(rt, 1, (size_t)n, f); fputc('\n', f);
            total += 3;
        }
        dgs_batch_free(b);
    }
    free(rq); free(rt); free(ms);
    fclose(f);
    return total;
}

/* The strings of a batch as 4-bit codes, two columns per byte (include/dagcon_b200.h: DGC_BATCH_PACKED4): what a packer
 * on the host side of the C ABI does before the upload.  out needs sum(16 + (len+1)/2 rounded up to 16) * 2 bytes.
 * Returns the packed arena's size, or -1 when a string holds a character outside "-ACGTN.". */
int64_t dgs_pack4(const uint8_t* arena, uint32_t n_alns, const uint32_t* len, const uint64_t* qoff, const uint64_t* toff,
                  uint8_t* out, uint64_t* new_qoff, uint64_t* new_toff) {
    uint8_t code[256];
    uint64_t pos = 0;
    uint32_t k;
    memset(code, 0xFF, sizeof code);
    code['-'] = 0; code['A'] = 1; code['C'] = 2; code['G'] = 3; code['T'] = 4; code['N'] = 5; code['.'] = 6;
    for (k = 0; k < n_alns; k++) {
        int s;
        for (s = 0; s < 2; s++) {
            const uint8_t* src = arena + (s ? toff[k] : qoff[k]);
            const uint32_t n = len[k];
            uint32_t c;
            uint8_t bad = 0;
            if (s) new_toff[k] = pos; else new_qoff[k] = pos;
            for (c = 0; c + 1 < n; c += 2) {
                const uint8_t a = code[src[c]], b = code[src[c + 1]];
                bad |= a | b;
                out[pos + (c >> 1)] = (uint8_t)(a | (b << 4));
            }
            if (c < n) { const uint8_t a = code[src[c]]; bad |= a; out[pos + (c >> 1)] = a; }
            if (bad & 0x80) return -1;
            pos += (((uint64_t)n + 1) / 2 + 15) & ~15ull;
        }
    }
    return (int64_t)pos;
}
