// dagcon_kernels.cuh -- sm_100a kernels of the DAGCon consensus hot path.
//
// Reference semantics (verdurin/pbdagcon @ 3c382f2, paths relative to the reference tree):
//   k_normalize_trim  normalizeGaps + trimAln            src/cpp/Alignment.cpp:131-242
//   k_target_plan     min-coverage filter + arena sizing src/cpp/main.cpp:66-72,118,132
//   k_compact         CnsResult packing for the D2H copy
// The warp-per-alignment normalizer is in normalize_warp.cuh; the alignment graph (k_build_graph2, k_merge_nodes2,
// k_best_path2: AlnGraphBoost.cpp:16-459) and its record layout are in graph_rec.cuh.
#pragma once
#include <stdint.h>
#include <float.h>
#include <cuda_runtime.h>

typedef uint8_t u8;
typedef uint16_t u16;
typedef uint32_t u32;
typedef unsigned long long u64;

#define NIL 0xFFFFFFFFu
#define NF_BACKBONE 1u
#define NF_DELETED 2u
#define NF_QUEUED 4u
#define NF_SCORED 8u

#define AF_SKIP 1u  // raw qstr.length() < minLen (main.cpp:132)

#define T_OK 0u
#define T_SKIPPED 1u
#define T_BAD_ALIGNMENT 2u
#define T_TOO_MANY 3u
#define T_INTERNAL 4u
#define T_POOL 5u  // internal: a list pool ran dry; the dispatcher rebuilds the target with a worst-case pool

#define F_NO_NORMALIZE 1u
#define F_NO_TRIM 2u
#define F_NO_MERGE 4u
#define F_SINGLE 8u
#define F_NO_PREFETCH 256u
#define F_PACKED4 0x10000u  // internal (NormArgs.flags): the batch's strings are 4-bit codes (DGC_BATCH_PACKED4)

// 4-bit wire codes -> ASCII: "-ACGTN." (code 7 is never sent; it reads as 'N')
#define PACK4_LUT_LO 0x4743412Du   // codes 0..3: '-' 'A' 'C' 'G' (little-endian bytes)
#define PACK4_LUT_HI 0x4E2E4E54u   // codes 4..7: 'T' 'N' '.' 'N'
__device__ __forceinline__ u32 pack4_ascii(u32 code) {
    return ((code & 4u ? PACK4_LUT_HI : PACK4_LUT_LO) >> (8u * (code & 3u))) & 0xFFu;
}

__device__ __forceinline__ u32 n_base(u32 bfv) { return bfv & 0xFFu; }
__device__ __forceinline__ u32 n_flags(u32 bfv) { return (bfv >> 8) & 0xFFu; }
__device__ __forceinline__ u32 n_vis(u32 bfv) { return bfv >> 16; }

// ------------------------------------------------------------------------------------------
// K1 (serial form): normalizeGaps + trimAln, one thread per alignment.  Runs the alignments the warp kernel
// (normalize_warp.cuh) hands back (AF_RETRY) and everything when normalisation is switched off.
// normalizeGaps is a left-to-right transducer with look-ahead: column i may pull the next
// target (query) base leftwards into its gap (Alignment.cpp:167-197).  Because a pull always
// takes the *next* non-gap character, the mutated state is fully described by two cursors
// JT/JQ = "next target/query base not yet consumed"; everything between the main cursor and
// JT (JQ) is a gap in the current state.  That makes the pass O(len) with three forward
// cursors over the raw strings and no scratch besides the output.
struct ByteStream {
    const u64* w;
    u64 off, ci, cw;
    u32 packed;   // the string is 4-bit codes, two columns per byte (DGC_BATCH_PACKED4)
    __device__ __forceinline__ void init(const u8* base, u64 o, u32 packed4 = 0) {
        w = (const u64*)base;
        off = o;
        ci = ~0ull;
        cw = 0;
        packed = packed4;
    }
    __device__ __forceinline__ u32 get(u64 r) {
        u64 a = off + (packed ? r >> 1 : r), wi = a >> 3;
        if (wi != ci) {
            cw = __ldg(w + wi);
            ci = wi;
        }
        const u32 b = (u32)(cw >> ((a & 7u) * 8u)) & 0xFFu;
        return packed ? pack4_ascii((b >> (4u * (u32)(r & 1u))) & 7u) : b;
    }
};

struct Look {
    ByteStream q, t;
    u64 r, e;  // raw column, expanded index of its first column
    u64 epos;  // expanded index of the pending base, ~0 = none left
    u32 ch, nc;
};
#define DASH 45u
#define DOT 46u
__device__ __forceinline__ void look_find(Look& c, u64 n, bool want_t) {
    while (c.r < n) {
        u32 qb = c.q.get(c.r), tb = c.t.get(c.r);
        if (qb == DOT) qb = DASH;
        if (tb == DOT) tb = DASH;
        u32 m = (qb != tb && qb != DASH && tb != DASH) ? 1u : 0u;
        if (want_t) {
            if (tb != DASH) { c.ch = tb; c.nc = 1 + m; c.epos = c.e; return; }
        } else {
            if (qb != DASH) { c.ch = qb; c.nc = 1 + m; c.epos = c.e + m; return; }
        }
        c.e += 1 + m;
        c.r++;
    }
    c.epos = ~0ull;
}
__device__ __forceinline__ void look_adv(Look& c, u64 n, bool want_t) {
    c.e += c.nc;
    c.r++;
    look_find(c, n, want_t);
}

struct NormArgs {
    const u8* arena;
    const u32* aln_start;
    const u32* aln_len;
    const u64* aln_q_off;
    const u64* aln_t_off;
    const u64* aln_norm_off;  // column offset of this alignment's slot in `norm` (capacity 2*len)
    u32 n_alns, min_len, trim, flags;  // n_alns: end of the alignment range
    u32 aln_first;                     // start of the alignment range (staged uploads run one piece at a time)
    u32 retry_only;  // serial kernel: handle only the alignments the warp kernel flagged AF_RETRY
    u16* norm;  // q | t << 8 per column
    u64* o_off;
    u32* o_len;
    u32* o_start;
    u32* o_flags;
    u32* o_nI;     // insertion columns kept (new vertices)
    u32* o_nE;     // upper bound on new edges (addEdge calls that can miss)
    u32* o_tkept;  // backbone positions consumed
};

__global__ void __launch_bounds__(128) k_normalize_trim(NormArgs a) {
    u32 ai = a.aln_first + blockIdx.x * blockDim.x + threadIdx.x;
    if (ai >= a.n_alns) return;
    if (a.retry_only && a.o_flags[ai] != 2u /* AF_RETRY */) return;
    const u64 n = a.aln_len[ai];
    const u64 base = a.aln_norm_off[ai];
    u16* out = a.norm + base;
    u64 m = 0;
    if (n < a.min_len) {
        a.o_off[ai] = base; a.o_len[ai] = 0; a.o_start[ai] = a.aln_start[ai]; a.o_flags[ai] = AF_SKIP;
        a.o_nI[ai] = 0; a.o_nE[ai] = 0; a.o_tkept[ai] = 0;
        return;
    }
    ByteStream mq, mt;
    mq.init(a.arena, a.aln_q_off[ai], a.flags & F_PACKED4);
    mt.init(a.arena, a.aln_t_off[ai], a.flags & F_PACKED4);
    if (a.flags & F_NO_NORMALIZE) {
        for (u64 r = 0; r < n; r++) out[m++] = (u16)(mq.get(r) | (mt.get(r) << 8));
    } else {
        Look jt, jq;
        jt.q = mq; jt.t = mt; jt.r = 0; jt.e = 0;
        jq.q = mq; jq.t = mt; jq.r = 0; jq.e = 0;
        look_find(jt, n, true);
        look_find(jq, n, false);
        u64 e = 0;
        for (u64 r = 0; r < n; r++) {
            u32 qb = mq.get(r), tb = mt.get(r);
            if (qb == DOT) qb = DASH;
            if (tb == DOT) tb = DASH;
            const bool mism = (qb != tb && qb != DASH && tb != DASH);
            const u32 ncols = mism ? 2u : 1u;
            for (u32 h = 0; h < ncols; h++) {
                const u64 i = e + h;
                // mismatch -> deletion column then insertion column (Alignment.cpp:147-159)
                const u32 qo = mism ? (h == 0 ? DASH : qb) : qb;
                const u32 to = mism ? (h == 0 ? tb : DASH) : tb;
                u32 tc = (to != DASH && jt.epos == i) ? to : DASH;
                u32 qc = (qo != DASH && jq.epos == i) ? qo : DASH;
                const bool last = (r == n - 1) && (h == ncols - 1);
                if (!last) {
                    if (tc != DASH) look_adv(jt, n, true);
                    else if (jt.epos != ~0ull && jt.ch == qc) { tc = jt.ch; look_adv(jt, n, true); }
                    if (qc != DASH) look_adv(jq, n, false);
                    else if (jq.epos != ~0ull && jq.ch == tc) { qc = jq.ch; look_adv(jq, n, false); }
                }
                if (qc != DASH || tc != DASH) out[m++] = (u16)(qc | (tc << 8));
            }
            e += ncols;
        }
    }
    // trimAln (Alignment.cpp:219-242)
    u64 lo = 0, hi = m;
    u32 lb = 0, rb = 0;
    if (!(a.flags & F_NO_TRIM)) {
        while (lb < a.trim && lo < m) { if ((out[lo++] >> 8) != DASH) lb++; }
        while (rb < a.trim && hi > lo) { if ((out[--hi] >> 8) != DASH) rb++; }
    }
    // sizes of what addAln will create (same column classes as AlnGraphBoost.cpp:75-104)
    u32 nI = 0, nE = 1, tk = 0;
    bool prev_ins = false, jumped = true;  // enter -> first vertex counts as a possible new edge
    for (u64 c = lo; c < hi; c++) {
        u32 col = out[c], q = col & 0xFFu, t = col >> 8;
        if (q == t) { if (prev_ins || jumped) nE++; prev_ins = false; jumped = false; tk++; }
        else if (q == DASH) { jumped = true; tk++; }
        else if (t == DASH) { nE++; nI++; prev_ins = true; jumped = false; }
    }
    a.o_off[ai] = base + lo;
    a.o_len[ai] = (u32)(hi - lo);
    a.o_start[ai] = a.aln_start[ai] + lb;
    a.o_flags[ai] = 0;
    a.o_nI[ai] = nI;
    a.o_nE[ai] = nE;
    a.o_tkept[ai] = tk;
}

#include "normalize_warp.cuh"

// ------------------------------------------------------------------------------------------
// K1b: per target, the min-coverage filter (main.cpp:66-72,118), alignment validation, and the
// exact vertex count / edge bound that size the arena.
struct PlanArgs {
    u32 n_targets, min_cov;  // n_targets: end of the target range
    u32 t_first;             // start of the target range
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    const u32* a_start;
    const u32* a_flags;
    const u32* a_nI;
    const u32* a_nE;
    const u32* a_tkept;
    const u32* a_len;
    u32* tgt_status;
    u32* tgt_V;
    u32* tgt_E;
    u64* tgt_cols;
};
__global__ void k_target_plan(PlanArgs a) {
    u32 t = a.t_first + blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= a.n_targets) return;
    u32 a0 = a.tgt_aln_begin[t], a1 = a.tgt_aln_begin[t + 1], A = a1 - a0, L = a.tgt_tlen[t];
    u32 st = T_OK;
    u64 sI = 0, sE = 0, cols = 0;
    if (A == 0 || A < a.min_cov) st = T_SKIPPED;
    else if (A > 65000u) st = T_TOO_MANY;
    else {
        for (u32 k = a0; k < a1; k++) {
            if (a.a_flags[k] & AF_SKIP) continue;
            u64 s = a.a_start[k];
            if (s < 1 || s - 1 + a.a_tkept[k] > L) st = T_BAD_ALIGNMENT;
            sI += a.a_nI[k]; sE += a.a_nE[k]; cols += a.a_len[k];
        }
        if ((u64)L + 2 + sI > 0x7FFFFFF0ull || (u64)L + 1 + sE + A + 16 > 0x7FFFFFF0ull) st = T_TOO_MANY;
    }
    a.tgt_status[t] = st;
    a.tgt_V[t] = st == T_OK ? (u32)(L + 2 + sI) : 0;
    a.tgt_E[t] = st == T_OK ? (u32)(L + 1 + sE + A + 16) : 0;
    a.tgt_cols[t] = st == T_OK ? cols : 0;
}

// ------------------------------------------------------------------------------------------
// Block-wide tile scan shared by the graph build (graph_rec.cuh): 256 threads x 4 columns.
// 512 threads x 2 columns: measured on B200 (5000 targets of 8 kb x 30x) 256x4 44.9 ms, 384x2 40.4, 512x2 39.1, 512x4 39.8,
// 768x2 47.2, 1024x2 43.6, 256x8 65.6, 128x4 57.0
#ifndef BT
#define BT 512
#endif
#ifndef CPT
#define CPT 2
#endif
#define TILE (BT * CPT)
__device__ __forceinline__ void block_scan2(u32 vs, u32 vm, u32& ex_s, u32& ex_m, u32& tot_s, u32& tot_m,
                                            u32* sa, u32* sb) {
    const u32 lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    u32 s = vs, m = vm;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        u32 ts = __shfl_up_sync(0xFFFFFFFFu, s, d), tm = __shfl_up_sync(0xFFFFFFFFu, m, d);
        if (lane >= (u32)d) { s += ts; m = max(m, tm); }
    }
    u32 pm = __shfl_up_sync(0xFFFFFFFFu, m, 1);
    if (lane == 0) pm = 0;
    if (lane == 31) { sa[wid] = s; sb[wid] = m; }
    __syncthreads();
    u32 ps = 0, pmx = 0, ts = 0, tm = 0;
#pragma unroll
    for (u32 w = 0; w < BT / 32; w++) {
        u32 x = sa[w], y = sb[w];
        if (w < wid) { ps += x; pmx = max(pmx, y); }
        ts += x; tm = max(tm, y);
    }
    ex_s = ps + s - vs;
    ex_m = max(pmx, pm);
    tot_s = ts;
    tot_m = tm;
    __syncthreads();
}

// Staged uploads: the planner's per-target offsets live in pinned (device-visible) host memory.  They are pulled by
// a kernel instead of cudaMemcpyAsync because an H2D memcpy would queue on the copy engine BEHIND the gigabytes of
// alignment strings still travelling on the copy stream, and the build of the first piece would wait for all of them.
struct PullArgs {
    u32* dst[8];
    const u32* src[8];
    u32 words[8];
};
__global__ void __launch_bounds__(256) k_pull_words(PullArgs a) {
#pragma unroll 1
    for (int j = 0; j < 8; j++)
        for (u32 i = blockIdx.x * blockDim.x + threadIdx.x; i < a.words[j]; i += gridDim.x * blockDim.x) a.dst[j][i] = a.src[j][i];
}

#include "graph_rec.cuh"

// ------------------------------------------------------------------------------------------
// Output compaction: exclusive scans over targets (one block), then a gather.
__global__ void __launch_bounds__(1024) k_scan_outputs(u32 n, const u32* status, const u32* cons_len, const u32* nfrag,
                                                       u64* cons_coff, u32* frag_coff, u64* totals) {
    // totals[0..1] = consensus bytes / fragments emitted by earlier chunks; updated on exit
    __shared__ u64 s_c[1024];
    __shared__ u32 s_f[1024];
    __shared__ u64 s_tot[2];
    const u64 base_c = totals[0], base_f = totals[1];
    const u32 tid = threadIdx.x, per = (n + 1023) / 1024;
    const u32 b = min(n, tid * per), e = min(n, b + per);
    u64 sc = 0;
    u32 sf = 0;
    for (u32 i = b; i < e; i++) if (status[i] == T_OK) { sc += cons_len[i]; sf += nfrag[i]; }
    s_c[tid] = sc; s_f[tid] = sf;
    __syncthreads();
    if (tid == 0) {
        u64 ac = 0; u32 af = 0;
        for (u32 i = 0; i < 1024; i++) { u64 x = s_c[i]; u32 y = s_f[i]; s_c[i] = ac; s_f[i] = af; ac += x; af += y; }
        s_tot[0] = ac; s_tot[1] = af;
    }
    __syncthreads();
    sc = base_c + s_c[tid]; sf = (u32)base_f + s_f[tid];
    for (u32 i = b; i < e; i++) {
        cons_coff[i] = sc; frag_coff[i] = sf;
        if (status[i] == T_OK) { sc += cons_len[i]; sf += nfrag[i]; }
    }
    __syncthreads();
    if (tid == 0) { totals[0] = base_c + s_tot[0]; totals[1] = base_f + s_tot[1]; }
}

__global__ void __launch_bounds__(128) k_compact(u32 n, const u32* status, const u32* cons_len, const u32* nfrag,
                                                 const u8* cons, const u64* cons_off, const int* frags,
                                                 const u64* frag_off, const u64* cons_coff, const u32* frag_coff,
                                                 u8* out_cons, int* out_frags) {
    const u32 t = blockIdx.x;
    if (t >= n || status[t] != T_OK) return;
    const u8* src = cons + cons_off[t];
    u8* dst = out_cons + cons_coff[t];
    const u32 len = cons_len[t];
    for (u32 i = threadIdx.x; i < len; i += blockDim.x) dst[i] = src[i];
    const int* fs = frags + 2 * frag_off[t];
    int* fd = out_frags + 2 * (u64)frag_coff[t];
    for (u32 i = threadIdx.x; i < 2 * nfrag[t]; i += blockDim.x) fd[i] = fs[i];
}
