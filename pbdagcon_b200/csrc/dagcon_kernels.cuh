// dagcon_kernels.cuh -- sm_100a kernels of the DAGCon consensus hot path.
//
// Reference semantics (verdurin/pbdagcon @ 3c382f2, paths relative to the reference tree):
//   k_normalize_trim  normalizeGaps + trimAln            src/cpp/Alignment.cpp:131-242
//   k_target_plan     min-coverage filter + arena sizing src/cpp/main.cpp:66-72,118,132
//   k_build_graph     AlnGraphBoost ctor, addAln, addEdge src/cpp/AlnGraphBoost.cpp:16-127
//   k_merge_nodes     mergeNodes / mergeInNodes / mergeOutNodes / markForReaper   :129-273
//   k_best_path       bestPath + consensus(minWeight[,minLen])                    :285-459
//   k_compact         CnsResult packing for the D2H copy
//
// Data layout in HBM (per target, offsets from k_target_plan's counts):
//   Node[V]  32 B  {deg(in16|out16), base|flags|vis, in_head, out_head | weight, bbmap, score, best}
//   Edge[E]  32 B  {src, dst, count, visited, next_out, prev_out, next_in, prev_in}
//   cov[L+2] int   coverage of backbone vertices (only they are ever incremented, :76,87)
// Vertex ids are the reference's: 0 enter, 1..L backbone, L+1 exit, then insertion vertices in
// creation order.  Adjacency lists are doubly linked through the edge records so that append
// (add_edge) and order-preserving erase (clear_vertex) are O(1); head.prev is the tail.
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

#define F_NO_NORMALIZE 1u
#define F_NO_TRIM 2u
#define F_NO_MERGE 4u
#define F_SINGLE 8u

struct __align__(16) Node {
    u32 deg;  // in_deg | out_deg << 16
    u32 bfv;  // base | flags << 8 | visited-in-edge count << 16
    u32 in_head, out_head;
    int weight;
    u32 bbmap;  // _bbMap[v]; 0 for enter/exit (std::map default, AlnGraphBoost.cpp:407)
    union {
        float score;
        int cnt;
    };
    u32 best;  // BFS queue link, then (best path) the best out edge
};
struct __align__(16) Edge {
    u32 src, dst;
    int count;
    u32 visited;
    u32 next_out, prev_out, next_in, prev_in;
};
static_assert(sizeof(Node) == 32 && sizeof(Edge) == 32, "record size");

__device__ __forceinline__ u32 n_in_deg(const Node& n) { return n.deg & 0xFFFFu; }
__device__ __forceinline__ u32 n_out_deg(const Node& n) { return n.deg >> 16; }
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
    __device__ __forceinline__ void init(const u8* base, u64 o) {
        w = (const u64*)base;
        off = o;
        ci = ~0ull;
        cw = 0;
    }
    __device__ __forceinline__ u32 get(u64 r) {
        u64 a = off + r, wi = a >> 3;
        if (wi != ci) {
            cw = __ldg(w + wi);
            ci = wi;
        }
        return (u32)(cw >> ((a & 7u) * 8u)) & 0xFFu;
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
    u32 n_alns, min_len, trim, flags;
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
    u32 ai = blockIdx.x * blockDim.x + threadIdx.x;
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
    mq.init(a.arena, a.aln_q_off[ai]);
    mt.init(a.arena, a.aln_t_off[ai]);
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
    u32 n_targets, min_cov;
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
    u32 t = blockIdx.x * blockDim.x + threadIdx.x;
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
// K2: graph construction, one 256-thread block per target.  Alignments are replayed in input
// order (the adjacency-list order of every vertex is the order in which addEdge first created
// each edge, AlnGraphBoost.cpp:109-127); inside one alignment every vertex is entered and left
// at most once, so all of its columns are processed in parallel (4 per thread, 1024 per tile):
// block scans give each column its backbone position, its insertion-vertex id and its
// predecessor on the path.
// The per-backbone-position sums that every match/deletion column touches (coverage, weight,
// count of the ctor edge p->p+1, last target base) live in shared memory as 16-bit counters
// for the whole replay and are written to the vertex/edge records once at the end: without
// that every alignment re-streams the target's backbone records through DRAM (ncu, round 1:
// 38 GB of DRAM traffic per 1000 targets, 17x the arena size).
struct BuildArgs {
    u32 first;  // first target of this chunk
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    const u64* tgt_bb_off;  // or nullptr
    const u8* arena;
    const u32* tgt_status;
    const u64* node_off;
    const u64* edge_off;
    const u64* cov_off;
    const u16* norm;
    const u64* a_off;
    const u32* a_len;
    const u32* a_start;
    const u32* a_flags;
    Node* nodes;
    Edge* edges;
    int* cov;
    u32* tgt_ne;  // edges in use after the build
    u32* tgt_nv;
    u32 smem_positions;  // backbone positions (L+2) the dynamic shared memory can hold
    u16* gacc;           // global fallback for longer backbones: 4 u16 per position, at 4*cov_off
};

#define BT 256
#define CPT 4
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

__device__ __forceinline__ void list_append_out(Node* N, Edge* E, u32 u, u32 e) {
    u32 h = N[u].out_head;
    if (h == NIL) { N[u].out_head = e; E[e].prev_out = e; }
    else { u32 tail = E[h].prev_out; E[tail].next_out = e; E[e].prev_out = tail; E[h].prev_out = e; }
}
__device__ __forceinline__ void list_append_in(Node* N, Edge* E, u32 v, u32 e) {
    u32 h = N[v].in_head;
    if (h == NIL) { N[v].in_head = e; E[e].prev_in = e; }
    else { u32 tail = E[h].prev_in; E[tail].next_in = e; E[e].prev_in = tail; E[h].prev_in = e; }
}
__device__ __forceinline__ u32 find_out(const Node* N, const Edge* E, u32 u, u32 v) {
    for (u32 e = N[u].out_head; e != NIL; e = E[e].next_out)
        if (E[e].dst == v) return e;
    return NIL;
}
__device__ __forceinline__ u32 find_in(const Node* N, const Edge* E, u32 v, u32 u) {
    for (u32 e = N[v].in_head; e != NIL; e = E[e].next_in)
        if (E[e].src == u) return e;
    return NIL;
}

__global__ void __launch_bounds__(BT) k_build_graph(BuildArgs a) {
    const u32 t = a.first + blockIdx.x;
    if (a.tgt_status[t] != T_OK) return;
    const u32 tid = threadIdx.x;
    const u32 L = a.tgt_tlen[t];
    Node* N = a.nodes + a.node_off[t];
    Edge* E = a.edges + a.edge_off[t];
    int* cov = a.cov + a.cov_off[t];
    const u32 ENTER = 0, EXIT = L + 1;
    __shared__ u32 sa[BT / 32], sb[BT / 32], s_cur[TILE];
    __shared__ u32 s_carry_t, s_carry_i, s_last, s_ecur, s_ins;
    extern __shared__ __align__(16) u16 s_acc[];
    // per-position accumulators: coverage, match weight, ctor-edge count, last target base
    u16* accCov = (L + 2 <= a.smem_positions) ? s_acc : a.gacc + 4 * a.cov_off[t];
    u16* accW = accCov + (L + 2);
    u16* accC = accW + (L + 2);
    u16* accB = accC + (L + 2);

    // ctor: AlnGraphBoost(size_t) :41-62 / AlnGraphBoost(const std::string&) :16-39
    const u8* bb = a.tgt_bb_off ? a.arena + a.tgt_bb_off[t] : nullptr;
    for (u32 p = tid; p <= L + 1; p += BT) {
        Node nd;
        nd.in_head = p >= 1 ? p - 1 : NIL;
        nd.out_head = p <= L ? p : NIL;
        nd.deg = (p >= 1 ? 1u : 0u) | (p <= L ? 0x10000u : 0u);
        const bool inner = p >= 1 && p <= L;
        nd.weight = inner ? 1 : 0;
        nd.bbmap = inner ? p : 0;
        u32 b = p == 0 ? (u32)'^' : p == L + 1 ? (u32)'$' : (bb ? (u32)bb[p - 1] : (u32)'N');
        nd.bfv = b | (NF_BACKBONE << 8);
        nd.cnt = 0;
        nd.best = NIL;
        N[p] = nd;
        accCov[p] = 0; accW[p] = 0; accC[p] = 0; accB[p] = 0;
        if (p <= L) {
            Edge ed;
            ed.src = p; ed.dst = p + 1; ed.count = 0; ed.visited = 0;
            ed.next_out = NIL; ed.prev_out = p; ed.next_in = NIL; ed.prev_in = p;
            E[p] = ed;
        }
    }
    if (tid == 0) { s_ecur = L + 1; s_ins = L + 2; }
    __syncthreads();

    const u32 a0 = a.tgt_aln_begin[t], a1 = a.tgt_aln_begin[t + 1];
    for (u32 k = a0; k < a1; k++) {
        if (a.a_flags[k] & AF_SKIP) continue;
        const u32 n = a.a_len[k], start = a.a_start[k];
        const u16* cols = a.norm + a.a_off[k];
        if (tid == 0) { s_carry_t = 0; s_carry_i = 0; s_last = ENTER; }
        __syncthreads();
        for (u32 c0 = 0; c0 < n; c0 += TILE) {
            const u32 cb = c0 + tid * CPT;
            u32 q[CPT], tc[CPT], kind[CPT];  // kind: 0 none, 1 match, 2 deletion, 3 insertion
            u32 nadv = 0, nins = 0, lastp = 0;
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                kind[j] = 0; q[j] = 0; tc[j] = 0;
                if (cb + j < n) {
                    const u32 col = cols[cb + j];
                    q[j] = col & 0xFFu; tc[j] = col >> 8;
                    // column classes exactly as addAln tests them (:75, :84, :93)
                    if (q[j] == tc[j]) kind[j] = 1;
                    else if (q[j] == DASH) kind[j] = 2;
                    else if (tc[j] == DASH) kind[j] = 3;
                }
                if (kind[j] == 1 || kind[j] == 2) nadv++;
                if (kind[j] == 3) nins++;
                if (kind[j] == 1 || kind[j] == 3) lastp = tid * CPT + j + 1;
            }
            u32 ex_s, ex_m, tot_s, tot_m;
            block_scan2(nadv | (nins << 16), lastp, ex_s, ex_m, tot_s, tot_m, sa, sb);
            u32 bbpos[CPT], cur[CPT];
            {
                u32 tb = start + s_carry_t + (ex_s & 0xFFFFu), ib = s_ins + s_carry_i + (ex_s >> 16);
#pragma unroll
                for (int j = 0; j < CPT; j++) {
                    bbpos[j] = tb;
                    cur[j] = kind[j] == 1 ? tb : (kind[j] == 3 ? ib : NIL);
                    if (kind[j] == 1 || kind[j] == 2) tb++;
                    if (kind[j] == 3) ib++;
                    s_cur[tid * CPT + j] = cur[j];
                }
            }
            __syncthreads();
            u32 prev = ex_m ? s_cur[ex_m - 1] : s_last;
            const u32 ecur = s_ecur;
            // vertex side of addAln + classification of addEdge(prev, cur)
            u32 exist[CPT], pv[CPT], nnew = 0;
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                exist[j] = NIL; pv[j] = NIL;
                const u32 p = bbpos[j];
                if (kind[j] == 1 || kind[j] == 2) {
                    accCov[p]++;                 // _g[_bbMap[curr]].coverage++
                    accB[p] = (u16)tc[j];        // .base = targetBase (last alignment in order wins)
                    if (kind[j] == 1) accW[p]++;
                } else if (kind[j] == 3) {
                    Node nd;
                    nd.in_head = NIL; nd.out_head = NIL; nd.deg = 0; nd.weight = 1; nd.bbmap = p;
                    nd.bfv = q[j]; nd.cnt = 0; nd.best = NIL;
                    N[cur[j]] = nd;
                }
                if (kind[j] == 1 || kind[j] == 3) {
                    pv[j] = prev;
                    if (kind[j] == 1 && prev <= L && cur[j] == prev + 1) accC[prev]++;   // ctor edge prev -> prev+1
                    else if (kind[j] == 3 || prev >= L + 2) { exist[j] = NIL - 1; nnew++; }   // touches a fresh vertex
                    else {
                        const u32 e = prev == ENTER ? find_in(N, E, cur[j], ENTER) : find_out(N, E, prev, cur[j]);
                        if (e == NIL) { exist[j] = NIL - 1; nnew++; } else exist[j] = e;
                    }
                    prev = cur[j];
                }
            }
            u32 ex_e, d0, tot_e, d1;
            block_scan2(nnew, 0u, ex_e, d0, tot_e, d1, sa, sb);  // also orders the vertex writes
            u32 e = ecur + ex_e;
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                if (exist[j] == NIL - 1) {
                    Edge ed;
                    ed.src = pv[j]; ed.dst = cur[j]; ed.count = 1; ed.visited = 0;
                    ed.next_out = NIL; ed.prev_out = e; ed.next_in = NIL; ed.prev_in = e;
                    E[e] = ed;
                    list_append_out(N, E, pv[j], e);
                    list_append_in(N, E, cur[j], e);
                    atomicAdd(&N[pv[j]].deg, 0x10000u);
                    atomicAdd(&N[cur[j]].deg, 1u);
                    e++;
                } else if (exist[j] != NIL) {
                    E[exist[j]].count++;
                }
            }
            __syncthreads();
            if (tid == 0) {
                s_carry_t += tot_s & 0xFFFFu;
                s_carry_i += tot_s >> 16;
                if (tot_m) s_last = s_cur[tot_m - 1];
                s_ecur = ecur + tot_e;
            }
            __syncthreads();
        }
        if (tid == 0) {  // addEdge(prevVtx, _exitVtx) :106
            const u32 prev = s_last;
            u32 ex = NIL;
            if (prev == L) accC[L]++;
            else {
                if (prev < L + 2) ex = find_out(N, E, prev, EXIT);
                if (ex != NIL) E[ex].count++;
                else {
                    const u32 e = s_ecur++;
                    Edge ed;
                    ed.src = prev; ed.dst = EXIT; ed.count = 1; ed.visited = 0;
                    ed.next_out = NIL; ed.prev_out = e; ed.next_in = NIL; ed.prev_in = e;
                    E[e] = ed;
                    list_append_out(N, E, prev, e);
                    list_append_in(N, E, EXIT, e);
                    N[prev].deg += 0x10000u;
                    N[EXIT].deg += 1u;
                }
            }
            s_ins += s_carry_i;
        }
        __syncthreads();
    }
    // fold the accumulators into the records
    for (u32 p = tid; p <= L + 1; p += BT) {
        const u32 cv = accCov[p];
        cov[p] = (int)cv;
        if (cv) N[p].bfv = (N[p].bfv & ~0xFFu) | (accB[p] & 0xFFu);
        const u32 w = accW[p];
        if (w) N[p].weight += (int)w;
        if (p <= L) { const u32 cc = accC[p]; if (cc) E[p].count = (int)cc; }
    }
    if (tid == 0) { a.tgt_ne[t] = s_ecur; a.tgt_nv[t] = s_ins; }
}

// ------------------------------------------------------------------------------------------
// K3/K4: the order-sensitive part.  One scalar worker (thread) per target; `tpw` targets share
// a warp (lanes 0, 32/tpw, ...) so that small batches still spread over all SMs while large
// ones get memory-level parallelism from the divergent lanes.
struct SerialArgs {
    u32 first, count, tpw;
    u32 flags, min_weight, min_len;
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    u32* tgt_status;
    const u64* node_off;
    const u64* edge_off;
    const u64* cov_off;
    const u32* tgt_E;  // edge capacity
    u32* tgt_ne;
    const u32* tgt_nv;
    Node* nodes;
    Edge* edges;
    int* cov;
    u32* gscratch;          // overflow of the per-worker group snapshot stack
    const u64* gscratch_off;
    const u32* gscratch_cap;
    // best path outputs
    u8* cons;
    const u64* cons_off;
    int* frags;             // (range0, range1) pairs
    const u64* frag_off;
    const u32* frag_cap;
    u32* tgt_cons_len;
    u32* tgt_nfrag;
};

struct TCtx {
    Node* N;
    Edge* E;
    u32 e_bump, e_cap, free_head;
    u32* S;
    u32 scap;
    u32* gS;
    u32 gscap;
    u32 status;
};
__device__ __forceinline__ u32 S_get(const TCtx& c, u32 i) { return i < c.scap ? c.S[i] : c.gS[i - c.scap]; }
__device__ __forceinline__ bool S_put(TCtx& c, u32 i, u32 v) {
    if (i < c.scap) { c.S[i] = v; return true; }
    if (i - c.scap < c.gscap) { c.gS[i - c.scap] = v; return true; }
    c.status = T_INTERNAL;
    return false;
}

// boost::add_edge + property copy (AlnGraphBoost.cpp:198-203, :252-257)
__device__ __forceinline__ void g_add_edge(TCtx& c, u32 u, u32 v, int count, u32 visited) {
    u32 e;
    if (c.free_head != NIL) { e = c.free_head; c.free_head = c.E[e].next_out; }
    else if (c.e_bump < c.e_cap) e = c.e_bump++;
    else { c.status = T_INTERNAL; return; }
    Node* N = c.N;
    Edge* E = c.E;
    Edge ed;
    ed.src = u; ed.dst = v; ed.count = count; ed.visited = visited;
    ed.next_out = NIL; ed.next_in = NIL; ed.prev_out = e; ed.prev_in = e;
    u32 h = N[u].out_head;
    if (h == NIL) N[u].out_head = e;
    else { u32 tail = E[h].prev_out; E[tail].next_out = e; ed.prev_out = tail; E[h].prev_out = e; }
    h = N[v].in_head;
    if (h == NIL) N[v].in_head = e;
    else { u32 tail = E[h].prev_in; E[tail].next_in = e; ed.prev_in = tail; E[h].prev_in = e; }
    E[e] = ed;
    N[u].deg += 0x10000u;
    N[v].deg += 1u;
    if (visited) N[v].bfv += 0x10000u;
}
__device__ __forceinline__ void g_unlink_in(Node* N, Edge* E, u32 v, u32 e) {
    const u32 p = E[e].prev_in, nx = E[e].next_in, h = N[v].in_head;
    if (e == h) { N[v].in_head = nx; if (nx != NIL) E[nx].prev_in = p; }
    else { E[p].next_in = nx; if (nx != NIL) E[nx].prev_in = p; else E[h].prev_in = p; }
}
__device__ __forceinline__ void g_unlink_out(Node* N, Edge* E, u32 u, u32 e) {
    const u32 p = E[e].prev_out, nx = E[e].next_out, h = N[u].out_head;
    if (e == h) { N[u].out_head = nx; if (nx != NIL) E[nx].prev_out = p; }
    else { E[p].next_out = nx; if (nx != NIL) E[nx].prev_out = p; else E[h].prev_out = p; }
}
// markForReaper: deleted = true; clear_vertex (AlnGraphBoost.cpp:269-273)
__device__ __forceinline__ void g_reap(TCtx& c, u32 ni) {
    Node* N = c.N;
    Edge* E = c.E;
    for (u32 e = N[ni].out_head; e != NIL;) {
        const u32 nx = E[e].next_out, v = E[e].dst;
        if (v != ni) {
            g_unlink_in(N, E, v, e);
            N[v].deg -= 1u;
            if (E[e].visited) N[v].bfv -= 0x10000u;
        }
        E[e].next_out = c.free_head; c.free_head = e;
        e = nx;
    }
    for (u32 e = N[ni].in_head; e != NIL;) {
        const u32 nx = E[e].next_in, s = E[e].src;
        if (s != ni) {
            g_unlink_out(N, E, s, e);
            N[s].deg -= 0x10000u;
            E[e].next_out = c.free_head; c.free_head = e;
        }
        e = nx;
    }
    N[ni].in_head = NIL; N[ni].out_head = NIL; N[ni].deg = 0;
    N[ni].bfv = (N[ni].bfv & 0xFFFFu) | (NF_DELETED << 8);
}

// Snapshot of the merge candidates of n, in list order (the std::map<char, std::vector<VtxDesc>>
// built at :167-181 / :222-236).  Frame: [n, cnt, last key + 129, parent frame, members...].
__device__ __forceinline__ u32 push_in_frame(TCtx& c, u32 n, u32& top, u32 parent) {
    Node* N = c.N;
    Edge* E = c.E;
    if (n_in_deg(N[n]) < 2) return NIL;
    u32 cnt = 0;
    for (u32 e = N[n].in_head; e != NIL; e = E[e].next_in) {
        const u32 s = E[e].src;
        if (n_out_deg(N[s]) == 1) { if (!S_put(c, top + 4 + cnt, s)) return NIL; cnt++; }
    }
    if (cnt < 2) return NIL;
    const u32 fb = top;
    S_put(c, fb, n); S_put(c, fb + 1, cnt); S_put(c, fb + 2, 0); S_put(c, fb + 3, parent);
    top += 4 + cnt;
    return fb;
}

// next key in ascending (signed) char order; returns false when the map is exhausted
__device__ __forceinline__ bool next_group(TCtx& c, u32 fb, u32& an, u32& members) {
    const u32 cnt = S_get(c, fb + 1);
    const int last = (int)S_get(c, fb + 2) - 129;
    int key = 1000;
    for (u32 i = 0; i < cnt; i++) {
        int b = (int)(signed char)n_base(c.N[S_get(c, fb + 4 + i)].bfv);
        if (b > last && b < key) key = b;
    }
    if (key == 1000) return false;
    S_put(c, fb + 2, (u32)(key + 129));
    an = NIL; members = 0;
    for (u32 i = 0; i < cnt; i++) {
        const u32 ni = S_get(c, fb + 4 + i);
        if ((int)(signed char)n_base(c.N[ni].bfv) == key) { if (an == NIL) an = ni; members++; }
    }
    return true;
}

// mergeInNodes :162-215 (recursion unrolled onto the snapshot stack)
__device__ void g_merge_in(TCtx& c, u32 n0) {
    Node* N = c.N;
    Edge* E = c.E;
    u32 top = 0;
    u32 fb = push_in_frame(c, n0, top, NIL);
    while (fb != NIL && c.status == T_OK) {
        u32 an, members;
        if (!next_group(c, fb, an, members)) { top = fb; fb = S_get(c, fb + 3); continue; }
        if (members < 2) continue;
        const u32 cnt = S_get(c, fb + 1);
        const int key = (int)S_get(c, fb + 2) - 129;
        const u32 an_out = N[an].out_head;  // *anoi, taken once (:186-188)
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S_get(c, fb + 4 + i);
            if (ni == an || (int)(signed char)n_base(N[ni].bfv) != key) continue;
            E[an_out].count += E[N[ni].out_head].count;
            N[an].weight += N[ni].weight;
        }
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S_get(c, fb + 4 + i);
            if (ni == an || (int)(signed char)n_base(N[ni].bfv) != key) continue;
            for (u32 ie = N[ni].in_head; ie != NIL; ie = E[ie].next_in) {
                const u32 n1 = E[ie].src;
                const u32 e = find_out(N, E, n1, an);
                if (e != NIL) E[e].count += E[ie].count;
                else g_add_edge(c, n1, an, E[ie].count, E[ie].visited);
            }
            g_reap(c, ni);
        }
        const u32 nfb = push_in_frame(c, an, top, fb);  // mergeInNodes(an) :212
        if (nfb != NIL) fb = nfb;
    }
}

// mergeOutNodes :217-267
__device__ void g_merge_out(TCtx& c, u32 n) {
    Node* N = c.N;
    Edge* E = c.E;
    if (n_out_deg(N[n]) < 2) return;
    u32 cnt = 0;
    for (u32 e = N[n].out_head; e != NIL; e = E[e].next_out) {
        const u32 d = E[e].dst;
        if (n_in_deg(N[d]) == 1) { if (!S_put(c, 4 + cnt, d)) return; cnt++; }
    }
    if (cnt < 2) return;
    S_put(c, 0, n); S_put(c, 1, cnt); S_put(c, 2, 0); S_put(c, 3, NIL);
    u32 an, members;
    while (c.status == T_OK && next_group(c, 0, an, members)) {
        if (members < 2) continue;
        const int key = (int)S_get(c, 2) - 129;
        const u32 an_in = N[an].in_head;
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S_get(c, 4 + i);
            if (ni == an || (int)(signed char)n_base(N[ni].bfv) != key) continue;
            E[an_in].count += E[N[ni].in_head].count;
            N[an].weight += N[ni].weight;
        }
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S_get(c, 4 + i);
            if (ni == an || (int)(signed char)n_base(N[ni].bfv) != key) continue;
            for (u32 oe = N[ni].out_head; oe != NIL; oe = E[oe].next_out) {
                const u32 n2 = E[oe].dst;
                const u32 e = find_out(N, E, an, n2);
                if (e != NIL) E[e].count += E[oe].count;
                else g_add_edge(c, an, n2, E[oe].count, E[oe].visited);
            }
            g_reap(c, ni);
        }
    }
}

// ---- lean fast paths.  The serial kernels are instruction-issue bound (ncu, round 1: 385
// instructions per popped vertex, 1 active lane), so the common case -- a vertex whose
// neighbourhood needs no merge -- is handled with 128-bit loads of the record halves and a
// neighbour snapshot in shared memory that the visited-marking loop reuses.
__device__ __forceinline__ uint4 ld4(const void* p) { return *reinterpret_cast<const uint4*>(p); }
#define NODE_A(N, v) ld4(&(N)[v])                           // deg, bfv, in_head, out_head
#define NODE_B(N, v) ld4(((const uint4*)&(N)[v]) + 1)       // weight, bbmap, score|cnt, best
#define EDGE_A(E, e) ld4(&(E)[e])                           // src, dst, count, visited
#define EDGE_B(E, e) ld4(((const uint4*)&(E)[e]) + 1)       // next_out, prev_out, next_in, prev_in

#define MAXK 62  // neighbours snapshotted on the fast path (4 arrays + an 8-word bitmap = 256 words)
#define TOO_BIG 0xFFFFFFFFu

// One pass over the out-list of u: se = edge id | visited << 31, sd = target, sx = target bfv,
// sy = target deg.  While at it, find the bases that occur more than once among the merge
// candidates (targets with in_degree 1): dup_lo/dup_hi are bitmaps over base 0..63 / 64..127.
// Bases >= 128 (negative as `char`, sorted first by std::map<char,...>) set `weird` and send the
// vertex down the generic path.
struct OutScan { u32 cnt; u64 dup_lo, dup_hi; bool weird; };
__device__ __forceinline__ OutScan scan_out(const Node* N, const Edge* E, u32 head, u32* se, u32* sd, u32* sx, u32* sy) {
    OutScan r;
    r.cnt = 0; r.dup_lo = 0; r.dup_hi = 0; r.weird = false;
    u64 seen_lo = 0, seen_hi = 0;
    for (u32 e = head; e != NIL;) {
        const uint4 ea = EDGE_A(E, e);
        const u32 nx = E[e].next_out;
        const uint4 da = NODE_A(N, ea.y);
        if (r.cnt >= MAXK) { r.cnt = TOO_BIG; return r; }
        se[r.cnt] = e | (ea.w ? 0x80000000u : 0u); sd[r.cnt] = ea.y; sx[r.cnt] = da.y; sy[r.cnt] = da.x;
        r.cnt++;
        if ((da.x & 0xFFFFu) == 1) {
            const u32 b = da.y & 0xFFu;
            const u64 m = 1ull << (b & 63);
            if (b < 64) { r.dup_lo |= seen_lo & m; seen_lo |= m; }
            else if (b < 128) { r.dup_hi |= seen_hi & m; seen_hi |= m; }
            else r.weird = true;
        }
        e = nx;
    }
    return r;
}
__device__ __forceinline__ bool bitmap_test_set(u32* bm, u32 b) {
    const u32 w = b >> 5, m = 1u << (b & 31);
    const u32 old = bm[w];
    bm[w] = old | m;
    return old & m;
}
// does any std::map group of mergeInNodes(n) hold more than one vertex (predecessors with
// out_degree 1 sharing a base)?
__device__ __forceinline__ bool in_group_exists(const Node* N, const Edge* E, u32 in_head) {
    u64 seen_lo = 0, seen_hi = 0;
    bool found = false;
    for (u32 e = in_head; e != NIL;) {
        const u32 s = E[e].src, nx = E[e].next_in;
        const uint4 sa = NODE_A(N, s);
        if ((sa.x >> 16) == 1) {
            const u32 b = sa.y & 0xFFu;
            const u64 m = 1ull << (b & 63);
            if (b < 64) { found |= (seen_lo & m) != 0; seen_lo |= m; }
            else if (b < 128) { found |= (seen_hi & m) != 0; seen_hi |= m; }
            else found = true;  // let the generic code sort it out
        }
        e = nx;
    }
    return found;
}

// mergeOutNodes :217-267 for one std::map key, driven by the shared-memory snapshot.  The two
// loops of the reference (sum counts/weights, then move edges and reap) touch disjoint state
// per merged vertex, so they are fused.
__device__ __forceinline__ void merge_out_group(TCtx& c, u32 key, u32 cnt, const u32* sd, const u32* sx, const u32* sy) {
    Node* N = c.N;
    Edge* E = c.E;
    u32 an = NIL, an_in = NIL;
    for (u32 i = 0; i < cnt; i++) {
        if ((sy[i] & 0xFFFFu) != 1 || (sx[i] & 0xFFu) != key) continue;
        const u32 ni = sd[i];
        if (an == NIL) { an = ni; an_in = N[an].in_head; continue; }
        const uint4 nia = NODE_A(N, ni);
        E[an_in].count += E[nia.z].count;
        N[an].weight += N[ni].weight;
        for (u32 oe = nia.w; oe != NIL;) {
            const uint4 oa = EDGE_A(E, oe);
            const u32 nx = E[oe].next_out;
            const u32 e = find_out(N, E, an, oa.y);
            if (e != NIL) E[e].count += (int)oa.z;
            else g_add_edge(c, an, oa.y, (int)oa.z, oa.w);
            oe = nx;
        }
        g_reap(c, ni);
    }
}

// mergeNodes :129-160
__device__ void g_merge_nodes(TCtx& c, u32 ENTER) {
    Node* N = c.N;
    Edge* E = c.E;
    u32* se = c.S;
    u32* sd = se + MAXK;
    u32* sx = sd + MAXK;
    u32* sy = sx + MAXK;
    u32 head = ENTER, tail = ENTER;
    N[ENTER].best = NIL;
    N[ENTER].bfv |= NF_QUEUED << 8;
    while (head != NIL && c.status == T_OK) {
        const u32 u = head;
        uint4 a = NODE_A(N, u);
        if ((a.x & 0xFFFFu) >= 2 && in_group_exists(N, E, a.z)) {
            g_merge_in(c, u);
            a = NODE_A(N, u);
        }
        OutScan sc;
        sc.cnt = 0; sc.dup_lo = 0; sc.dup_hi = 0; sc.weird = false;
        if (a.w != NIL) sc = scan_out(N, E, a.w, se, sd, sx, sy);
        if (sc.cnt == TOO_BIG || sc.weird) {  // very wide vertex (enter) or exotic bases: generic path
            g_merge_out(c, u);
            for (u32 e = N[u].out_head; e != NIL; e = E[e].next_out) {
                const u32 v = E[e].dst;
                if (!E[e].visited) { E[e].visited = 1; N[v].bfv += 0x10000u; }
                const u32 bfv = N[v].bfv;
                if (n_vis(bfv) == n_in_deg(N[v])) {  // notVisited == 0
                    if (n_flags(bfv) & NF_QUEUED) { c.status = T_INTERNAL; break; }
                    N[v].bfv = bfv | (NF_QUEUED << 8);
                    N[tail].best = v;
                    tail = v;
                }
            }
            head = N[u].best;
            continue;
        }
        if (sc.dup_lo | sc.dup_hi) {
            // groups in ascending key order, as std::map<char, ...> iterates them
            for (u64 m = sc.dup_lo; m && c.status == T_OK; m &= m - 1) merge_out_group(c, (u32)__ffsll((long long)m) - 1, sc.cnt, sd, sx, sy);
            for (u64 m = sc.dup_hi; m && c.status == T_OK; m &= m - 1) merge_out_group(c, 63u + (u32)__ffsll((long long)m), sc.cnt, sd, sx, sy);
            sc = scan_out(N, E, N[u].out_head, se, sd, sx, sy);  // merging only removed out-edges of u
        }
        for (u32 i = 0; i < sc.cnt; i++) {
            const u32 ev = se[i], v = sd[i];
            u32 bfv = sx[i];
            if (!(ev >> 31)) { E[ev].visited = 1; bfv += 0x10000u; }
            if ((bfv >> 16) == (sy[i] & 0xFFFFu)) {  // notVisited == 0
                if (bfv & (NF_QUEUED << 8)) { c.status = T_INTERNAL; break; }
                bfv |= NF_QUEUED << 8;
                N[tail].best = v;   // N[v].best is still NIL from the build
                tail = v;
            }
            if (bfv != sx[i]) N[v].bfv = bfv;
        }
        head = N[u].best;
    }
}

__device__ __forceinline__ bool serial_setup(const SerialArgs& a, u32& t, TCtx& c, u32* smem, u32 scap_total) {
    const u32 lane = threadIdx.x & 31, stride = 32 / a.tpw;
    if (lane % stride) return false;
    const u32 warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const u32 idx = warp * a.tpw + lane / stride;
    if (idx >= a.count) return false;
    t = a.first + idx;
    if (a.tgt_status[t] != T_OK) return false;
    const u32 workers = (blockDim.x >> 5) * a.tpw;
    const u32 w = (threadIdx.x >> 5) * a.tpw + lane / stride;
    c.scap = scap_total / workers;
    c.S = smem + w * c.scap;
    c.N = a.nodes + a.node_off[t];
    c.E = a.edges + a.edge_off[t];
    c.e_bump = a.tgt_ne[t];
    c.e_cap = a.tgt_E[t];
    c.free_head = NIL;
    c.gS = a.gscratch + a.gscratch_off[t];
    c.gscap = a.gscratch_cap[t];
    c.status = T_OK;
    return true;
}

#define SERIAL_THREADS 128
#define SERIAL_SMEM_WORDS 4096  // 16 KB per block: 1024 words per worker at tpw=1, 256 at tpw=4

__global__ void __launch_bounds__(SERIAL_THREADS) k_merge_nodes(SerialArgs a) {
    __shared__ u32 smem[SERIAL_SMEM_WORDS];
    u32 t;
    TCtx c;
    if (!serial_setup(a, t, c, smem, SERIAL_SMEM_WORDS)) return;
    g_merge_nodes(c, 0);
    a.tgt_ne[t] = c.e_bump;
    if (c.status != T_OK) a.tgt_status[t] = c.status;
}

// bestPath :375-459 and the two consensus() forms :285-373
__global__ void __launch_bounds__(SERIAL_THREADS) k_best_path(SerialArgs a) {
    u32 t;
    TCtx c;
    if (!serial_setup(a, t, c, nullptr, SERIAL_SMEM_WORDS)) return;
    Node* N = c.N;
    Edge* E = c.E;
    const int* cov = a.cov + a.cov_off[t];
    const u32 L = a.tgt_tlen[t], ENTER = 0, EXIT = L + 1, nV = a.tgt_nv[t];
    // reverse BFS from exit; `cnt` counts visited out-edges until the vertex is scored
    u32 head = EXIT, tail = EXIT;
    N[EXIT].best = NIL;
    while (head != NIL) {
        const u32 n = head;
        const uint4 na = NODE_A(N, n);
        float best_score = -FLT_MAX;
        u32 best_e = NIL;
        for (u32 e = na.w; e != NIL;) {
            const uint4 ea = EDGE_A(E, e);
            const u32 nx = E[e].next_out;
            const uint4 oa = NODE_A(N, ea.y);
            const uint4 ob = NODE_B(N, ea.y);
            const float sc = __uint_as_float(ob.z);
            float ns;
            if ((oa.y & (NF_BACKBONE << 8)) && (int)ob.x == 1) ns = __fsub_rn(sc, 10.0f);
            else ns = __fadd_rn(__fsub_rn((float)(int)ea.z, __fmul_rn((float)cov[ob.y], 0.5f)), sc);
            if (ns > best_score) { best_score = ns; best_e = e; }
            e = nx;
        }
        for (u32 e = na.z; e != NIL;) {
            const u32 s = E[e].src, nx = E[e].next_in;
            const u32 k = (u32)(++N[s].cnt);
            if (k == (N[s].deg >> 16)) {  // all out-edges visited
                N[s].best = NIL;
                N[tail].best = s;
                tail = s;
            }
            e = nx;
        }
        const u32 nx = N[n].best;
        N[n].score = best_e != NIL ? best_score : 0.0f;  // absent map key reads as 0.0f (:401)
        N[n].best = best_e;
        N[n].bfv = na.y | (NF_SCORED << 8);
        head = nx;
    }
    // walk the best path and cut it into min-weight runs
    u8* cons = a.cons + a.cons_off[t];
    int* fr = a.frags + 2 * a.frag_off[t];
    const u32 fcap = a.frag_cap[t];
    const u32 eb = n_base(N[ENTER].bfv), xb = n_base(N[EXIT].bfv);
    const int minw = (int)a.min_weight;
    const bool single = a.flags & F_SINGLE;
    int offs = 0, idx = 0, best_offs = 0, length = 0;
    bool met = false;
    u32 nfrag = 0, steps = 0, prev = ENTER, status = T_OK;
    for (;;) {
        const uint4 pa = NODE_A(N, prev);
        const uint4 pb = NODE_B(N, prev);
        const u32 b = pa.y & 0xFFu;
        const int w = (int)pb.x;
        if (b != eb && b != xb) {
            cons[idx] = (u8)b;
            if (!met && w >= minw) { offs = idx; met = true; }
            else if (met && w < minw) {
                met = false;
                if (single) { if (idx - offs > length) { best_offs = offs; length = idx - offs; } }
                else if ((u32)(idx - offs) >= a.min_len) {
                    if (nfrag < fcap) { fr[2 * nfrag] = offs; fr[2 * nfrag + 1] = idx; nfrag++; } else status = T_INTERNAL;
                }
            }
            idx++;
        }
        if (!(pa.y & (NF_SCORED << 8)) || pb.w == NIL) break;
        prev = E[pb.w].dst;
        if (++steps > nV) { status = T_INTERNAL; break; }
    }
    if (single) {
        if (met && idx - offs > length) { best_offs = offs; length = idx - offs; }
        fr[0] = best_offs; fr[1] = best_offs + length; nfrag = 1;
    } else if (met && (u32)(idx - offs) >= a.min_len) {
        if (nfrag < fcap) { fr[2 * nfrag] = offs; fr[2 * nfrag + 1] = idx; nfrag++; } else status = T_INTERNAL;
    }
    a.tgt_cons_len[t] = (u32)idx;
    a.tgt_nfrag[t] = nfrag;
    if (status != T_OK) { a.tgt_status[t] = status; a.tgt_nfrag[t] = 0; }
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
