// graph_v2.cuh -- the alignment graph as fat vertex records (K2 build, K3 mergeNodes, K4 bestPath/consensus).
//
// Reference semantics (verdurin/pbdagcon @ 3c382f2): src/cpp/AlnGraphBoost.cpp
//   ctor :16-62, addAln :64-107, addEdge :109-127, mergeNodes :129-160, mergeInNodes :162-215,
//   mergeOutNodes :217-267, markForReaper :269-273, consensus :285-373, bestPath :375-459.
//
// Layout.  A vertex IS its record; an edge lives in its source's out-slot (dst, count, visited) and is mirrored
// by a 4-byte in-slot (src) at its target -- there is no edge array.  (u,v) identifies an edge because neither
// addEdge nor the merges ever create a parallel edge (:114-121, :194-203, :248-257).
//   backbone vertex p (0 = enter, 1..L, L+1 = exit)   256 B: Hdr | 16 out-slots | 24 in-slots      at bb  + 256 p
//   insertion vertex v >= L+2 (creation order)        128 B: Hdr |  6 out-slots | 12 in-slots      at ins + 128 (v-L-2)
// Slots keep the adjacency-list ORDER the reference's tie-breaks depend on: append = write at [deg], order
// preserving erase = shift down.  A list that outgrows its inline slots moves as a whole into a block of the
// target's pool (Hdr.ext_*; block word 0 = capacity).  One 128-byte line therefore holds everything mergeNodes
// and bestPath need to know about a vertex, and the header of a successor -- fetched when its predecessor is
// processed -- brings that successor's whole record on chip before the sweep reaches it.
#pragma once

#define BB_REC 256u
#define INS_REC 128u
#define BB_OUT_CAP 16u
#define BB_IN_CAP 24u
#define INS_OUT_CAP 6u
#define INS_IN_CAP 12u
#define BB_IN_OFF (32u + 8u * BB_OUT_CAP)    // 160
#define INS_IN_OFF (32u + 8u * INS_OUT_CAP)  // 80
#define VISBIT 0x80000000u

struct __align__(16) Hdr {
    u32 bfv;      // base | flags << 8 | visited-in-edge count << 16
    u32 deg;      // in_deg | out_deg << 16
    u32 ext_out;  // 0: inline out-slots; else pool index (8-byte units) of the list's block
    u32 ext_in;
    int weight;
    int cov;      // coverage of _bbMap[v] (bestPath, :407); an insertion vertex holds its bbmap here until the build folds
    union {
        float score;
        int cnt;
    };
    u32 best;     // BFS queue link, then (bestPath) index of the best out-slot
};
struct OSlot {
    u32 dst;
    u32 cv;  // count | visited << 31
};
static_assert(sizeof(Hdr) == 32 && sizeof(OSlot) == 8, "record layout");

struct GV {           // one target's graph
    u8* bb;           // (L+2) backbone records
    u8* ins;          // insertion records
    u64* pool;        // list blocks
    u32 nbb;          // L + 2
    u32 pool_top, pool_cap;
    u32 status;
};

__device__ __forceinline__ u8* gv_rec(const GV& g, u32 h) {
    return h < g.nbb ? g.bb + (size_t)BB_REC * h : g.ins + (size_t)INS_REC * (h - g.nbb);
}
__device__ __forceinline__ Hdr* gv_hdr(const GV& g, u32 h) { return (Hdr*)gv_rec(g, h); }
__device__ __forceinline__ OSlot* gv_out(const GV& g, u32 h, u32 ext_out) {
    return ext_out ? (OSlot*)(g.pool + ext_out + 1) : (OSlot*)(gv_rec(g, h) + 32);
}
__device__ __forceinline__ u32* gv_in(const GV& g, u32 h, u32 ext_in) {
    return ext_in ? (u32*)(g.pool + ext_in + 1) : (u32*)(gv_rec(g, h) + (h < g.nbb ? BB_IN_OFF : INS_IN_OFF));
}
__device__ __forceinline__ u32 gv_out_cap(const GV& g, u32 h, u32 ext_out) {
    return ext_out ? (u32)g.pool[ext_out] : (h < g.nbb ? BB_OUT_CAP : INS_OUT_CAP);
}
__device__ __forceinline__ u32 gv_in_cap(const GV& g, u32 h, u32 ext_in) {
    return ext_in ? (u32)g.pool[ext_in] : (h < g.nbb ? BB_IN_CAP : INS_IN_CAP);
}
// a block of `units` 8-byte words behind one capacity word; 0 when the pool is exhausted
__device__ __forceinline__ u32 gv_alloc(GV& g, u32 units, u32 cap) {
    if (g.pool_top + units + 1 > g.pool_cap) { g.status = T_INTERNAL; return 0; }
    const u32 off = g.pool_top;
    g.pool_top += units + 1;
    g.pool[off] = cap;
    return off;
}

// boost::edge(u, v, g): first match in out(u); returns the slot index or NIL
__device__ __forceinline__ u32 gv_find_out(const OSlot* o, u32 n, u32 v) {
    for (u32 i = 0; i < n; i++)
        if (o[i].dst == v) return i;
    return NIL;
}

__device__ __forceinline__ OSlot* gv_edge_slot(const GV& g, u32 u, u32 v) {
    const Hdr* hu = gv_hdr(g, u);
    OSlot* o = gv_out(g, u, hu->ext_out);
    const u32 i = gv_find_out(o, hu->deg >> 16, v);
    return i == NIL ? nullptr : o + i;
}

// boost::add_edge(u, v) with the AlnEdge properties copied (:198-203, :252-257); u's and v's headers are updated
__device__ __forceinline__ void gv_add_edge(GV& g, u32 u, u32 v, u32 count, u32 visited) {
    Hdr* hu = gv_hdr(g, u);
    {
        const u32 n = hu->deg >> 16;
        u32 ext = hu->ext_out;
        OSlot* o = gv_out(g, u, ext);
        if (n == gv_out_cap(g, u, ext)) {
            const u32 ncap = max(8u, 2u * n);
            const u32 blk = gv_alloc(g, ncap, ncap);
            if (!blk) return;
            OSlot* no = (OSlot*)(g.pool + blk + 1);
            for (u32 i = 0; i < n; i++) no[i] = o[i];
            hu->ext_out = blk;
            o = no;
        }
        OSlot s;
        s.dst = v; s.cv = count | (visited ? VISBIT : 0u);
        o[n] = s;
        hu->deg += 0x10000u;
    }
    Hdr* hv = gv_hdr(g, v);
    {
        const u32 n = hv->deg & 0xFFFFu;
        u32 ext = hv->ext_in;
        u32* in = gv_in(g, v, ext);
        if (n == gv_in_cap(g, v, ext)) {
            const u32 ncap = max(16u, 2u * n);
            const u32 blk = gv_alloc(g, ncap / 2, ncap);
            if (!blk) return;
            u32* ni = (u32*)(g.pool + blk + 1);
            for (u32 i = 0; i < n; i++) ni[i] = in[i];
            hv->ext_in = blk;
            in = ni;
        }
        in[n] = u;
        hv->deg += 1u;
        if (visited) hv->bfv += 0x10000u;
    }
}

// order-preserving erase of the edge (u -> v) from v's in-list / u's out-list (half of boost::clear_vertex)
__device__ __forceinline__ void gv_erase_in(GV& g, u32 v, u32 u, bool was_visited) {
    Hdr* hv = gv_hdr(g, v);
    const u32 n = hv->deg & 0xFFFFu;
    u32* in = gv_in(g, v, hv->ext_in);
    u32 i = 0;
    while (i < n && in[i] != u) i++;
    if (i == n) { g.status = T_INTERNAL; return; }
    for (; i + 1 < n; i++) in[i] = in[i + 1];
    hv->deg -= 1u;
    if (was_visited) hv->bfv -= 0x10000u;
}
__device__ __forceinline__ void gv_erase_out(GV& g, u32 u, u32 v) {
    Hdr* hu = gv_hdr(g, u);
    const u32 n = hu->deg >> 16;
    OSlot* o = gv_out(g, u, hu->ext_out);
    u32 i = gv_find_out(o, n, v);
    if (i == NIL) { g.status = T_INTERNAL; return; }
    for (; i + 1 < n; i++) o[i] = o[i + 1];
    hu->deg -= 0x10000u;
}

// markForReaper (:269-273): deleted = true; clear_vertex(n)
__device__ __forceinline__ void gv_reap(GV& g, u32 ni) {
    Hdr* h = gv_hdr(g, ni);
    const u32 deg = h->deg;
    const OSlot* o = gv_out(g, ni, h->ext_out);
    for (u32 i = 0; i < (deg >> 16); i++) {
        const OSlot s = o[i];
        if (s.dst != ni) gv_erase_in(g, s.dst, ni, (s.cv & VISBIT) != 0);
    }
    const u32* in = gv_in(g, ni, h->ext_in);
    for (u32 i = 0; i < (deg & 0xFFFFu); i++) {
        const u32 s = in[i];
        if (s != ni) gv_erase_out(g, s, ni);
    }
    h->deg = 0;
    h->bfv = (h->bfv & 0xFFFFu) | (NF_DELETED << 8);
}

// ------------------------------------------------------------------------------------------
// K2: graph construction (ctor :16-62, addAln :64-107, addEdge :109-127), one 256-thread block per target.
// Alignments are replayed in input order (every adjacency list is ordered by the alignment that first created
// each edge); inside one alignment every vertex is entered and left at most once, so its columns are processed in
// parallel, 1024 per tile: block scans give each column its backbone position, its insertion-vertex id and its
// predecessor on the path.  Per backbone position the running sums (coverage, match weight, count of the ctor
// edge p->p+1, last target base, list lengths) stay in shared memory for the whole replay and are folded into
// the records once at the end; insertion records are written exactly once.
struct Build2Args {
    u32 first;
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    const u64* tgt_bb_off;  // or nullptr
    const u8* arena;
    const u32* tgt_status;
    const u64* bb_off;    // record index of the target's first backbone record
    const u64* ins_off;   // record index of its first insertion record
    const u64* pool_off;  // 8-byte units
    const u32* pool_cap;
    const u64* acc_off;   // positions before this target (global accumulator fallback)
    const u16* norm;
    const u64* a_off;
    const u32* a_len;
    const u32* a_start;
    const u32* a_flags;
    u8* bb;
    u8* ins;
    u64* pool;
    u32* tgt_status_out;
    u32* tgt_nv;
    u32* tgt_pool_top;
    u32 smem_positions;
    u8* gacc;  // 9 bytes per position
};

struct BuildAcc {  // per-position running state of the replay
    u16* cov;
    u16* w;
    u16* c;
    u8* b;
    u8* dout;
    u8* din;
};

// where the out-list of backbone vertex p lives when it holds n entries (n <= BB_OUT_CAP: still inline)
__device__ __forceinline__ OSlot* b2_out_list(u8* bbrec, u64* pool, u32 p, u32 n, u32 enter_ext) {
    if (p == 0) return (OSlot*)(pool + enter_ext + 1);
    if (n <= BB_OUT_CAP) return (OSlot*)(bbrec + (size_t)BB_REC * p + 32);
    return (OSlot*)(pool + ((const Hdr*)(bbrec + (size_t)BB_REC * p))->ext_out + 1);
}
__device__ __forceinline__ u32* b2_in_list(u8* bbrec, u64* pool, u32 p, u32 n, u32 exit_v, u32 exit_ext) {
    if (p == exit_v) return (u32*)(pool + exit_ext + 1);
    if (n <= BB_IN_CAP) return (u32*)(bbrec + (size_t)BB_REC * p + BB_IN_OFF);
    return (u32*)(pool + ((const Hdr*)(bbrec + (size_t)BB_REC * p))->ext_in + 1);
}

struct B2Ctx {
    u8* bbrec;
    u8* insrec;
    u64* pool;
    u32 L, A, enter_ext, exit_ext, pool_cap;
    u32* s_pool_top;
    u32* s_status;
    u32* s_enter_out;
    u32* s_exit_in;
    BuildAcc acc;
};

// append (p -> v, count 1) to backbone vertex p's out-list
__device__ __forceinline__ void b2_append_out(const B2Ctx& c, u32 p, u32 v) {
    u32 n;
    OSlot* o;
    if (p == 0) {
        n = (*c.s_enter_out)++;
        o = (OSlot*)(c.pool + c.enter_ext + 1);
    } else {
        n = c.acc.dout[p];
        if (n >= 254u) { *c.s_status = T_TOO_MANY; return; }
        c.acc.dout[p] = (u8)(n + 1);
        o = b2_out_list(c.bbrec, c.pool, p, n, c.enter_ext);
        if (n == BB_OUT_CAP) {  // the inline slots are full: the list moves into a pool block that can hold any degree
            const u32 cap = c.A + 2, blk = atomicAdd(c.s_pool_top, cap + 1);
            if (blk + cap + 1 > c.pool_cap) { *c.s_status = T_INTERNAL; return; }
            c.pool[blk] = cap;
            OSlot* no = (OSlot*)(c.pool + blk + 1);
            for (u32 i = 0; i < n; i++) no[i] = o[i];
            ((Hdr*)(c.bbrec + (size_t)BB_REC * p))->ext_out = blk;
            o = no;
        }
    }
    OSlot s;
    s.dst = v; s.cv = 1;
    o[n] = s;
}
__device__ __forceinline__ void b2_append_in(const B2Ctx& c, u32 p, u32 u) {
    u32 n;
    u32* in;
    if (p == c.L + 1) {
        n = (*c.s_exit_in)++;
        in = (u32*)(c.pool + c.exit_ext + 1);
    } else {
        n = c.acc.din[p];
        if (n >= 254u) { *c.s_status = T_TOO_MANY; return; }
        c.acc.din[p] = (u8)(n + 1);
        in = b2_in_list(c.bbrec, c.pool, p, n, c.L + 1, c.exit_ext);
        if (n == BB_IN_CAP) {
            const u32 cap = c.A + 2, units = (cap + 1) / 2, blk = atomicAdd(c.s_pool_top, units + 1);
            if (blk + units + 1 > c.pool_cap) { *c.s_status = T_INTERNAL; return; }
            c.pool[blk] = cap;
            u32* ni = (u32*)(c.pool + blk + 1);
            for (u32 i = 0; i < n; i++) ni[i] = in[i];
            ((Hdr*)(c.bbrec + (size_t)BB_REC * p))->ext_in = blk;
            in = ni;
        }
    }
    in[n] = u;
}

// addEdge(prev, cur) (:109-127) for one path step; prev/cur are vertex ids, `fresh` marks insertion vertices
// created by the alignment being replayed (their single in/out slot is written directly)
__device__ __forceinline__ void b2_edge(const B2Ctx& c, u32 prev, u32 cur) {
    const u32 nbb = c.L + 2;
    const bool prev_bb = prev < nbb, cur_bb = cur < nbb;
    if (prev_bb && cur_bb) {
        if (cur == prev + 1) { c.acc.c[prev]++; return; }  // the ctor edge p -> p+1
        // a backbone->backbone skip (deletion run, alignment start or end): it may exist already
        const u32 n = prev == 0 ? *c.s_enter_out : c.acc.dout[prev];
        OSlot* o = b2_out_list(c.bbrec, c.pool, prev, n, c.enter_ext);
        for (u32 i = 0; i < n; i++)
            if (o[i].dst == cur) { o[i].cv++; return; }
        b2_append_out(c, prev, cur);
        b2_append_in(c, cur, prev);
        return;
    }
    if (prev_bb) b2_append_out(c, prev, cur);
    else { OSlot s; s.dst = cur; s.cv = 1; *(OSlot*)(c.insrec + (size_t)INS_REC * (prev - nbb) + 32) = s; }
    if (cur_bb) b2_append_in(c, cur, prev);
    else *(u32*)(c.insrec + (size_t)INS_REC * (cur - nbb) + INS_IN_OFF) = prev;
}

__global__ void __launch_bounds__(BT) k_build_graph2(Build2Args a) {
    const u32 t = a.first + blockIdx.x;
    if (a.tgt_status[t] != T_OK) return;
    const u32 tid = threadIdx.x;
    const u32 L = a.tgt_tlen[t];
    const u32 ENTER = 0, EXIT = L + 1, nbb = L + 2;
    const u32 a0 = a.tgt_aln_begin[t], a1 = a.tgt_aln_begin[t + 1];
    __shared__ u32 sa[BT / 32], sb[BT / 32], s_cur[TILE];
    __shared__ u32 s_carry_t, s_carry_i, s_last, s_ins, s_pool_top, s_status, s_enter_out, s_exit_in;
    extern __shared__ __align__(16) u8 s_dyn[];
    u8* accbase = (nbb <= a.smem_positions) ? s_dyn : a.gacc + 9ull * a.acc_off[t];
    B2Ctx c;
    c.acc.cov = (u16*)accbase;
    c.acc.w = c.acc.cov + nbb;
    c.acc.c = c.acc.w + nbb;
    c.acc.b = (u8*)(c.acc.c + nbb);
    c.acc.dout = c.acc.b + nbb;
    c.acc.din = c.acc.dout + nbb;
    c.bbrec = a.bb + (size_t)BB_REC * a.bb_off[t];
    c.insrec = a.ins + (size_t)INS_REC * a.ins_off[t];
    c.pool = a.pool + a.pool_off[t];
    c.pool_cap = a.pool_cap[t];
    c.L = L; c.A = a1 - a0;
    c.enter_ext = 1;                       // block 0 of the pool is never handed out: ext == 0 means "inline"
    c.exit_ext = 1 + (c.A + 2) + 1;
    c.s_pool_top = &s_pool_top; c.s_status = &s_status; c.s_enter_out = &s_enter_out; c.s_exit_in = &s_exit_in;

    // ctor: AlnGraphBoost(size_t) :41-62 / AlnGraphBoost(const std::string&) :16-39
    for (u32 p = tid; p < nbb; p += BT) {
        c.acc.cov[p] = 0; c.acc.w[p] = 0; c.acc.c[p] = 0; c.acc.b[p] = 0;
        c.acc.dout[p] = p <= L ? 1 : 0;
        c.acc.din[p] = p >= 1 ? 1 : 0;
        u8* r = c.bbrec + (size_t)BB_REC * p;
        Hdr h;
        h.bfv = 0; h.deg = 0; h.ext_out = p == ENTER ? c.enter_ext : 0; h.ext_in = p == EXIT ? c.exit_ext : 0;
        h.weight = 0; h.cov = 0; h.cnt = 0; h.best = NIL;
        *(Hdr*)r = h;
        if (p >= 1 && p <= L) {
            OSlot s; s.dst = p + 1; s.cv = 0;
            *(OSlot*)(r + 32) = s;
        }
        if (p >= 1 && p <= L) *(u32*)(r + BB_IN_OFF) = p - 1;
    }
    if (tid == 0) {
        s_ins = nbb; s_status = T_OK;
        s_enter_out = 1; s_exit_in = 1;
        s_pool_top = c.exit_ext + (c.A + 2 + 1) / 2 + 1;
        if (s_pool_top > c.pool_cap) s_status = T_INTERNAL;
        else {
            c.pool[c.enter_ext] = c.A + 2;
            c.pool[c.exit_ext] = c.A + 2;
            OSlot s; s.dst = 1; s.cv = 0;                 // enter -> 1 (or -> exit when L == 0)
            *(OSlot*)(c.pool + c.enter_ext + 1) = s;
            *(u32*)(c.pool + c.exit_ext + 1) = L;          // L -> exit
        }
    }
    __syncthreads();
    if (s_status != T_OK) { if (tid == 0) a.tgt_status_out[t] = s_status; return; }

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
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                const u32 p = bbpos[j];
                if (kind[j] == 1 || kind[j] == 2) {
                    c.acc.cov[p]++;                 // _g[_bbMap[curr]].coverage++
                    c.acc.b[p] = (u8)tc[j];         // .base = targetBase (the last alignment in order wins)
                    if (kind[j] == 1) c.acc.w[p]++;
                } else if (kind[j] == 3) {
                    Hdr h;
                    h.bfv = q[j]; h.deg = 0x10001u; h.ext_out = 0; h.ext_in = 0; h.weight = 1; h.cov = (int)p; h.cnt = 0; h.best = NIL;
                    *(Hdr*)(c.insrec + (size_t)INS_REC * (cur[j] - nbb)) = h;
                }
                if (kind[j] == 1 || kind[j] == 3) {
                    b2_edge(c, prev, cur[j]);
                    prev = cur[j];
                }
            }
            __syncthreads();
            if (tid == 0) {
                s_carry_t += tot_s & 0xFFFFu;
                s_carry_i += tot_s >> 16;
                if (tot_m) s_last = s_cur[tot_m - 1];
            }
            __syncthreads();
        }
        if (tid == 0) {  // addEdge(prevVtx, _exitVtx) :106
            b2_edge(c, s_last, EXIT);
            s_ins += s_carry_i;
        }
        __syncthreads();
    }
    // fold the running sums into the backbone headers, then give every insertion vertex the coverage of the
    // backbone vertex it maps to (bestPath reads _g[_bbMap[v]].coverage, :407)
    const u8* bbstr = a.tgt_bb_off ? a.arena + a.tgt_bb_off[t] : nullptr;
    for (u32 p = tid; p < nbb; p += BT) {
        Hdr* hp = (Hdr*)(c.bbrec + (size_t)BB_REC * p);
        Hdr h = *hp;
        const u32 cv = c.acc.cov[p];
        u32 b = p == ENTER ? (u32)'^' : p == EXIT ? (u32)'$' : (bbstr ? (u32)bbstr[p - 1] : (u32)'N');
        if (cv) b = c.acc.b[p];
        const bool inner = p >= 1 && p <= L;
        h.bfv = b | (NF_BACKBONE << 8);
        const u32 dout = p == ENTER ? s_enter_out : c.acc.dout[p], din = p == EXIT ? s_exit_in : c.acc.din[p];
        h.deg = din | (dout << 16);
        h.weight = (inner ? 1 : 0) + (int)c.acc.w[p];
        h.cov = (int)cv;
        *hp = h;
        if (p <= L) {
            OSlot* o = h.ext_out ? (OSlot*)(c.pool + h.ext_out + 1) : (OSlot*)((u8*)hp + 32);
            o[0].cv = c.acc.c[p];
        }
    }
    const u32 nv = s_ins;
    for (u32 v = nbb + tid; v < nv; v += BT) {
        Hdr* hv = (Hdr*)(c.insrec + (size_t)INS_REC * (v - nbb));
        const u32 bm = (u32)hv->cov;
        hv->cov = bm < nbb ? (int)c.acc.cov[bm] : 0;
    }
    if (tid == 0) {
        a.tgt_nv[t] = nv;
        a.tgt_pool_top[t] = s_pool_top;
        if (s_status != T_OK) a.tgt_status_out[t] = s_status;
    }
}

// ------------------------------------------------------------------------------------------
// K3: mergeNodes (:129-160) -- the order-sensitive part; one scalar worker per target (see serial_setup).
struct Serial2Args {
    u32 first, count, tpw;
    u32 flags, min_weight, min_len;
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    u32* tgt_status;
    const u64* bb_off;
    const u64* ins_off;
    const u64* pool_off;
    const u32* pool_cap;
    u32* tgt_pool_top;
    const u32* tgt_nv;
    u8* bb;
    u8* ins;
    u64* pool;
    u32* gscratch;          // overflow of the per-worker group snapshot stack
    const u64* gscratch_off;
    const u32* gscratch_cap;
    u8* cons;
    const u64* cons_off;
    int* frags;
    const u64* frag_off;
    const u32* frag_cap;
    u32* tgt_cons_len;
    u32* tgt_nfrag;
};

struct T2Ctx {
    GV g;
    u32* S;
    u32 scap;
    u32* gS;
    u32 gscap;
};
__device__ __forceinline__ u32 S2_get(const T2Ctx& c, u32 i) { return i < c.scap ? c.S[i] : c.gS[i - c.scap]; }
__device__ __forceinline__ bool S2_put(T2Ctx& c, u32 i, u32 v) {
    if (i < c.scap) { c.S[i] = v; return true; }
    if (i - c.scap < c.gscap) { c.gS[i - c.scap] = v; return true; }
    c.g.status = T_INTERNAL;
    return false;
}

// Snapshot of the merge candidates of n, in list order (the std::map<char, std::vector<VtxDesc>> built at
// :167-181).  Frame: [n, cnt, last key + 129, parent frame, members...].
__device__ __forceinline__ u32 m2_push_in_frame(T2Ctx& c, u32 n, u32& top, u32 parent) {
    const Hdr* h = gv_hdr(c.g, n);
    const u32 nin = h->deg & 0xFFFFu;
    if (nin < 2) return NIL;
    const u32* in = gv_in(c.g, n, h->ext_in);
    u32 cnt = 0;
    for (u32 i = 0; i < nin; i++) {
        const u32 s = in[i];
        if ((gv_hdr(c.g, s)->deg >> 16) == 1) { if (!S2_put(c, top + 4 + cnt, s)) return NIL; cnt++; }
    }
    if (cnt < 2) return NIL;
    const u32 fb = top;
    S2_put(c, fb, n); S2_put(c, fb + 1, cnt); S2_put(c, fb + 2, 0); S2_put(c, fb + 3, parent);
    top += 4 + cnt;
    return fb;
}

// next key in ascending (signed) char order; false when the map is exhausted
__device__ __forceinline__ bool m2_next_group(T2Ctx& c, u32 fb, u32& an, u32& members) {
    const u32 cnt = S2_get(c, fb + 1);
    const int last = (int)S2_get(c, fb + 2) - 129;
    int key = 1000;
    for (u32 i = 0; i < cnt; i++) {
        const int b = (int)(signed char)n_base(gv_hdr(c.g, S2_get(c, fb + 4 + i))->bfv);
        if (b > last && b < key) key = b;
    }
    if (key == 1000) return false;
    S2_put(c, fb + 2, (u32)(key + 129));
    an = NIL; members = 0;
    for (u32 i = 0; i < cnt; i++) {
        const u32 ni = S2_get(c, fb + 4 + i);
        if ((int)(signed char)n_base(gv_hdr(c.g, ni)->bfv) == key) { if (an == NIL) an = ni; members++; }
    }
    return true;
}

// first out-slot / single out-slot of a vertex
__device__ __forceinline__ OSlot* m2_out0(const GV& g, u32 v) { return gv_out(g, v, gv_hdr(g, v)->ext_out); }

// mergeInNodes :162-215 (the recursion :212 runs on the snapshot stack)
__device__ void m2_merge_in(T2Ctx& c, u32 n0) {
    GV& g = c.g;
    u32 top = 0;
    u32 fb = m2_push_in_frame(c, n0, top, NIL);
    while (fb != NIL && g.status == T_OK) {
        u32 an, members;
        if (!m2_next_group(c, fb, an, members)) { top = fb; fb = S2_get(c, fb + 3); continue; }
        if (members < 2) continue;
        const u32 cnt = S2_get(c, fb + 1);
        const int key = (int)S2_get(c, fb + 2) - 129;
        OSlot* an_out = m2_out0(g, an);  // *anoi, taken once (:186-188)
        Hdr* han = gv_hdr(g, an);
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S2_get(c, fb + 4 + i);
            if (ni == an || (int)(signed char)n_base(gv_hdr(g, ni)->bfv) != key) continue;
            if ((gv_hdr(g, ni)->deg >> 16) == 0) { g.status = T_INTERNAL; return; }
            an_out->cv += m2_out0(g, ni)->cv & ~VISBIT;
            han->weight += gv_hdr(g, ni)->weight;
        }
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S2_get(c, fb + 4 + i);
            const Hdr* hni = gv_hdr(g, ni);
            if (ni == an || (int)(signed char)n_base(hni->bfv) != key) continue;
            const u32 nin = hni->deg & 0xFFFFu;
            for (u32 k = 0; k < nin; k++) {
                // the in-list of ni is stable here: edges are only added towards `an`
                const u32 n1 = gv_in(g, ni, hni->ext_in)[k];
                Hdr* h1 = gv_hdr(g, n1);
                OSlot* o1 = gv_out(g, n1, h1->ext_out);
                const u32 d1 = h1->deg >> 16;
                const u32 e_ni = gv_find_out(o1, d1, ni);
                if (e_ni == NIL) { g.status = T_INTERNAL; return; }
                const u32 cvn = o1[e_ni].cv;
                const u32 e_an = gv_find_out(o1, d1, an);
                if (e_an != NIL) o1[e_an].cv += cvn & ~VISBIT;
                else gv_add_edge(g, n1, an, cvn & ~VISBIT, cvn & VISBIT);
            }
            gv_reap(g, ni);
        }
        const u32 nfb = m2_push_in_frame(c, an, top, fb);  // mergeInNodes(an) :212
        if (nfb != NIL) fb = nfb;
    }
}

// mergeOutNodes :217-267 for one std::map key among the out-neighbours snapshotted in sd/sx/sy
__device__ __forceinline__ void m2_merge_out_group(T2Ctx& c, u32 key, u32 cnt, const u32* sd, const u32* sx, const u32* sy) {
    GV& g = c.g;
    u32 an = NIL;
    Hdr* han = nullptr;
    for (u32 i = 0; i < cnt; i++) {
        if ((sy[i] & 0xFFFFu) != 1 || (sx[i] & 0xFFu) != key) continue;
        const u32 ni = sd[i];
        if (an == NIL) { an = ni; han = gv_hdr(g, an); continue; }
        Hdr* hni = gv_hdr(g, ni);
        // _g[*anii].count += _g[*ii].count (:241): the first in-edge of each; an edge's count lives in its source's slot
        {
            if ((han->deg & 0xFFFFu) == 0 || (hni->deg & 0xFFFFu) == 0) { g.status = T_INTERNAL; return; }
            OSlot* ea = gv_edge_slot(g, gv_in(g, an, han->ext_in)[0], an);
            OSlot* ei = gv_edge_slot(g, gv_in(g, ni, hni->ext_in)[0], ni);
            if (!ea || !ei) { g.status = T_INTERNAL; return; }
            ea->cv += ei->cv & ~VISBIT;
        }
        han->weight += hni->weight;
        const u32 nout = hni->deg >> 16;
        for (u32 k = 0; k < nout; k++) {
            const OSlot s = gv_out(g, ni, hni->ext_out)[k];
            OSlot* oa = gv_out(g, an, han->ext_out);
            const u32 e = gv_find_out(oa, han->deg >> 16, s.dst);
            if (e != NIL) oa[e].cv += s.cv & ~VISBIT;
            else gv_add_edge(g, an, s.dst, s.cv & ~VISBIT, s.cv & VISBIT);
        }
        gv_reap(g, ni);
    }
}

#define M2_MAXK 60  // out-neighbours snapshotted per popped vertex (4 arrays of 60 words + spare = 256 words per worker)

// One pass over the out-list of u: sd = target, sc = slot cv, sx = target bfv, sy = target deg.  The loads of the
// targets' headers are independent, so they overlap; they also pull the successors' records on chip before the
// sweep reaches them.  dup_lo/dup_hi: bases (0..63 / 64..127) that occur more than once among the merge
// candidates (targets with in_degree 1); bases >= 128 (negative as `char`) send the vertex down the generic path.
struct OutScan2 { u32 cnt; u64 dup_lo, dup_hi; bool weird; };
__device__ __forceinline__ OutScan2 m2_scan_out(const GV& g, u32 u, const Hdr& hu, u32* sd, u32* sc, u32* sx, u32* sy) {
    OutScan2 r;
    r.cnt = hu.deg >> 16; r.dup_lo = 0; r.dup_hi = 0; r.weird = false;
    if (r.cnt > M2_MAXK) { r.cnt = TOO_BIG; return r; }
    const OSlot* o = gv_out(g, u, hu.ext_out);
    for (u32 i = 0; i < r.cnt; i++) { const OSlot s = o[i]; sd[i] = s.dst; sc[i] = s.cv; }
    for (u32 i = 0; i < r.cnt; i++) { const uint2 d = *(const uint2*)gv_hdr(g, sd[i]); sx[i] = d.x; sy[i] = d.y; }
    u64 seen_lo = 0, seen_hi = 0;
    for (u32 i = 0; i < r.cnt; i++) {
        if ((sy[i] & 0xFFFFu) == 1) {
            const u32 b = sx[i] & 0xFFu;
            const u64 m = 1ull << (b & 63);
            if (b < 64) { r.dup_lo |= seen_lo & m; seen_lo |= m; }
            else if (b < 128) { r.dup_hi |= seen_hi & m; seen_hi |= m; }
            else r.weird = true;
        }
    }
    return r;
}
// does any std::map group of mergeInNodes(n) hold more than one vertex (predecessors with out_degree 1 sharing a base)?
__device__ __forceinline__ bool m2_in_group_exists(const GV& g, u32 u, const Hdr& hu) {
    const u32 nin = hu.deg & 0xFFFFu;
    const u32* in = gv_in(g, u, hu.ext_in);
    u64 seen_lo = 0, seen_hi = 0;
    bool found = false;
    for (u32 i = 0; i < nin; i++) {
        const uint2 d = *(const uint2*)gv_hdr(g, in[i]);
        if ((d.y >> 16) == 1) {
            const u32 b = d.x & 0xFFu;
            const u64 m = 1ull << (b & 63);
            if (b < 64) { found |= (seen_lo & m) != 0; seen_lo |= m; }
            else if (b < 128) { found |= (seen_hi & m) != 0; seen_hi |= m; }
            else found = true;  // let the generic code sort it out
        }
    }
    return found;
}

// generic mergeOutNodes (:217-267) for very wide vertices / exotic bases: snapshot on the stack like mergeInNodes
__device__ void m2_merge_out_generic(T2Ctx& c, u32 n) {
    GV& g = c.g;
    const Hdr* hn = gv_hdr(g, n);
    const u32 nout = hn->deg >> 16;
    if (nout < 2) return;
    u32 cnt = 0;
    for (u32 i = 0; i < nout; i++) {
        const u32 d = gv_out(g, n, hn->ext_out)[i].dst;
        if ((gv_hdr(g, d)->deg & 0xFFFFu) == 1) { if (!S2_put(c, 4 + cnt, d)) return; cnt++; }
    }
    if (cnt < 2) return;
    S2_put(c, 0, n); S2_put(c, 1, cnt); S2_put(c, 2, 0); S2_put(c, 3, NIL);
    u32 an, members;
    while (g.status == T_OK && m2_next_group(c, 0, an, members)) {
        if (members < 2) continue;
        const int key = (int)S2_get(c, 2) - 129;
        Hdr* han = gv_hdr(g, an);
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S2_get(c, 4 + i);
            Hdr* hni = gv_hdr(g, ni);
            if (ni == an || (int)(signed char)n_base(hni->bfv) != key) continue;
            if ((han->deg & 0xFFFFu) == 0 || (hni->deg & 0xFFFFu) == 0) { g.status = T_INTERNAL; return; }
            OSlot* ea = gv_edge_slot(g, gv_in(g, an, han->ext_in)[0], an);
            OSlot* ei = gv_edge_slot(g, gv_in(g, ni, hni->ext_in)[0], ni);
            if (!ea || !ei) { g.status = T_INTERNAL; return; }
            ea->cv += ei->cv & ~VISBIT;
            han->weight += hni->weight;
            const u32 no = hni->deg >> 16;
            for (u32 k = 0; k < no; k++) {
                const OSlot s = gv_out(g, ni, hni->ext_out)[k];
                OSlot* oa = gv_out(g, an, han->ext_out);
                const u32 e = gv_find_out(oa, han->deg >> 16, s.dst);
                if (e != NIL) oa[e].cv += s.cv & ~VISBIT;
                else gv_add_edge(g, an, s.dst, s.cv & ~VISBIT, s.cv & VISBIT);
            }
            gv_reap(g, ni);
        }
    }
}

// mergeNodes :129-160
__device__ void m2_merge_nodes(T2Ctx& c, u32 ENTER) {
    GV& g = c.g;
    u32* sd = c.S;
    u32* sc = sd + M2_MAXK;
    u32* sx = sc + M2_MAXK;
    u32* sy = sx + M2_MAXK;
    u32 head = ENTER, tail = ENTER;
    gv_hdr(g, ENTER)->best = NIL;
    gv_hdr(g, ENTER)->bfv |= NF_QUEUED << 8;
    while (head != NIL && g.status == T_OK) {
        const u32 u = head;
        Hdr* pu = gv_hdr(g, u);
        Hdr hu = *pu;
        if ((hu.deg & 0xFFFFu) >= 2 && m2_in_group_exists(g, u, hu)) {
            m2_merge_in(c, u);
            hu = *pu;
        }
        OutScan2 sc2;
        sc2.cnt = 0; sc2.dup_lo = 0; sc2.dup_hi = 0; sc2.weird = false;
        if (hu.deg >> 16) sc2 = m2_scan_out(g, u, hu, sd, sc, sx, sy);
        if (sc2.cnt == TOO_BIG || sc2.weird) {  // very wide vertex (enter) or exotic bases: generic path
            m2_merge_out_generic(c, u);
            hu = *pu;
            const u32 nout = hu.deg >> 16;
            for (u32 i = 0; i < nout && g.status == T_OK; i++) {
                OSlot* o = gv_out(g, u, hu.ext_out) + i;
                const u32 v = o->dst;
                Hdr* hv = gv_hdr(g, v);
                if (!(o->cv & VISBIT)) { o->cv |= VISBIT; hv->bfv += 0x10000u; }
                const u32 bfv = hv->bfv;
                if (n_vis(bfv) == (hv->deg & 0xFFFFu)) {  // notVisited == 0
                    if (n_flags(bfv) & NF_QUEUED) { g.status = T_INTERNAL; break; }
                    hv->bfv = bfv | (NF_QUEUED << 8);
                    gv_hdr(g, tail)->best = v;
                    tail = v;
                }
            }
            head = pu->best;
            continue;
        }
        if (sc2.dup_lo | sc2.dup_hi) {
            // groups in ascending key order, as std::map<char, ...> iterates them
            for (u64 m = sc2.dup_lo; m && g.status == T_OK; m &= m - 1) m2_merge_out_group(c, (u32)__ffsll((long long)m) - 1, sc2.cnt, sd, sx, sy);
            for (u64 m = sc2.dup_hi; m && g.status == T_OK; m &= m - 1) m2_merge_out_group(c, 63u + (u32)__ffsll((long long)m), sc2.cnt, sd, sx, sy);
            hu = *pu;
            sc2 = m2_scan_out(g, u, hu, sd, sc, sx, sy);  // merging only removed out-edges of u (and changed their targets)
        }
        OSlot* o = gv_out(g, u, hu.ext_out);
        for (u32 i = 0; i < sc2.cnt; i++) {
            const u32 v = sd[i];
            u32 bfv = sx[i];
            if (!(sc[i] & VISBIT)) { o[i].cv = sc[i] | VISBIT; bfv += 0x10000u; }
            if ((bfv >> 16) == (sy[i] & 0xFFFFu)) {  // notVisited == 0
                if (bfv & (NF_QUEUED << 8)) { g.status = T_INTERNAL; break; }
                bfv |= NF_QUEUED << 8;
                gv_hdr(g, tail)->best = v;   // hdr(v).best is still NIL from the build
                tail = v;
            }
            if (bfv != sx[i]) gv_hdr(g, v)->bfv = bfv;
        }
        head = pu->best;
    }
}

__device__ __forceinline__ bool serial2_setup(const Serial2Args& a, u32& t, T2Ctx& c, u32* smem, u32 scap_total) {
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
    c.S = smem ? smem + w * c.scap : nullptr;
    c.g.bb = a.bb + (size_t)BB_REC * a.bb_off[t];
    c.g.ins = a.ins + (size_t)INS_REC * a.ins_off[t];
    c.g.pool = a.pool + a.pool_off[t];
    c.g.nbb = a.tgt_tlen[t] + 2;
    c.g.pool_top = a.tgt_pool_top[t];
    c.g.pool_cap = a.pool_cap[t];
    c.g.status = T_OK;
    c.gS = a.gscratch + a.gscratch_off[t];
    c.gscap = a.gscratch_cap[t];
    return true;
}

__global__ void __launch_bounds__(SERIAL_THREADS) k_merge_nodes2(Serial2Args a) {
    __shared__ u32 smem[SERIAL_SMEM_WORDS];
    u32 t;
    T2Ctx c;
    if (!serial2_setup(a, t, c, smem, SERIAL_SMEM_WORDS)) return;
    m2_merge_nodes(c, 0);
    a.tgt_pool_top[t] = c.g.pool_top;
    if (c.g.status != T_OK) a.tgt_status[t] = c.g.status;
}

// ------------------------------------------------------------------------------------------
// K4: bestPath (:375-459) and the two consensus() forms (:285-373); one scalar worker per target.
__global__ void __launch_bounds__(SERIAL_THREADS) k_best_path2(Serial2Args a) {
    u32 t;
    T2Ctx c;
    if (!serial2_setup(a, t, c, nullptr, SERIAL_SMEM_WORDS)) return;
    const GV& g = c.g;
    const u32 L = a.tgt_tlen[t], ENTER = 0, EXIT = L + 1, nV = a.tgt_nv[t];
    // reverse topological sweep from exit; Hdr.cnt counts a vertex's scored successors until it is scored itself
    u32 head = EXIT, tail = EXIT;
    gv_hdr(g, EXIT)->best = NIL;
    while (head != NIL) {
        const u32 n = head;
        Hdr* pn = gv_hdr(g, n);
        const Hdr hn = *pn;
        float best_score = -FLT_MAX;
        u32 best_e = NIL;
        const OSlot* o = gv_out(g, n, hn.ext_out);
        const u32 nout = hn.deg >> 16;
        for (u32 i = 0; i < nout; i++) {
            const OSlot s = o[i];
            const Hdr hd = *gv_hdr(g, s.dst);
            float ns;
            // bestPath :402-409: a backbone vertex no read agrees with costs 10; otherwise count - coverage/2
            if ((hd.bfv & (NF_BACKBONE << 8)) && hd.weight == 1) ns = __fsub_rn(hd.score, 10.0f);
            else ns = __fadd_rn(__fsub_rn((float)(int)(s.cv & ~VISBIT), __fmul_rn((float)hd.cov, 0.5f)), hd.score);
            if (ns > best_score) { best_score = ns; best_e = i; }  // strict >: the first maximum in list order wins
        }
        const u32* in = gv_in(g, n, hn.ext_in);
        const u32 nin = hn.deg & 0xFFFFu;
        for (u32 i = 0; i < nin; i++) {
            const u32 s = in[i];
            Hdr* hs = gv_hdr(g, s);
            const u32 k = (u32)(++hs->cnt);
            if (k == (hs->deg >> 16)) {  // all out-edges lead to scored vertices
                hs->best = NIL;
                gv_hdr(g, tail)->best = s;
                tail = s;
            }
        }
        const u32 nx = pn->best;
        pn->score = best_e != NIL ? best_score : 0.0f;  // an absent map key reads as 0.0f (:401)
        pn->best = best_e;
        pn->bfv = hn.bfv | (NF_SCORED << 8);
        head = nx;
    }
    // walk the best path and cut it into min-weight runs
    u8* cons = a.cons + a.cons_off[t];
    int* fr = a.frags + 2 * a.frag_off[t];
    const u32 fcap = a.frag_cap[t];
    const u32 eb = n_base(gv_hdr(g, ENTER)->bfv), xb = n_base(gv_hdr(g, EXIT)->bfv);
    const int minw = (int)a.min_weight;
    const bool single = a.flags & F_SINGLE;
    int offs = 0, idx = 0, best_offs = 0, length = 0;
    bool met = false;
    u32 nfrag = 0, steps = 0, prev = ENTER, status = T_OK;
    for (;;) {
        const Hdr hp = *gv_hdr(g, prev);
        const u32 b = hp.bfv & 0xFFu;
        const int w = hp.weight;
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
        if (!(hp.bfv & (NF_SCORED << 8)) || hp.best == NIL) break;
        prev = gv_out(g, prev, hp.ext_out)[hp.best].dst;
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
