// graph_rec.cuh -- the alignment graph as fat vertex records (K2 build, K3 mergeNodes, K4 bestPath/consensus).
//
// Reference semantics (verdurin/pbdagcon @ 3c382f2): src/cpp/AlnGraphBoost.cpp
//   ctor :16-62, addAln :64-107, addEdge :109-127, mergeNodes :129-160, mergeInNodes :162-215,
//   mergeOutNodes :217-267, markForReaper :269-273, consensus :285-373, bestPath :375-459.
//
// Layout.  A vertex IS its record; an edge lives in its source's out-slot (dst, count, visited) and is mirrored
// by a 4-byte in-slot (src) at its target -- there is no edge array.  (u,v) identifies an edge because neither
// addEdge nor the merges ever create a parallel edge (:114-121, :194-203, :248-257).
// One target's graph is one arena of 128-byte units followed by a pool of 8-byte words:
//   insertion vertex v >= L+2 (creation order)   1 unit : Hdr | 12 in-slots | 6 out-slots          handle v + (L+2)
//   backbone vertex p (0 enter, 1..L, L+1 exit)   2 units: Hdr | 12 in-slots | 22 out-slots         handle 2p
// A handle is the unit index, so a record's address is base + 128*handle for both kinds; slots hold handles.
// Slots keep the adjacency-list ORDER the reference's tie-breaks depend on: append = write at [deg], order
// preserving erase = shift down.  Hdr.out_ptr / Hdr.in_ptr give the list's position in 8-byte words from the
// arena base: the inline slots at first, a pool block (word -1 = capacity) once a list outgrows them.
// A vertex and its lists thus sit in one or two 128-byte lines, a warp reads a whole list with one load, and
// the neighbours' headers with one more.
#pragma once

#define REC_UNIT 128u
#define REC_IN_OFF 32u
#define REC_OUT_OFF 80u
#define REC_IN_CAP 12u
#define INS_OUT_CAP 6u
#define BB_OUT_CAP 22u
#define VISBIT 0x80000000u

struct __align__(16) Hdr {
    u32 bfv;      // base | flags << 8 | visited-in-edge count << 16
    u32 deg;      // in_deg | out_deg << 16
    u32 out_ptr;  // 8-byte word index (from the arena base) of out-slot 0
    u32 in_ptr;   // 8-byte word index of in-slot 0
    int weight;
    int cov;      // coverage of _bbMap[v] (bestPath, :407)
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

__host__ __device__ __forceinline__ u32 gv_handle(u32 v, u32 nbb) { return v < nbb ? 2u * v : v + nbb; }
__host__ __device__ __forceinline__ u32 gv_vertex(u32 h, u32 nbb) { return h < 2u * nbb ? h >> 1 : h - nbb; }
__host__ __device__ __forceinline__ u32 gv_inline_out(u32 h) { return 16u * h + REC_OUT_OFF / 8; }
__host__ __device__ __forceinline__ u32 gv_inline_in(u32 h) { return 16u * h + REC_IN_OFF / 8; }

struct GV {           // one target's graph
    u8* base;
    u32 nbb;          // L + 2
    u32 pool_top, pool_end;  // 8-byte word indices from base
    u32 status;
    u32 nunits;       // record units in front of the pool: handles are < nunits; 0 switches the prefetches off
    u32 pool_failed;  // the pool ran dry: the target is rebuilt with a worst-case pool (T_POOL), whatever else went wrong after
};

// The two sweeps (mergeNodes forwards, bestPath backwards) are chains of dependent loads: a vertex is first seen as
// a neighbour's slot (header needed at once), popped a step or two later (its own slots needed), and its neighbours'
// headers after that.  Each of those was a separate DRAM miss (ncu r01b: 45 % of the stall samples on three loads).
// So whenever a lane first touches a vertex it asks L2 for the whole record and, once the record's first slots have
// arrived, for the records those slots name: the chain then runs one step behind its own prefetches.
// Into L1 as well: the warp that asks is the only reader of its target's records (measured: 256 targets 74.7 -> 69.7 ms,
// 10 000 targets unchanged against prefetch.global.L2; -DGV_PF_L2_ONLY restores that form).
#ifdef GV_PF_L2_ONLY
__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }
#else
__device__ __forceinline__ void pf_l2(const void* p) { asm volatile("prefetch.global.L1 [%0];" ::"l"(p)); }
#endif
__device__ __forceinline__ void gv_prefetch_rec(const GV& g, u32 h) {
    if (h < g.nunits) {  // a slot past a list's length holds stale bytes: never form an address outside the arena
        const u8* r = g.base + (size_t)h * REC_UNIT;
        pf_l2(r);
        pf_l2(r + 64);
    }
}

__device__ __forceinline__ Hdr* gb_hdr(u8* base, u32 h) { return (Hdr*)(base + (size_t)h * REC_UNIT); }
__device__ __forceinline__ OSlot* gb_out(u8* base, u32 out_ptr) { return (OSlot*)(base + (size_t)out_ptr * 8); }
__device__ __forceinline__ u32* gb_in(u8* base, u32 in_ptr) { return (u32*)(base + (size_t)in_ptr * 8); }
__device__ __forceinline__ Hdr* gv_hdr(const GV& g, u32 h) { return (Hdr*)(g.base + (size_t)h * REC_UNIT); }
__device__ __forceinline__ OSlot* gv_out(const GV& g, u32 out_ptr) { return (OSlot*)(g.base + (size_t)out_ptr * 8); }
__device__ __forceinline__ u32* gv_in(const GV& g, u32 in_ptr) { return (u32*)(g.base + (size_t)in_ptr * 8); }
__device__ __forceinline__ u32 gv_out_cap(const GV& g, u32 h, u32 out_ptr) {
    return out_ptr == gv_inline_out(h) ? (h < 2u * g.nbb ? BB_OUT_CAP : INS_OUT_CAP) : (u32)((const u64*)g.base)[out_ptr - 1];
}
__device__ __forceinline__ u32 gv_in_cap(const GV& g, u32 h, u32 in_ptr) {
    return in_ptr == gv_inline_in(h) ? REC_IN_CAP : (u32)((const u64*)g.base)[in_ptr - 1];
}
// a pool block of `words` 8-byte words behind one capacity word; returns the index of its first data word or 0
__device__ __forceinline__ u32 gv_alloc(GV& g, u32 words, u32 cap) {
    if (g.pool_top + words + 1 > g.pool_end) { g.status = T_POOL; g.pool_failed = 1; return 0; }
    const u32 off = g.pool_top;
    g.pool_top += words + 1;
    ((u64*)g.base)[off] = cap;
    return off + 1;
}

// boost::edge(u, v, g): first match in out(u); returns the slot index or NIL
__device__ __forceinline__ u32 gv_find_out(const OSlot* o, u32 n, u32 v) {
#pragma unroll 1
    for (u32 i = 0; i < n; i++)
        if (o[i].dst == v) return i;
    return NIL;
}
__device__ __forceinline__ OSlot* gv_edge_slot(const GV& g, u32 u, u32 v) {
    const Hdr* hu = gv_hdr(g, u);
    OSlot* o = gv_out(g, hu->out_ptr);
    const u32 i = gv_find_out(o, hu->deg >> 16, v);
    return i == NIL ? nullptr : o + i;
}

// boost::add_edge(u, v) with the AlnEdge properties copied (:198-203, :252-257); u's and v's headers are updated
__device__ __noinline__ void gv_add_edge(GV& g, u32 u, u32 v, u32 count, u32 visited) {
    Hdr* hu = gv_hdr(g, u);
    {
        const u32 n = hu->deg >> 16;
        OSlot* o = gv_out(g, hu->out_ptr);
        if (n == gv_out_cap(g, u, hu->out_ptr)) {
            const u32 ncap = max(8u, 2u * n);
            const u32 blk = gv_alloc(g, ncap, ncap);
            if (!blk) return;
            OSlot* no = gv_out(g, blk);
#pragma unroll 1
            for (u32 i = 0; i < n; i++) no[i] = o[i];
            hu->out_ptr = blk;
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
        u32* in = gv_in(g, hv->in_ptr);
        if (n == gv_in_cap(g, v, hv->in_ptr)) {
            const u32 ncap = max(16u, 2u * n);
            const u32 blk = gv_alloc(g, ncap / 2, ncap);
            if (!blk) return;
            u32* ni = gv_in(g, blk);
#pragma unroll 1
            for (u32 i = 0; i < n; i++) ni[i] = in[i];
            hv->in_ptr = blk;
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
    u32* in = gv_in(g, hv->in_ptr);
    u32 i = 0;
#pragma unroll 1
    while (i < n && in[i] != u) i++;
    if (i == n) { g.status = T_INTERNAL; return; }
#pragma unroll 1
    for (; i + 1 < n; i++) in[i] = in[i + 1];
    hv->deg -= 1u;
    if (was_visited) hv->bfv -= 0x10000u;
}
__device__ __forceinline__ void gv_erase_out(GV& g, u32 u, u32 v) {
    Hdr* hu = gv_hdr(g, u);
    const u32 n = hu->deg >> 16;
    OSlot* o = gv_out(g, hu->out_ptr);
    u32 i = gv_find_out(o, n, v);
    if (i == NIL) { g.status = T_INTERNAL; return; }
#pragma unroll 1
    for (; i + 1 < n; i++) o[i] = o[i + 1];
    hu->deg -= 0x10000u;
}

// markForReaper (:269-273): deleted = true; clear_vertex(n)
__device__ __noinline__ void gv_reap(GV& g, u32 ni) {
    Hdr* h = gv_hdr(g, ni);
    const u32 deg = h->deg;
    const OSlot* o = gv_out(g, h->out_ptr);
#pragma unroll 1
    for (u32 i = 0; i < (deg >> 16); i++) {
        const OSlot s = o[i];
        if (s.dst != ni) gv_erase_in(g, s.dst, ni, (s.cv & VISBIT) != 0);
    }
    const u32* in = gv_in(g, h->in_ptr);
#pragma unroll 1
    for (u32 i = 0; i < (deg & 0xFFFFu); i++) {
        const u32 s = in[i];
        if (s != ni) gv_erase_out(g, s, ni);
    }
    h->deg = 0;
    h->bfv = (h->bfv & 0xFFFFu) | (NF_DELETED << 8);
}

// ------------------------------------------------------------------------------------------
// K2: graph construction (ctor :16-62, addAln :64-107, addEdge :109-127), one 512-thread block per target.
// Alignments are replayed in input order (every adjacency list is ordered by the alignment that first created
// each edge); inside one alignment every vertex is entered and left at most once, so its columns are processed in
// parallel, 1024 per tile (512 threads x 2): block scans give each column its backbone position, its insertion-vertex id and its
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
    const u64* rec_off;   // first 128-byte unit of the target's arena
    const u32* tgt_V;     // planned vertex count: the pool starts behind (L+2) + V units
    const u32* pool_cap;  // 8-byte words
    const u64* acc_off;   // positions before this target (global accumulator fallback)
    const u16* norm;
    const u64* a_off;
    const u32* a_len;
    const u32* a_start;
    const u32* a_flags;
    const u32* a_tkept;   // backbone positions each alignment consumes (normalise kernels): coverage is known before the replay
    u8* rec;
    u32* tgt_status_out;
    u32* tgt_nv;
    u32* tgt_pool_top;
    u32 smem_positions;
    u8* gacc;  // 9 bytes per position
};

// DT = type of the per-position list lengths: u8 (9 bytes per position; a backbone vertex has at most A + 1 in- or
// out-edges, so this is exact for targets of up to 252 alignments) or u16 (11 bytes per position; any target the
// 16-bit degrees of the records can hold, i.e. up to 65 000 alignments).  The dispatcher picks the variant per launch.
template <typename DT>
struct BuildAcc {  // per-position running state of the replay
    u16* cov;
    u16* w;
    u16* c;
    u8* b;
    DT* dout;
    DT* din;
};
template <typename DT> __host__ __device__ constexpr u32 acc_max_deg() { return sizeof(DT) == 1 ? 254u : 65534u; }
template <typename DT> __host__ __device__ constexpr u32 acc_bytes_per_pos() { return 7u + 2u * (u32)sizeof(DT); }

template <typename DT>
struct B2Ctx {
    u8* base;
    u32 L, A, nbb, enter_out, exit_in, pool_end;  // enter_out / exit_in: word index of slot 0 of the two wide lists
    u32* s_pool_top;
    u32* s_status;
    u32* s_enter_out;
    u32* s_exit_in;
    BuildAcc<DT> acc;
};
template <typename DT>
__device__ __forceinline__ Hdr* b2_hdr(const B2Ctx<DT>& c, u32 h) { return (Hdr*)(c.base + (size_t)h * REC_UNIT); }

// append (p -> dst handle, count 1) to backbone vertex p's out-list; returns the slot it went to (NIL: it did not)
template <typename DT>
__device__ __forceinline__ u32 b2_append_out(const B2Ctx<DT>& c, u32 p, u32 dsth) {
    u32 n, ptr;
    if (p == 0) {
        n = (*c.s_enter_out)++;
        ptr = c.enter_out;
    } else {
        n = c.acc.dout[p];
        if (n >= acc_max_deg<DT>()) { *c.s_status = T_TOO_MANY; return NIL; }
        c.acc.dout[p] = (DT)(n + 1);
        ptr = n <= BB_OUT_CAP ? gv_inline_out(2 * p) : b2_hdr(c, 2 * p)->out_ptr;
        const u32 have = n <= BB_OUT_CAP ? BB_OUT_CAP : (u32)((const u64*)c.base)[ptr - 1];
        if (n == have) {  // full: the list moves into a pool block twice the size (a backbone vertex has at most A+1 out-edges)
            const u32 cap = min(2u * have, max(c.A + 2, have + 1)), blk = atomicAdd(c.s_pool_top, cap + 1);
            if (blk + cap + 1 > c.pool_end) { *c.s_status = T_POOL; c.acc.dout[p] = (DT)n; return NIL; }
            ((u64*)c.base)[blk] = cap;
            const OSlot* o = (const OSlot*)(c.base + (size_t)ptr * 8);
            OSlot* no = (OSlot*)(c.base + (size_t)(blk + 1) * 8);
            for (u32 i = 0; i < n; i++) no[i] = o[i];
            b2_hdr(c, 2 * p)->out_ptr = blk + 1;
            ptr = blk + 1;
        }
    }
    OSlot s;
    s.dst = dsth; s.cv = 1;
    ((OSlot*)(c.base + (size_t)ptr * 8))[n] = s;
    return n;
}
template <typename DT>
__device__ __forceinline__ void b2_append_in(const B2Ctx<DT>& c, u32 p, u32 srch) {
    u32 n, ptr;
    if (p == c.L + 1) {
        n = (*c.s_exit_in)++;
        ptr = c.exit_in;
    } else {
        n = c.acc.din[p];
        if (n >= acc_max_deg<DT>()) { *c.s_status = T_TOO_MANY; return; }
        c.acc.din[p] = (DT)(n + 1);
        ptr = n <= REC_IN_CAP ? gv_inline_in(2 * p) : b2_hdr(c, 2 * p)->in_ptr;
        const u32 have = n <= REC_IN_CAP ? REC_IN_CAP : (u32)((const u64*)c.base)[ptr - 1];
        if (n == have) {
            const u32 cap = min(2u * have, max(c.A + 2, have + 1)), words = (cap + 1) / 2, blk = atomicAdd(c.s_pool_top, words + 1);
            if (blk + words + 1 > c.pool_end) { *c.s_status = T_POOL; c.acc.din[p] = (DT)n; return; }
            ((u64*)c.base)[blk] = cap;
            const u32* in = (const u32*)(c.base + (size_t)ptr * 8);
            u32* ni = (u32*)(c.base + (size_t)(blk + 1) * 8);
            for (u32 i = 0; i < n; i++) ni[i] = in[i];
            b2_hdr(c, 2 * p)->in_ptr = blk + 1;
            ptr = blk + 1;
        }
    }
    ((u32*)(c.base + (size_t)ptr * 8))[n] = srch;
}

// addEdge(prev, cur) (:109-127) for one path step; prev/cur are vertex ids.  An insertion vertex on the path was
// created by the alignment being replayed: its single in-slot / out-slot is written directly.
template <typename DT>
__device__ __forceinline__ void b2_edge(const B2Ctx<DT>& c, u32 prev, u32 cur) {
    const u32 nbb = c.nbb;
    const bool prev_bb = prev < nbb, cur_bb = cur < nbb;
    const u32 ph = gv_handle(prev, nbb), ch = gv_handle(cur, nbb);
    if (prev_bb && cur_bb) {
        if (cur == prev + 1) { c.acc.c[prev]++; return; }  // the ctor edge p -> p+1
        // a backbone->backbone skip (deletion run, alignment start or end): it may exist already.  Searching prev's
        // out-list means reading a record that left the caches many alignments ago, with the whole block waiting at the
        // tile's barrier.  With byte-wide list lengths (<= 252 alignments) the ctor-edge count above stays below 256, and
        // its upper byte remembers prev's FIRST skip edge: (positions skipped) << 5 | slot, 0 = none yet, 0xFF = there is
        // one that does not fit this code.  A deletion of the same length at the same place -- the usual case -- is then
        // one fire-and-forget add to a known slot; prev has one successor per alignment, so the byte has one writer.
        if (sizeof(DT) == 1 && prev == 0) {
            // enter -> the vertex an alignment starts on: enter's out-list is as long as the target has alignments and sits in
            // the pool.  The upper byte of the START vertex's match weight (<= 252 as well; the thread that counts the first
            // column's match is the one that adds this edge) remembers the edge's slot + 1.
            const u32 ww = c.acc.w[cur], hi = ww >> 8;
            if (hi) { atomicAdd(&((OSlot*)(c.base + (size_t)c.enter_out * 8))[hi - 1].cv, 1u); return; }
            const u32 slot = b2_append_out(c, 0u, ch);
            b2_append_in(c, cur, ph);
            if (slot < 255u) { c.acc.w[cur] = (u16)(ww | ((slot + 1) << 8)); return; }
            return;   // (cannot happen: at most 252 alignments)
        }
        if (sizeof(DT) == 1 && prev != 0) {
            const u32 cw = c.acc.c[prev], hi = cw >> 8, k1 = cur - prev - 1;
            if (hi == 0) {   // no skip edge yet: the list holds the ctor edge and edges to insertion vertices only
                const u32 slot = b2_append_out(c, prev, ch);
                b2_append_in(c, cur, ph);
                c.acc.c[prev] = (u16)(cw | ((k1 <= 6 && slot <= 31) ? ((k1 << 5) | slot) << 8 : 0xFF00u));
                return;
            }
            if (hi != 0xFFu && (hi >> 5) == k1 && c.acc.dout[prev] <= BB_OUT_CAP) {   // still in the record itself
                atomicAdd(&((OSlot*)(c.base + (size_t)gv_inline_out(ph) * 8))[hi & 31u].cv, 1u);
                return;
            }
        }
        const u32 n = prev == 0 ? *c.s_enter_out : c.acc.dout[prev];
        const u32 ptr = prev == 0 ? c.enter_out : (n <= BB_OUT_CAP ? gv_inline_out(ph) : b2_hdr(c, ph)->out_ptr);
        OSlot* o = (OSlot*)(c.base + (size_t)ptr * 8);
#pragma unroll 1
        for (u32 i = 0; i < n; i += 4) {   // first match in list order; four independent loads per round trip
            const u32 i1 = min(i + 1, n - 1), i2 = min(i + 2, n - 1), i3 = min(i + 3, n - 1);
            const u32 d0 = o[i].dst, d1 = o[i1].dst, d2 = o[i2].dst, d3 = o[i3].dst;
            const u32 j = d0 == ch ? i : d1 == ch ? i1 : d2 == ch ? i2 : d3 == ch ? i3 : NIL;
            if (j != NIL) { o[j].cv++; return; }
        }
        b2_append_out(c, prev, ch);
        b2_append_in(c, cur, ph);
        return;
    }
    if (prev_bb) b2_append_out(c, prev, ch);
    else { OSlot s; s.dst = ch; s.cv = 1; *(OSlot*)(c.base + (size_t)ph * REC_UNIT + REC_OUT_OFF) = s; }
    if (cur_bb) b2_append_in(c, cur, ph);
    else *(u32*)(c.base + (size_t)ch * REC_UNIT + REC_IN_OFF) = ph;
}

#ifndef BUILD_MIN_BLOCKS
#define BUILD_MIN_BLOCKS 3  // 40 registers: three 512-thread blocks per SM also when no shared memory limits them (gacc)
#endif
template <typename DT>
__global__ void __launch_bounds__(BT, BUILD_MIN_BLOCKS) k_build_graph2(Build2Args a) {
    const u32 t = a.first + blockIdx.x;
    if (a.tgt_status[t] != T_OK) return;
    const u32 tid = threadIdx.x;
    const u32 L = a.tgt_tlen[t];
    const u32 ENTER = 0, EXIT = L + 1, nbb = L + 2;
    const u32 a0 = a.tgt_aln_begin[t], a1 = a.tgt_aln_begin[t + 1];
    __shared__ u32 sa[BT / 32], sb[BT / 32], s_cur[TILE];
    __shared__ u32 s_pool_top, s_status, s_enter_out, s_exit_in;
    extern __shared__ __align__(16) u8 s_dyn[];
    // acc_off[t] is a multiple of 16 positions (the planner rounds every target up), so both layouts stay aligned
    u8* accbase = (nbb <= a.smem_positions) ? s_dyn : a.gacc + (u64)acc_bytes_per_pos<DT>() * a.acc_off[t];
    B2Ctx<DT> c;
    c.acc.cov = (u16*)accbase;
    c.acc.w = c.acc.cov + nbb;
    c.acc.c = c.acc.w + nbb;
    if (sizeof(DT) == 2) {
        c.acc.dout = (DT*)(c.acc.c + nbb);
        c.acc.din = c.acc.dout + nbb;
        c.acc.b = (u8*)(c.acc.din + nbb);
    } else {
        c.acc.b = (u8*)(c.acc.c + nbb);
        c.acc.dout = (DT*)(c.acc.b + nbb);
        c.acc.din = c.acc.dout + nbb;
    }
    c.base = a.rec + (size_t)REC_UNIT * a.rec_off[t];
    c.L = L; c.A = a1 - a0; c.nbb = nbb;
    const u32 pool_begin = 16u * (nbb + a.tgt_V[t]);
    c.pool_end = pool_begin + a.pool_cap[t];
    c.enter_out = pool_begin + 1;                          // [cap][A+2 out-slots]
    c.exit_in = pool_begin + 1 + (c.A + 2) + 1;            // [cap][(A+2) in-slots]
    c.s_pool_top = &s_pool_top; c.s_status = &s_status; c.s_enter_out = &s_enter_out; c.s_exit_in = &s_exit_in;

    // ctor: AlnGraphBoost(size_t) :41-62 / AlnGraphBoost(const std::string&) :16-39
    for (u32 p = tid; p < nbb; p += BT) {
        ((int*)c.acc.cov)[p] = 0;   // cov and w (adjacent u16 arrays) start as one int per position: the coverage differences below
        c.acc.c[p] = 0; c.acc.b[p] = 0;
        c.acc.dout[p] = p <= L ? 1 : 0;
        c.acc.din[p] = p >= 1 ? 1 : 0;
        u8* r = c.base + (size_t)REC_UNIT * 2 * p;
        Hdr h;
        h.bfv = 0; h.deg = 0;
        h.out_ptr = p == ENTER ? c.enter_out : gv_inline_out(2 * p);
        h.in_ptr = p == EXIT ? c.exit_in : gv_inline_in(2 * p);
        h.weight = 0; h.cov = 0; h.cnt = 0; h.best = NIL;
        *(Hdr*)r = h;
        if (p >= 1 && p <= L) {
            OSlot s; s.dst = 2 * (p + 1); s.cv = 0;
            *(OSlot*)(r + REC_OUT_OFF) = s;
            *(u32*)(r + REC_IN_OFF) = 2 * (p - 1);
        }
    }
    if (tid == 0) {
        s_status = T_OK;
        s_enter_out = 1; s_exit_in = 1;
        s_pool_top = c.exit_in + (c.A + 2 + 1) / 2;
        if (s_pool_top > c.pool_end) s_status = T_POOL;
        else {
            ((u64*)c.base)[c.enter_out - 1] = c.A + 2;
            ((u64*)c.base)[c.exit_in - 1] = c.A + 2;
            OSlot s; s.dst = 2; s.cv = 0;                              // enter -> 1 (-> exit when L == 0)
            *(OSlot*)(c.base + (size_t)c.enter_out * 8) = s;
            *(u32*)(c.base + (size_t)c.exit_in * 8) = 2 * L;            // L -> exit
        }
    }
    __syncthreads();
    if (s_status != T_OK) { if (tid == 0) a.tgt_status_out[t] = s_status; return; }

    // Coverage of every backbone position, before the replay: an alignment covers [start, start + tkept) -- every one of
    // its match and deletion columns consumes the next position (:75-92) -- so the counts are a prefix sum of +1 / -1
    // marks.  An insertion vertex can then be written ONCE with the coverage of the backbone vertex it maps to
    // (bestPath reads _g[_bbMap[v]].coverage, :407); the pass that patched every insertion record after the replay
    // (a second DRAM round trip per record) is gone.
    {
        int* dcov = (int*)c.acc.cov;
        for (u32 k = a0 + tid; k < a1; k += BT) {
            if (a.a_flags[k] & AF_SKIP) continue;
            const u32 tk = a.a_tkept[k], st = a.a_start[k];
            if (tk) { atomicAdd(&dcov[st], 1); atomicAdd(&dcov[st + tk], -1); }   // st + tk <= L + 1 (k_target_plan)
        }
        __syncthreads();
        const u32 per = (nbb + BT - 1) / BT, lo = min(nbb, tid * per), hi = min(nbb, lo + per);
        u32 sum = 0;
        for (u32 p = lo; p < hi; p++) sum += (u32)dcov[p];
        u32 ex_s, ex_m, tot_s, tot_m;
        block_scan2(sum, 0, ex_s, ex_m, tot_s, tot_m, sa, sb);
        u32 run = ex_s;
        for (u32 p = lo; p < hi; p++) { run += (u32)dcov[p]; dcov[p] = (int)run; }
        __syncthreads();
        // int per position -> the u16 cov array in place, BT positions per round: a round writes bytes that only rounds
        // up to itself have read
        for (u32 r0 = 0; r0 < nbb; r0 += BT) {
            const u32 p = r0 + tid;
            const int v = p < nbb ? dcov[p] : 0;
            __syncthreads();
            if (p < nbb) c.acc.cov[p] = (u16)v;
        }
        for (u32 p = tid; p < nbb; p += BT) c.acc.w[p] = 0;   // every read of the ints is behind the last barrier
        __syncthreads();
    }

    u32 ins_base = nbb;  // id of the next insertion vertex
    for (u32 k = a0; k < a1; k++) {
        if (a.a_flags[k] & AF_SKIP) continue;
        const u32 n = a.a_len[k], start = a.a_start[k];
        const u16* cols = a.norm + a.a_off[k];
        // block-uniform replay state, kept in registers by every thread: no serial section, no barrier to publish it
        u32 carry_t = 0, carry_i = 0, last = ENTER;
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
                u32 tb = start + carry_t + (ex_s & 0xFFFFu), ib = ins_base + carry_i + (ex_s >> 16);
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
            u32 prev = ex_m ? s_cur[ex_m - 1] : last;
            const u32 tile_last = tot_m ? s_cur[tot_m - 1] : last;  // read before the next tile's scan lets s_cur be rewritten
#pragma unroll
            for (int j = 0; j < CPT; j++) {
                const u32 p = bbpos[j];
                if (kind[j] == 1 || kind[j] == 2) {
                    // _g[_bbMap[curr]].coverage++ is in acc.cov already (counted before the replay)
                    c.acc.b[p] = (u8)tc[j];         // .base = targetBase (the last alignment in order wins)
                    if (kind[j] == 1) c.acc.w[p]++;
                } else if (kind[j] == 3) {
                    const u32 hh = cur[j] + nbb;
                    Hdr h;
                    h.bfv = q[j]; h.deg = 0x10001u; h.out_ptr = gv_inline_out(hh); h.in_ptr = gv_inline_in(hh);
                    h.weight = 1; h.cov = p < nbb ? (int)c.acc.cov[p] : 0; h.cnt = 0; h.best = NIL;   // coverage of _bbMap[v]
                    *b2_hdr(c, hh) = h;
                }
                if (kind[j] == 1 || kind[j] == 3) {
                    b2_edge(c, prev, cur[j]);
                    prev = cur[j];
                }
            }
            // The next tile (or alignment) passes the scan's barriers before it touches s_cur or any list again, so
            // this tile's list writes and s_cur reads are ordered against it without a barrier of their own.
            carry_t += tot_s & 0xFFFFu;
            carry_i += tot_s >> 16;
            last = tile_last;
        }
        if (n == 0) __syncthreads();           // an empty alignment ran no scan: order its exit edge by hand
        if (tid == 0) b2_edge(c, last, EXIT);  // addEdge(prevVtx, _exitVtx) :106
        ins_base += carry_i;
        if (n == 0) __syncthreads();
    }
    __syncthreads();
    // fold the running sums into the backbone headers
    const u8* bbstr = a.tgt_bb_off ? a.arena + a.tgt_bb_off[t] : nullptr;
    for (u32 p = tid; p < nbb; p += BT) {
        Hdr* hp = b2_hdr(c, 2 * p);
        Hdr h = *hp;
        const u32 cv = c.acc.cov[p];
        u32 b = p == ENTER ? (u32)'^' : p == EXIT ? (u32)'$' : (bbstr ? (u32)bbstr[p - 1] : (u32)'N');
        if (cv) b = c.acc.b[p];
        const bool inner = p >= 1 && p <= L;
        h.bfv = b | (NF_BACKBONE << 8);
        const u32 dout = p == ENTER ? s_enter_out : c.acc.dout[p], din = p == EXIT ? s_exit_in : c.acc.din[p];
        h.deg = din | (dout << 16);
        h.weight = (inner ? 1 : 0) + (int)(sizeof(DT) == 1 ? (c.acc.w[p] & 0xFFu) : c.acc.w[p]);   // upper byte: b2_edge's enter-edge note
        h.cov = (int)cv;
        *hp = h;
        if (p <= L) ((OSlot*)(c.base + (size_t)h.out_ptr * 8))[0].cv = sizeof(DT) == 1 ? (c.acc.c[p] & 0xFFu) : c.acc.c[p];   // upper byte: b2_edge's skip-edge note
    }
    const u32 nv = ins_base;
    if (tid == 0) {
        a.tgt_nv[t] = nv;
        a.tgt_pool_top[t] = s_pool_top;
        if (s_status != T_OK) a.tgt_status_out[t] = s_status;
    }
}

// ------------------------------------------------------------------------------------------
// K3: mergeNodes (:129-160) -- the order-sensitive part, one WARP per target.  The breadth-first sweep itself is
// warp-parallel: lane j holds out-slot j (or in-slot j) of the popped vertex and loads its neighbour's header, so
// candidate detection (match_any on the base), visited marking and the in-order queue appends are a handful of
// warp instructions regardless of the degree.  The merges themselves mutate several lists in an order-sensitive
// way; they run on lane 0 with the scalar list primitives above.
struct Serial2Args {
    u32 first, count;
    u32 flags, min_weight, min_len;
    const u32* tgt_tlen;
    const u32* tgt_aln_begin;
    u32* tgt_status;
    const u64* rec_off;
    const u32* tgt_V;
    const u32* pool_cap;
    u32* tgt_pool_top;
    const u32* tgt_nv;
    u8* rec;
    const u32* order;       // order[first + i] = target of the launch's i-th warp (longest graphs first), or nullptr
    u32* gscratch;          // the group snapshot stack of mergeInNodes
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
};
__device__ __forceinline__ u32 S2_get(const T2Ctx& c, u32 i) { return c.S[i]; }
__device__ __forceinline__ bool S2_put(T2Ctx& c, u32 i, u32 v) {
    if (i < c.scap) { c.S[i] = v; return true; }
    c.g.status = T_INTERNAL;
    return false;
}

// Snapshot of the merge candidates of n, in list order (the std::map<char, std::vector<VtxDesc>> built at
// :167-181).  Frame: [n, cnt, last key + 129, parent frame, members...].
__device__ __forceinline__ u32 m2_push_in_frame(T2Ctx& c, u32 n, u32& top, u32 parent) {
    const Hdr* h = gv_hdr(c.g, n);
    const u32 nin = h->deg & 0xFFFFu;
    if (nin < 2) return NIL;
    const u32* in = gv_in(c.g, h->in_ptr);
    u32 cnt = 0;
#pragma unroll 1
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
#pragma unroll 1
    for (u32 i = 0; i < cnt; i++) {
        const int b = (int)(signed char)n_base(gv_hdr(c.g, S2_get(c, fb + 4 + i))->bfv);
        if (b > last && b < key) key = b;
    }
    if (key == 1000) return false;
    S2_put(c, fb + 2, (u32)(key + 129));
    an = NIL; members = 0;
#pragma unroll 1
    for (u32 i = 0; i < cnt; i++) {
        const u32 ni = S2_get(c, fb + 4 + i);
        if ((int)(signed char)n_base(gv_hdr(c.g, ni)->bfv) == key) { if (an == NIL) an = ni; members++; }
    }
    return true;
}

// mergeInNodes :162-215 (the recursion :212 runs on the snapshot stack)
__device__ __noinline__ void m2_merge_in(T2Ctx& c, u32 n0) {
    GV& g = c.g;
    u32 top = 0;
    u32 fb = m2_push_in_frame(c, n0, top, NIL);
    while (fb != NIL && g.status == T_OK) {
        u32 an, members;
        if (!m2_next_group(c, fb, an, members)) { top = fb; fb = S2_get(c, fb + 3); continue; }
        if (members < 2) continue;
        const u32 cnt = S2_get(c, fb + 1);
        const int key = (int)S2_get(c, fb + 2) - 129;
        Hdr* han = gv_hdr(g, an);
        if ((han->deg >> 16) == 0) { g.status = T_INTERNAL; return; }
        OSlot* an_out = gv_out(g, han->out_ptr);  // *anoi, taken once (:186-188)
#pragma unroll 1
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S2_get(c, fb + 4 + i);
            const Hdr* hni = gv_hdr(g, ni);
            if (ni == an || (int)(signed char)n_base(hni->bfv) != key) continue;
            if ((hni->deg >> 16) == 0) { g.status = T_INTERNAL; return; }
            an_out->cv += gv_out(g, hni->out_ptr)->cv & ~VISBIT;
            han->weight += hni->weight;
        }
#pragma unroll 1
        for (u32 i = 0; i < cnt; i++) {
            const u32 ni = S2_get(c, fb + 4 + i);
            const Hdr* hni = gv_hdr(g, ni);
            if (ni == an || (int)(signed char)n_base(hni->bfv) != key) continue;
            const u32 nin = hni->deg & 0xFFFFu;
#pragma unroll 1
            for (u32 k = 0; k < nin; k++) {
                // the in-list of ni is stable here: edges are only added towards `an`
                const u32 n1 = gv_in(g, hni->in_ptr)[k];
                Hdr* h1 = gv_hdr(g, n1);
                OSlot* o1 = gv_out(g, h1->out_ptr);
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

// mergeOutNodes :217-267 for one std::map key: `an` is the first member, `others` the rest of the group in list order
__device__ __noinline__ void m2_merge_out_member(GV& g, u32 an, u32 ni) {
    Hdr* han = gv_hdr(g, an);
    Hdr* hni = gv_hdr(g, ni);
    // _g[*anii].count += _g[*ii].count (:241): the first in-edge of each; an edge's count lives in its source's slot
    if ((han->deg & 0xFFFFu) == 0 || (hni->deg & 0xFFFFu) == 0) { g.status = T_INTERNAL; return; }
    OSlot* ea = gv_edge_slot(g, gv_in(g, han->in_ptr)[0], an);
    OSlot* ei = gv_edge_slot(g, gv_in(g, hni->in_ptr)[0], ni);
    if (!ea || !ei) { g.status = T_INTERNAL; return; }
    ea->cv += ei->cv & ~VISBIT;
    han->weight += hni->weight;
    const u32 nout = hni->deg >> 16;
#pragma unroll 1
    for (u32 k = 0; k < nout; k++) {
        const OSlot s = gv_out(g, hni->out_ptr)[k];
        OSlot* oa = gv_out(g, han->out_ptr);
        const u32 e = gv_find_out(oa, han->deg >> 16, s.dst);
        if (e != NIL) oa[e].cv += s.cv & ~VISBIT;
        else gv_add_edge(g, an, s.dst, s.cv & ~VISBIT, s.cv & VISBIT);
    }
    gv_reap(g, ni);
}

// generic mergeOutNodes (:217-267) for vertices with more than 32 out-edges (enter): snapshot on the stack
__device__ __noinline__ void m2_merge_out_generic(T2Ctx& c, u32 n) {
    GV& g = c.g;
    const Hdr* hn = gv_hdr(g, n);
    const u32 nout = hn->deg >> 16;
    if (nout < 2) return;
    u32 cnt = 0;
#pragma unroll 1
    for (u32 i = 0; i < nout; i++) {
        const u32 d = gv_out(g, hn->out_ptr)[i].dst;
        if ((gv_hdr(g, d)->deg & 0xFFFFu) == 1) { if (!S2_put(c, 4 + cnt, d)) return; cnt++; }
    }
    if (cnt < 2) return;
    S2_put(c, 0, n); S2_put(c, 1, cnt); S2_put(c, 2, 0); S2_put(c, 3, NIL);
    u32 an, members;
    while (g.status == T_OK && m2_next_group(c, 0, an, members)) {
        if (members < 2) continue;
        const int key = (int)S2_get(c, 2) - 129;
#pragma unroll 1
        for (u32 i = 0; i < cnt && g.status == T_OK; i++) {
            const u32 ni = S2_get(c, 4 + i);
            if (ni == an || (int)(signed char)n_base(gv_hdr(g, ni)->bfv) != key) continue;
            m2_merge_out_member(g, an, ni);
        }
    }
}

#define FULLMASK 0xFFFFFFFFu
// Appends the vertices `v` of the lanes in `ready` (in lane order) to the intrusive FIFO (Hdr.best links).
__device__ __forceinline__ void m2_queue_append(const GV& g, u32 ready, bool mine, u32 v, u32& tail, u32 lane) {
    if (!ready) return;
    const u32 before = ready & ((1u << lane) - 1u);
    const u32 src = before ? 31u - (u32)__clz((int)before) : 0u;
    const u32 pv = __shfl_sync(FULLMASK, v, src);
    if (mine) gv_hdr(g, before ? pv : tail)->best = v;
    tail = __shfl_sync(FULLMASK, v, 31u - (u32)__clz((int)ready));
}

// mergeInNodes(u) (:162-215) by the whole warp, for the shape almost every merge has: every member of a group is a plain
// chain vertex (one in-edge, one out-edge -- the edge to u), the members' predecessors n1 are distinct and none of them
// has an edge to the survivor yet.  Lane j holds in-slot j of u (`s`, header words `d`, taken before any merge: the
// std::map snapshot of :167-181); `dup` = the lane's vertex is in a group of two or more.  Groups go in ascending
// (signed) base order, members in list order.  Per member ni != an, all members at once:
//   :186-191  the count of ni's out-edge and ni's weight go to an
//   :196-208  n1 -> an does not exist: add_edge(n1, an) with the count / visited flag of n1 -> ni, i.e. n1's out-list gets
//             (an) at its end and an's in-list gets n1, in member order
//   :209      markForReaper(ni): ni leaves n1's out-list and u's in-list, order of the rest kept
//   :212      mergeInNodes(an): an's predecessors are the members' n1; only if two of them qualify does the scalar code run
// Anything else (a member with more edges, shared or already-connected predecessors, lists that would outgrow their
// place) goes to the scalar m2_merge_in, which takes its own snapshot of what is left.  Returns the target's status.
__device__ __noinline__ u32 m2_merge_in_warp(T2Ctx& c, const u32 u, const u32 nin, const u32 s, const uint2 d, const bool dup) {
    GV& g = c.g;
    const u32 lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
    const bool have = lane < nin;
    const u32 skey = (u32)((int)(signed char)(d.x & 0xFFu) + 128);
    u32 removed = 0, vis_removed = 0;
    bool pending = dup, fallback = false;
    for (;;) {
        const u32 k = __reduce_min_sync(FULLMASK, pending ? skey : 0xFFFFFFFFu);
        if (k == 0xFFFFFFFFu) break;
        const bool in_grp = pending && skey == k;
        const u32 M = __ballot_sync(FULLMASK, in_grp);
        const u32 la = (u32)__ffs((int)M) - 1;              // the survivor `an` = first member in list order
        const u32 an = __shfl_sync(FULLMASK, s, la);
        u32 n1 = NIL, cvo = 0, wgt = 0, e_ni = NIL, d1 = 0, cvn = 0, bfv_m = 0;
        uint2 dn1 = make_uint2(0u, 0u);
        OSlot* o1 = nullptr;
        bool ok = true;
        if (in_grp) {
            const bool is_a = lane == la;
            const Hdr m = *gv_hdr(g, s);
            bfv_m = m.bfv;
            ok = m.deg == 0x10001u;
            if (ok) {
                wgt = (u32)m.weight;
                const OSlot so = *gv_out(g, m.out_ptr);
                cvo = so.cv;
                ok = so.dst == u;
                n1 = *gv_in(g, m.in_ptr);
                if (is_a) ok = ok && m.in_ptr == gv_inline_in(s) && (u32)__popc(M) <= REC_IN_CAP;   // an's in-list grows in place
                const Hdr* h1 = gv_hdr(g, n1);
                dn1 = *(const uint2*)h1;
                d1 = dn1.y >> 16;
                o1 = gv_out(g, h1->out_ptr);
                ok = ok && d1 <= 32;
                if (ok) {
#pragma unroll 1
                    for (u32 i = 0; i < d1; i++) {
                        const u32 dst = o1[i].dst;
                        if (dst == s && e_ni == NIL) e_ni = i;
                        if (dst == an && !is_a) ok = false;      // edge(n1, an) exists (:201): scalar
                    }
                    if (e_ni == NIL) ok = false; else cvn = o1[e_ni].cv;
                }
            }
            if (__popc(__match_any_sync(M, n1)) != 1) ok = false;   // two members behind one predecessor: scalar
        }
        if (__any_sync(FULLMASK, in_grp && !ok)) { fallback = true; break; }
        if (in_grp) {
            const bool is_a = lane == la;
            const u32 cnt_sum = __reduce_add_sync(M, is_a ? 0u : (cvo & ~VISBIT));
            const u32 w_sum = __reduce_add_sync(M, is_a ? 0u : wgt);
            const u32 nvis = (u32)__popc(__ballot_sync(M, !is_a && (cvn & VISBIT)));
            const u32 gone_vis = __ballot_sync(M, !is_a && (cvo & VISBIT));
            const u32 others = M & ~(1u << la);
            Hdr* han = gv_hdr(g, an);
            if (is_a) {
                gv_out(g, han->out_ptr)->cv = cvo + cnt_sum;
                han->weight = (int)(wgt + w_sum);
                han->deg = (u32)__popc(M) | 0x10000u;
                han->bfv = bfv_m + (nvis << 16);
            } else {
#pragma unroll 1
                for (u32 i = e_ni; i + 1 < d1; i++) o1[i] = o1[i + 1];
                OSlot ns;
                ns.dst = an; ns.cv = cvn;
                o1[d1 - 1] = ns;
                gv_in(g, gv_inline_in(an))[1 + __popc(others & lt)] = n1;
                *(uint2*)gv_hdr(g, s) = make_uint2((bfv_m & 0xFFFFu) | (NF_DELETED << 8), 0u);
            }
            removed |= others;
            vis_removed |= gone_vis;
        }
        removed = __shfl_sync(FULLMASK, removed, la);
        vis_removed = __shfl_sync(FULLMASK, vis_removed, la);
        __syncwarp();
        // mergeInNodes(an) :212 -- an's in-list is the members' n1 in lane order; their out-degrees did not change
        bool rec = false;
        if (in_grp) {
            const bool c1 = (dn1.y >> 16) == 1;
            const u32 g1 = __match_any_sync(M, c1 ? (dn1.x & 0xFFu) : 0x100u + lane);
            rec = c1 && __popc(g1) > 1;
        }
        if (__any_sync(FULLMASK, rec)) {
            if (lane == 0) m2_merge_in(c, an);
            __syncwarp();
            if (__shfl_sync(FULLMASK, g.status, 0) != T_OK) break;
        }
        pending = pending && !in_grp;
    }
    // u's in-list without the merged members, order kept; their (visited) edges no longer count at u
    if (removed) {
        Hdr* pu = gv_hdr(g, u);
        const u32 surv = __ballot_sync(FULLMASK, have) & ~removed;
        const u32 in_ptr = pu->in_ptr, deg = pu->deg, bfv = pu->bfv;
        __syncwarp();
        if (have && !((removed >> lane) & 1u)) gv_in(g, in_ptr)[__popc(surv & lt)] = s;
        if (lane == 0) { pu->deg = (deg & 0xFFFF0000u) | (u32)__popc(surv); pu->bfv = bfv - ((u32)__popc(vis_removed) << 16); }
    }
    __syncwarp();
    if (fallback) {
        if (lane == 0) m2_merge_in(c, u);
        __syncwarp();
    }
    return __shfl_sync(FULLMASK, g.status, 0);
}

__device__ __forceinline__ bool warp2_setup(const Serial2Args& a, u32& t, T2Ctx& c) {
    const u32 idx = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (idx >= a.count) return false;
    t = a.order ? a.order[a.first + idx] : a.first + idx;
    if (a.tgt_status[t] != T_OK) return false;
    c.g.base = a.rec + (size_t)REC_UNIT * a.rec_off[t];
    c.g.nbb = a.tgt_tlen[t] + 2;
    const u32 pool_begin = 16u * (c.g.nbb + a.tgt_V[t]);
    c.g.pool_top = a.tgt_pool_top[t];
    c.g.pool_end = pool_begin + a.pool_cap[t];
    c.g.status = T_OK;
    c.g.pool_failed = 0;
    c.g.nunits = (a.flags & F_NO_PREFETCH) ? 0u : pool_begin / 16u;
    c.S = a.gscratch + a.gscratch_off[t];
    c.scap = a.gscratch_cap[t];
    return true;
}

#ifndef WARP_KERNEL_THREADS
#define WARP_KERNEL_THREADS 128
#endif
#ifndef MERGE_MIN_BLOCKS
#define MERGE_MIN_BLOCKS 9   // resident 128-thread blocks per SM the sweeps' register allocation must allow: 56 registers,
                             // 36 warps per SM.  Measured on B200, 10 000 targets, mergeNodes alone: 8 blocks 168 ms, 9-10 blocks
                             // 133-135 ms, 12 blocks (40 registers, spills) 143 ms, 16 blocks 161 ms; both sweeps fused: 9 blocks
                             // 189 ms, 10 blocks 193 ms, 11 blocks 204 ms
#endif

// mergeNodes :129-160
__device__ __forceinline__ u32 merge_sweep(const Serial2Args& a, const u32 t, T2Ctx& c) {
    GV& g = c.g;
    const u32 lane = threadIdx.x & 31;
    const u32 ENTER = 0;
    u32 head = ENTER, tail = ENTER, status = T_OK;
    if (lane == 0) { gv_hdr(g, ENTER)->best = NIL; gv_hdr(g, ENTER)->bfv |= NF_QUEUED << 8; }
    __syncwarp();
    while (head != NIL && status == T_OK) {
        const u32 u = head;
        Hdr* pu = gv_hdr(g, u);
        Hdr hu = *pu;
        // ---- mergeInNodes(u): is there a group of >= 2 predecessors with out_degree 1 and the same base?
        u32 nin = hu.deg & 0xFFFFu;
        if (nin >= 2) {
            if (nin > 32) {  // only exit gets that wide: the scalar code decides and merges
                if (lane == 0) m2_merge_in(c, u);
                __syncwarp();
                status = __shfl_sync(FULLMASK, g.status, 0);
                hu = *pu;
            } else {
                const bool have = lane < nin;
                const u32 s = have ? gv_in(g, hu.in_ptr)[lane] : 0u;
                uint2 d = make_uint2(0u, 0u);
                if (have) d = *(const uint2*)gv_hdr(g, s);
                const bool cand = have && (d.y >> 16) == 1;
                const u32 grp = __match_any_sync(FULLMASK, cand ? (d.x & 0xFFu) : 0x100u + lane);
                const bool dup = cand && __popc(grp) > 1;
                if (__any_sync(FULLMASK, dup)) {
                    status = m2_merge_in_warp(c, u, nin, s, d, dup);
                    hu = *pu;
                }
            }
        }
        // ---- mergeOutNodes(u), then mark the out-edges visited and enqueue the successors that became ready
        u32 nout = hu.deg >> 16;
        if (nout == 1) {
            // One successor: nothing for mergeOutNodes to group (:231 needs two).  Most vertices are such chain links
            // (an insertion run, a backbone stretch every read agrees on); they take this warp-uniform short cut
            // instead of the list machinery below: visit the edge, enqueue the successor if it became ready.
            OSlot* o = gv_out(g, hu.out_ptr);
            const OSlot s0 = *o;
            Hdr* hv = gv_hdr(g, s0.dst);
            const uint2 d0 = *(const uint2*)hv;
            const bool fresh = !(s0.cv & VISBIT);
            u32 bfv = d0.x + (fresh ? 0x10000u : 0u);
            const bool ready = (bfv >> 16) == (d0.y & 0xFFFFu);  // notVisited == 0 (:150-157)
            if (ready && (bfv & (NF_QUEUED << 8))) status = T_INTERNAL;
            if (ready) bfv |= NF_QUEUED << 8;
            __syncwarp();
            if (lane == 0) {
                if (fresh) o->cv = s0.cv | VISBIT;
                if (bfv != d0.x) hv->bfv = bfv;
                if (ready) gv_hdr(g, tail)->best = s0.dst;
            }
            if (ready) tail = s0.dst;
            __syncwarp();
            // u's queue link was in its header when it was popped (NIL = u was the tail: what follows is what this step appended)
            head = hu.best != NIL ? hu.best : (ready ? s0.dst : NIL);
            continue;
        }
        if (nout > 32) {  // enter (one edge per alignment start): scalar
            if (lane == 0) {
                m2_merge_out_generic(c, u);
                hu = *pu;
                nout = hu.deg >> 16;
#pragma unroll 1
                for (u32 i = 0; i < nout && g.status == T_OK; i++) {
                    OSlot* o = gv_out(g, hu.out_ptr) + i;
                    Hdr* hv = gv_hdr(g, o->dst);
                    if (!(o->cv & VISBIT)) { o->cv |= VISBIT; hv->bfv += 0x10000u; }
                    const u32 bfv = hv->bfv;
                    if (n_vis(bfv) == (hv->deg & 0xFFFFu)) {  // notVisited == 0
                        if (n_flags(bfv) & NF_QUEUED) { g.status = T_INTERNAL; break; }
                        hv->bfv = bfv | (NF_QUEUED << 8);
                        gv_hdr(g, tail)->best = o->dst;
                        tail = o->dst;
                    }
                }
            }
            __syncwarp();
            tail = __shfl_sync(FULLMASK, tail, 0);
            status = __shfl_sync(FULLMASK, g.status, 0);
            head = pu->best;
            continue;
        }
        bool have = lane < nout;
        OSlot sl;
        sl.dst = 0; sl.cv = 0;
        uint2 d = make_uint2(0u, 0u);
        if (have) {
            sl = gv_out(g, hu.out_ptr)[lane];
            const u8* rv = g.base + (size_t)sl.dst * REC_UNIT;
            d = *(const uint2*)rv;
#ifndef GV_NO_PREFETCH
            const uint4 o01 = *(const uint4*)(rv + REC_OUT_OFF);  // the successor's first two out-slots (inline position)
            if (g.nunits) pf_l2(rv + 32);
            if (d.y >> 16) gv_prefetch_rec(g, o01.x);
            if ((d.y >> 16) > 1) gv_prefetch_rec(g, o01.z);
#endif
        }
        {
            const bool cand = have && (d.y & 0xFFFFu) == 1;
            const u32 key = d.x & 0xFFu;
            const u32 grp = __match_any_sync(FULLMASK, cand ? key : 0x100u + lane);
            bool dup = cand && __popc(grp) > 1;
            if (__any_sync(FULLMASK, dup)) {
                const u32 skey = (u32)((int)(signed char)key + 128);  // std::map<char,...> iterates signed chars ascending
                // ---- can the whole merge run warp-parallel?  Yes when every member is a plain chain vertex (one
                // in-edge, one unvisited out-edge, inline lists) and a
                // group has at most 6 members (the survivor's inline out-slots then hold every distinct target).
                u32 x = NIL, cvx = 0, wgt = 0;
                bool ok = true;
                if (dup) {
                    u8* r = g.base + (size_t)sl.dst * REC_UNIT;
                    const u32 optr = ((const Hdr*)r)->out_ptr;
                    wgt = (u32)((const Hdr*)r)->weight;
                    const OSlot s0 = *(const OSlot*)(r + REC_OUT_OFF);
                    x = s0.dst; cvx = s0.cv;
                    ok = d.y == 0x10001u && optr == gv_inline_out(sl.dst) && !(cvx & VISBIT) && __popc(grp) <= 6;
                }
                // (no member can be the target of another member: its only in-edge, d.y == 0x10001, is the one from u)
                if (!__any_sync(FULLMASK, dup && !ok)) {
                    u32 removed = 0;
                    bool pending = dup;
                    for (;;) {
                        const u32 k = __reduce_min_sync(FULLMASK, pending ? skey : 0xFFFFFFFFu);
                        if (k == 0xFFFFFFFFu) break;
                        const bool in_grp = pending && skey == k;
                        const u32 M = __ballot_sync(FULLMASK, in_grp);
                        const u32 la = (u32)__ffs((int)M) - 1;          // the survivor `an` = first member in list order
                        const u32 an = __shfl_sync(FULLMASK, sl.dst, la);
                        u32 Mx = 0, xdeg = 0, xinp = 0;
                        bool leader = false;
                        if (in_grp) {
                            const bool is_a = lane == la;
                            // :238-243  _g[*anii].count += _g[*ii].count;  _g[an].weight += _g[*ni].weight
                            const u32 cvs = __reduce_add_sync(M, is_a ? 0u : (sl.cv & ~VISBIT));
                            const u32 ws = __reduce_add_sync(M, is_a ? 0u : wgt);
                            // :246-259  the members' out-edges move to `an`: one slot per distinct target, counts summed,
                            // new targets in the order of the member that brings them
                            Mx = __match_any_sync(M, x);
                            leader = lane == (u32)__ffs((int)Mx) - 1;
                            const u32 Cx = __reduce_add_sync(Mx, cvx);
                            const u32 newl = __ballot_sync(M, leader && !is_a);
                            u8* ra = g.base + (size_t)an * REC_UNIT;
                            if (is_a) {
                                sl.cv += cvs;
                                ((Hdr*)ra)->weight = (int)(wgt + ws);
                                ((Hdr*)ra)->deg = 1u | ((1u + (u32)__popc(newl)) << 16);
                                ((OSlot*)(ra + REC_OUT_OFF))[0].cv = Cx;
                            } else if (leader) {
                                OSlot ns;
                                ns.dst = x; ns.cv = Cx;
                                ((OSlot*)(ra + REC_OUT_OFF))[1 + __popc(newl & ((1u << lane) - 1u))] = ns;
                            }
                            if (leader) {   // every leader's header at once
                                const Hdr* hx = gv_hdr(g, x);
                                xdeg = hx->deg; xinp = hx->in_ptr;
                                const u32 nmx = (u32)__popc(Mx);
                                if (is_a && nmx == 1) leader = false;       // only `an` itself points there: the in-list stays as it is
                                else if ((xdeg & 0xFFFFu) == nmx) {          // every in-edge of x comes from this group (no parallel
                                    *gv_in(g, xinp) = an;                    // edges): what is left of the list is [an]
                                    gv_hdr(g, x)->deg = (xdeg & 0xFFFF0000u) | 1u;
                                    leader = false;
                                }
                            }
                            // markForReaper(member) :260
                            if (!is_a) *(uint2*)gv_hdr(g, sl.dst) = make_uint2((d.x & 0xFFFFu) | (NF_DELETED << 8), 0u);
                        }
                        // The in-list of each distinct target: the merged members leave (clear_vertex keeps the order of
                        // the rest), `an` is appended unless it was a source already.  One target at a time, by the WHOLE
                        // warp: lane i takes entry i, a ballot + prefix count compacts the list in place (the scalar loop of
                        // one leader lane that stood here was 10.7 % of the launch's instructions at 1.2 active lanes).
                        for (u32 lm = __ballot_sync(FULLMASK, leader); lm; lm &= lm - 1) {
                            const u32 j = (u32)__ffs((int)lm) - 1;
                            const u32 xj = __shfl_sync(FULLMASK, x, j), hdeg = __shfl_sync(FULLMASK, xdeg, j);
                            const u32 leaving = __shfl_sync(FULLMASK, Mx, j) & ~(1u << la);
                            u32* inx = gv_in(g, __shfl_sync(FULLMASK, xinp, j));
                            const u32 xin = hdeg & 0xFFFFu;
                            u32 w = 0;
#pragma unroll 1
                            for (u32 b0 = 0; b0 < xin; b0 += 32) {
                                const bool hv = b0 + lane < xin;
                                const u32 e = hv ? inx[b0 + lane] : NIL;
                                bool gone = false;
                                for (u32 mm = leaving; mm; mm &= mm - 1) gone |= e == __shfl_sync(FULLMASK, sl.dst, __ffs((int)mm) - 1);
                                const bool keep = hv && !gone;
                                const u32 bal = __ballot_sync(FULLMASK, keep);
                                if (keep) inx[w + __popc(bal & ((1u << lane) - 1u))] = e;   // never beyond the entry's old place
                                w += __popc(bal);
                            }
                            if (w + (u32)__popc(leaving) != xin) status = T_INTERNAL;
                            if (j != la) { if (lane == 0) inx[w] = an; w++; }
                            if (lane == 0) gv_hdr(g, xj)->deg = (hdeg & 0xFFFF0000u) | w;
                            __syncwarp();
                        }
                        removed |= M & ~(1u << la);
                        pending = pending && !in_grp;
                    }
                    // u's out-list without the merged members, order kept
                    const u32 surv = __ballot_sync(FULLMASK, have) & ~removed;
                    if (have && !((removed >> lane) & 1u)) gv_out(g, hu.out_ptr)[__popc(surv & ((1u << lane) - 1u))] = sl;
                    if (lane == 0) pu->deg = (hu.deg & 0xFFFFu) | ((u32)__popc(surv) << 16);
                    if (__any_sync(FULLMASK, status != T_OK)) status = T_INTERNAL;
                } else {
                    // groups in ascending key order, members in list order, on lane 0
                    for (;;) {
                        const u32 k = __reduce_min_sync(FULLMASK, dup ? skey : 0xFFFFFFFFu);
                        if (k == 0xFFFFFFFFu) break;
                        const bool in_grp = dup && skey == k;
                        u32 members = __ballot_sync(FULLMASK, in_grp);
                        const u32 an = __shfl_sync(FULLMASK, sl.dst, __ffs((int)members) - 1);
                        members &= members - 1;
                        while (members) {
                            const u32 ni = __shfl_sync(FULLMASK, sl.dst, __ffs((int)members) - 1);
                            members &= members - 1;
                            if (lane == 0 && g.status == T_OK) m2_merge_out_member(g, an, ni);
                        }
                        if (in_grp) dup = false;
                    }
                    __syncwarp();
                    status = __shfl_sync(FULLMASK, g.status, 0);
                }
                __syncwarp();
                hu = *pu;  // merging only removed out-edges of u (and changed their targets)
                nout = hu.deg >> 16;
                have = lane < nout;
                sl.dst = 0; sl.cv = 0;
                d = make_uint2(0u, 0u);
                if (have) { sl = gv_out(g, hu.out_ptr)[lane]; d = *(const uint2*)gv_hdr(g, sl.dst); }
            }
        }
        {
            const bool fresh = have && !(sl.cv & VISBIT);
            if (fresh) gv_out(g, hu.out_ptr)[lane].cv = sl.cv | VISBIT;
            u32 bfv = d.x + (fresh ? 0x10000u : 0u);
            const bool ready = have && (bfv >> 16) == (d.y & 0xFFFFu);  // notVisited == 0
            if (ready && (bfv & (NF_QUEUED << 8))) status = T_INTERNAL;
            if (ready) bfv |= NF_QUEUED << 8;
            if (have && bfv != d.x) gv_hdr(g, sl.dst)->bfv = bfv;
            const u32 rdy = __ballot_sync(FULLMASK, ready);
            m2_queue_append(g, rdy, ready, sl.dst, tail, lane);
            if (__any_sync(FULLMASK, status != T_OK)) status = T_INTERNAL;
            const u32 first = __shfl_sync(FULLMASK, sl.dst, rdy ? (u32)__ffs((int)rdy) - 1 : 0u);
            head = hu.best != NIL ? hu.best : (rdy ? first : NIL);
        }
        __syncwarp();
    }
    if (lane == 0) {
        a.tgt_pool_top[t] = g.pool_top;
        if (status == T_OK) status = g.status;
        if (g.pool_failed) status = T_POOL;
        if (status != T_OK) a.tgt_status[t] = status;
    }
    __syncwarp();
    return __shfl_sync(FULLMASK, status, 0);
}
__global__ void __launch_bounds__(WARP_KERNEL_THREADS, MERGE_MIN_BLOCKS) k_merge_nodes2(Serial2Args a) {
    u32 t;
    T2Ctx c;
    if (!warp2_setup(a, t, c)) return;
    merge_sweep(a, t, c);
}

// ------------------------------------------------------------------------------------------
// K4: bestPath (:375-459) and the two consensus() forms (:285-373), one warp per target.
// Reverse topological sweep from exit: lane j scores out-edge j of the popped vertex (its target's header is one
// 32-byte load), a warp max + ballot picks the FIRST maximum in list order (the reference's strict `>` scan),
// then lane j counts in-edge j down at its source and the sources that became ready are appended in list order.
__device__ __forceinline__ u32 f2ord(float f) {  // order-preserving map float -> u32
    const u32 u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void best_sweep(const Serial2Args& a, const u32 t, const T2Ctx& c) {
    const GV& g = c.g;
    const u32 lane = threadIdx.x & 31;
    const u32 L = a.tgt_tlen[t], ENTER = 0, EXIT = 2 * (L + 1), nV = a.tgt_nv[t];
    u32 head = EXIT, tail = EXIT;
    if (lane == 0) gv_hdr(g, EXIT)->best = NIL;
    __syncwarp();
    while (head != NIL) {
        const u32 n = head;
        Hdr* pn = gv_hdr(g, n);
        const Hdr hn = *pn;
        const u32 nout = hn.deg >> 16, nin = hn.deg & 0xFFFFu;
        // ---- the best out-edge: first maximum of the per-edge scores in list order
        float best_score = -FLT_MAX;
        u32 best_dst = NIL;
        if (nout == 1) {  // a chain link: its one edge is the best edge; every lane computes it, no reduction
            const OSlot s = *gv_out(g, hn.out_ptr);
            const Hdr hd = *gv_hdr(g, s.dst);
            best_dst = s.dst;
            if ((hd.bfv & (NF_BACKBONE << 8)) && hd.weight == 1) best_score = __fsub_rn(hd.score, 10.0f);
            else best_score = __fadd_rn(__fsub_rn((float)(int)(s.cv & ~VISBIT), __fmul_rn((float)hd.cov, 0.5f)), hd.score);
            if (!(best_score > -FLT_MAX)) best_dst = NIL;  // the reference's strict `>` against its initial value
        } else
        for (u32 b0 = 0; b0 < nout; b0 += 32) {
            const bool have = b0 + lane < nout;
            float ns = -FLT_MAX;
            u32 dst = NIL;
            if (have) {
                const OSlot s = gv_out(g, hn.out_ptr)[b0 + lane];
                const Hdr hd = *gv_hdr(g, s.dst);
                dst = s.dst;
                // bestPath :402-409: a backbone vertex no read agrees with costs 10; otherwise count - coverage/2
                if ((hd.bfv & (NF_BACKBONE << 8)) && hd.weight == 1) ns = __fsub_rn(hd.score, 10.0f);
                else ns = __fadd_rn(__fsub_rn((float)(int)(s.cv & ~VISBIT), __fmul_rn((float)hd.cov, 0.5f)), hd.score);
            }
            const u32 key = have ? f2ord(ns) : 0u;
            const u32 mx = __reduce_max_sync(FULLMASK, key);
            const u32 who = __ffs((int)__ballot_sync(FULLMASK, have && key == mx)) - 1;
            const float cs = __shfl_sync(FULLMASK, ns, who);
            const u32 cd = __shfl_sync(FULLMASK, dst, who);
            if (cs > best_score) { best_score = cs; best_dst = cd; }  // strict >: an earlier chunk's maximum stays
        }
        // ---- count this vertex off at its predecessors; those with all successors scored join the queue
        u32 first_app = NIL;  // first vertex this step appends: the next pop when n was the queue's tail
        if (nin == 1) {  // a chain link again: one predecessor, handled warp-uniformly without the append machinery
            const u32 s = gv_in(g, hn.in_ptr)[0];
            Hdr* hs = gv_hdr(g, s);
            const u32 sdeg = hs->deg;
            const u32 k = (u32)(hs->cnt + 1);
            const bool ready = k == (sdeg >> 16);
            __syncwarp();
            if (lane == 0) {
                hs->cnt = (int)k;
                if (ready) { hs->best = NIL; gv_hdr(g, tail)->best = s; }
            }
            if (ready) { tail = s; first_app = s; }
        } else
        for (u32 b0 = 0; b0 < nin; b0 += 32) {
            const bool have = b0 + lane < nin;
            u32 s = NIL;
            bool ready = false;
            if (have) {
                s = gv_in(g, hn.in_ptr)[b0 + lane];
                Hdr* hs = gv_hdr(g, s);
                const u32 sdeg = hs->deg;
                const u32 k = (u32)(hs->cnt + 1);
#ifndef GV_NO_PREFETCH
                const uint2 i01 = *(const uint2*)((const u8*)hs + REC_IN_OFF);  // the predecessor's first two in-slots (inline position)
                if (g.nunits) pf_l2((const u8*)hs + 64);                        // its out-slots, for when it is popped
#endif
                hs->cnt = (int)k;
                ready = k == (sdeg >> 16);
                if (ready) hs->best = NIL;
#ifndef GV_NO_PREFETCH
                if (sdeg & 0xFFFFu) gv_prefetch_rec(g, i01.x);
                if ((sdeg & 0xFFFFu) > 1) gv_prefetch_rec(g, i01.y);
#endif
            }
            __syncwarp();
            const u32 rdy = __ballot_sync(FULLMASK, ready);
            m2_queue_append(g, rdy, ready, s, tail, lane);
            if (rdy && first_app == NIL) first_app = __shfl_sync(FULLMASK, s, (u32)__ffs((int)rdy) - 1);
        }
        const u32 nx = hn.best != NIL ? hn.best : first_app;  // the queue link was in n's header when it was popped
        __syncwarp();
        if (lane == 0) {
            pn->score = best_dst != NIL ? best_score : 0.0f;  // an absent map key reads as 0.0f (:401)
            pn->best = best_dst;                              // from here on: the successor on the best path
            pn->bfv = hn.bfv | (NF_SCORED << 8);
        }
        __syncwarp();
        head = nx;
    }
    if (lane != 0) return;
    // walk the best path and cut it into min-weight runs (consensus :285-373)
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
        // the path is a chain of dependent loads, but it mostly runs along the backbone, whose records sit at known places:
        // ask for the backbone vertex a few positions ahead (records this far from exit left L2 long ago)
        if (prev < 2 * (L + 1) && g.nunits) pf_l2(g.base + (size_t)min(prev + 16u, 2 * (L + 1)) * REC_UNIT);
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
        prev = hp.best;
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
__global__ void __launch_bounds__(WARP_KERNEL_THREADS) k_best_path2(Serial2Args a) {
    u32 t;
    T2Ctx c;
    if (!warp2_setup(a, t, c)) return;
    best_sweep(a, t, c);
}

// K3 + K4 in one launch: a warp that has merged its target goes straight on to the best path.  The reverse sweep
// starts at exit, where the forward sweep has just ended (those records are still in L2), and the warps that finish
// mergeNodes early no longer wait at a kernel boundary for the longest chain of the launch.
__global__ void __launch_bounds__(WARP_KERNEL_THREADS, MERGE_MIN_BLOCKS) k_merge_best2(Serial2Args a) {
    u32 t;
    T2Ctx c;
    if (!warp2_setup(a, t, c)) return;
    if (merge_sweep(a, t, c) != T_OK) return;
    best_sweep(a, t, c);
}
