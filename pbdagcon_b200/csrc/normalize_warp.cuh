// normalize_warp.cuh -- K1 fast path: normalizeGaps + trimAln with one WARP per alignment.
//
// Reference semantics: src/cpp/Alignment.cpp:131-217 (normalizeGaps) and :219-242 (trimAln).
//
// normalizeGaps is a left-to-right transducer: after '.'->'-' and the mismatch split (:147-159), column i may
// pull the next not-yet-consumed target base into its target gap if it equals the query base, then the next
// not-yet-consumed query base into its query gap if it equals the (new) target base (:167-197); gap/gap columns
// are dropped (:199-206).  Because a pull always takes the NEXT pending base, the whole state between two
// columns is the pair of "leads" (bases consumed ahead of the cursor) of the two strands, and the leads fall
// back to (0,0) within a few columns.  So the 32 lanes each run the transducer over their own 32 raw columns
// assuming leads (0,0) at the chunk start; a lane whose left neighbour ended with other leads re-runs with the
// true ones until the chain is consistent (usually 0-1 repair rounds, at most 32).  The pending bases are looked
// up in O(1) in per-tile compacted base streams (ts/qs) that reach NW_MG columns past the tile; an alignment
// whose look-ahead would leave that window is handed to the serial kernel (AF_RETRY).
//
// Per tile the warp stages 1280 raw columns of q and t with coalesced 4-byte loads (funnel-shifted to undo the
// strings' byte phase) into shared memory; output columns (q | t<<8, u16) go straight to the alignment's slot.
#pragma once

#define NW_TR 1024u  // raw columns normalised per tile: 32 per lane
#define NW_MG 256u   // look-ahead margin staged behind the tile: 8 per lane
#define NW_W (NW_TR + NW_MG)
#ifndef NW_WARPS
#define NW_WARPS 4
#endif
#ifndef NW_MIN_BLOCKS
#define NW_MIN_BLOCKS 7
#endif
#define NW_STG 66u   // u16 per lane in the output staging area: 64 columns + padding against bank conflicts
#define AF_RETRY 2u  // needs the serial kernel (look-ahead left the staged window)

struct NwSmem {
    union {  // the staged raw strings are in registers by the time the transducer writes its output rows
        struct {
            uint4 q[NW_W / 16];
            uint4 t[NW_W / 16];
        };
        u16 stage[32 * NW_STG];  // every lane's output columns of the tile (a raw column yields at most two)
    };
    u8 ts[NW_W];  // target bases of the staged window, in order
    u8 qs[NW_W];  // query bases
};

__device__ __forceinline__ u32 nw_byte(const u32* w, int j) { return (w[j >> 2] >> ((j & 3) * 8)) & 0xFFu; }
// '.' -> '-' in all four bytes of a staged word (Alignment.cpp:133-146 does it to both strings before anything else): a byte
// of y is zero where the character is '.', z has 0x80 there, and '.' ^ '-' == 3
__device__ __forceinline__ u32 nw_undot4(u32 w) {
    const u32 y = w ^ 0x2E2E2E2Eu;
    const u32 z = ~(((y & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | y) & 0x80808080u;
    return w ^ ((z >> 7) | (z >> 6));
}

// how many of the word's bytes [first, first + 4) below `nvalid` are bases, i.e. not '-' (four at a time: a byte of x is
// zero where the character is '-'; bit 7 of ((v & 0x7F) + 0x7F) | v says "non-zero")
__device__ __forceinline__ u32 nw_count_bases(u32 w, u32 nvalid, u32 first) {
    const u32 x = w ^ 0x2D2D2D2Du;   // (the word is undotted already)
    u32 nz = (((x & 0x7F7F7F7Fu) + 0x7F7F7F7Fu) | x) & 0x80808080u;
    if (nvalid < first + 4u) nz &= nvalid > first ? (1u << (8u * (nvalid - first))) - 1u : 0u;
    return (u32)__popc(nz);
}

// One expanded column of the transducer, branch-free (the lanes of a warp see different columns; every data
// dependent branch here would serialise them).  ct/cq: bases consumed so far, ot/oq: bases located before this
// column (tile-relative).  Returns the output column, or 0 when it is gap/gap.
__device__ __forceinline__ u32 nw_step(u32 qo, u32 to, bool last, u32& ct, u32& cq, u32& ot, u32& oq, const u8* ts, const u8* qs,
                                       u32 availT, u32 availQ, u32 more, u32& overflow) {
    const bool t_here = to != DASH, q_here = qo != DASH;
    const bool t_own = t_here & (ct == ot), q_own = q_here & (cq == oq);  // this column's base is still unconsumed
    // pull the pending target base into a target gap if it equals the (real) query base   Alignment.cpp:169-181
    const bool t_try = !t_own & q_own & !last;
    u32 pend = 256u;
    if (t_try) { if (ct < availT) pend = ts[ct]; else overflow |= more; }
    const bool t_pull = pend == qo;
    const bool t_has = t_own | t_pull;
    const u32 tc = t_own ? to : (t_pull ? qo : DASH);
    ct += (u32)(!last & t_has);
    // then the pending query base into a query gap if it equals the (new) target base     Alignment.cpp:184-196
    const bool q_try = !q_own & t_has & !last;
    pend = 256u;
    if (q_try) { if (cq < availQ) pend = qs[cq]; else overflow |= more; }
    const bool q_pull = pend == tc;
    const bool q_has = q_own | q_pull;
    const u32 qc = q_own ? qo : (q_pull ? tc : DASH);
    cq += (u32)(!last & q_has);
    ot += (u32)t_here;
    oq += (u32)q_here;
    return (t_has | q_has) ? (qc | (tc << 8)) : 0u;
}

// The lane's chunk (<= 32 raw columns held in 8+8 registers).  lead = (ct-ot) | (cq-oq) << 16 on entry and exit.
// The word loop is NOT unrolled (the registers are rotated instead of indexed): fully unrolled, the three
// instantiations are ~12k straight-line instructions and the kernel becomes instruction-fetch bound.
template <bool EMIT>
__device__ __forceinline__ u32 nw_chunk(u32 (&qw)[8], u32 (&tw)[8], u32 nvalid, bool has_last, u32 ot, u32 oq, u32& lead,
                                        const u8* ts, const u8* qs, u32 availT, u32 availQ, u32 more, u32& overflow, u16* out) {
    u32 ct = ot + (lead & 0xFFFFu), cq = oq + (lead >> 16), nout = 0;
#pragma unroll 1
    for (u32 w = 0; w < 8; w++) {
        const u32 qword = qw[0], tword = tw[0];
#pragma unroll
        for (int k = 0; k < 7; k++) { qw[k] = qw[k + 1]; tw[k] = tw[k + 1]; }
        qw[7] = qword; tw[7] = tword;  // a full turn leaves the registers as they were
#pragma unroll
        for (int b = 0; b < 4; b++) {
            const u32 j = w * 4 + b;
            const bool valid = j < nvalid;
            // columns past the chunk behave as gap/gap: no state change, no output
            const u32 qb = valid ? (qword >> (8 * b)) & 0xFFu : DASH, tb = valid ? (tword >> (8 * b)) & 0xFFu : DASH;   // the words were undotted when they were loaded
            const bool mism = (qb != tb) & (qb != DASH) & (tb != DASH);
            const bool lastcol = has_last & (j + 1 == nvalid);
            // a mismatch is a deletion column followed by an insertion column (Alignment.cpp:147-159)
            u32 c = nw_step(mism ? DASH : qb, tb, lastcol & !mism, ct, cq, ot, oq, ts, qs, availT, availQ, more, overflow);
            if (c) { if (EMIT) out[nout] = (u16)c; nout++; }
            if (mism) {
                c = nw_step(qb, DASH, lastcol, ct, cq, ot, oq, ts, qs, availT, availQ, more, overflow);
                if (c) { if (EMIT) out[nout] = (u16)c; nout++; }
            }
        }
    }
    lead = (ct - ot) | ((cq - oq) << 16);
    return nout;
}

// coalesced staging of `nbytes` (<= NW_W) bytes starting at src into a 16-byte aligned shared buffer
__device__ __forceinline__ void nw_stage(const u8* src, u32 nbytes, u32* dst, u32 lane) {
    const unsigned long long a = (unsigned long long)src;
    const u32 ph = (u32)(a & 3ull);
    const u32* w = (const u32*)(a - ph);
    const u32 nw = (nbytes + 3) >> 2;
    for (u32 i = lane; i < nw; i += 32) {
        const u32 lo = __ldg(w + i), hi = ph ? __ldg(w + i + 1) : 0u;
        dst[i] = __funnelshift_r(lo, hi, 8 * ph);
    }
}

// the same for a string sent as 4-bit codes (DGC_BATCH_PACKED4): `src` = the byte holding the window's first column (the
// windows start on even columns), `ncols` columns.  One prmt turns four codes into four ASCII bytes.
__device__ __forceinline__ void nw_stage_packed(const u8* src, u32 ncols, u32* dst, u32 lane) {
    const unsigned long long a = (unsigned long long)src;
    const u32 ph = (u32)(a & 3ull);
    const u32* w = (const u32*)(a - ph);
    const u32 nw = (((ncols + 1) >> 1) + 3) >> 2;   // packed words: 8 columns each
    for (u32 i = lane; i < nw; i += 32) {
        const u32 lo = __ldg(w + i), hi = ph ? __ldg(w + i + 1) : 0u;
        const u32 x = __funnelshift_r(lo, hi, 8 * ph);
        uint2 o;
        o.x = __byte_perm(PACK4_LUT_LO, PACK4_LUT_HI, x & 0x7777u);
        o.y = __byte_perm(PACK4_LUT_LO, PACK4_LUT_HI, (x >> 16) & 0x7777u);
        ((uint2*)dst)[i] = o;
    }
}

__device__ __forceinline__ u32 warp_incl_scan(u32 v, u32 lane) {
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const u32 x = __shfl_up_sync(0xFFFFFFFFu, v, d);
        if (lane >= (u32)d) v += x;
    }
    return v;
}

__global__ void __launch_bounds__(NW_WARPS * 32, NW_MIN_BLOCKS) k_normalize_warp(NormArgs a) {
    __shared__ NwSmem sm_all[NW_WARPS];
    const u32 lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const u32 ai = a.aln_first + blockIdx.x * NW_WARPS + wid;
    if (ai >= a.n_alns) return;
    NwSmem& sm = sm_all[wid];
    const u32 n = a.aln_len[ai];
    const u64 base = a.aln_norm_off[ai];
    u16* out = a.norm + base;
    if (n < a.min_len) {  // main.cpp:132
        if (lane == 0) {
            a.o_off[ai] = base; a.o_len[ai] = 0; a.o_start[ai] = a.aln_start[ai]; a.o_flags[ai] = AF_SKIP;
            a.o_nI[ai] = 0; a.o_nE[ai] = 0; a.o_tkept[ai] = 0;
        }
        return;
    }
    const u8* qsrc = a.arena + a.aln_q_off[ai];
    const u8* tsrc = a.arena + a.aln_t_off[ai];
    u32 m = 0;           // output columns so far
    u32 carry = 0;       // leads at the tile boundary
    u32 overflow = 0;
    for (u32 c0 = 0; c0 < n; c0 += NW_TR) {
        const u32 nwin = min(NW_W, n - c0);
        const u32 more = c0 + NW_W < n ? 1u : 0u;
        __syncwarp();
        if (a.flags & F_PACKED4) {
            nw_stage_packed(qsrc + (c0 >> 1), nwin, (u32*)sm.q, lane);
            nw_stage_packed(tsrc + (c0 >> 1), nwin, (u32*)sm.t, lane);
        } else {
            nw_stage(qsrc + c0, nwin, (u32*)sm.q, lane);
            nw_stage(tsrc + c0, nwin, (u32*)sm.t, lane);
        }
        __syncwarp();
        // ---- the lane's 32 tile columns and 8 margin columns in registers
        u32 qw[8], tw[8], qm[2], tm[2];
        {
            const uint4 x0 = sm.q[2 * lane], x1 = sm.q[2 * lane + 1], y0 = sm.t[2 * lane], y1 = sm.t[2 * lane + 1];
            qw[0] = x0.x; qw[1] = x0.y; qw[2] = x0.z; qw[3] = x0.w; qw[4] = x1.x; qw[5] = x1.y; qw[6] = x1.z; qw[7] = x1.w;
            tw[0] = y0.x; tw[1] = y0.y; tw[2] = y0.z; tw[3] = y0.w; tw[4] = y1.x; tw[5] = y1.y; tw[6] = y1.z; tw[7] = y1.w;
            const uint2 xm = ((const uint2*)sm.q)[NW_TR / 8 + lane], ym = ((const uint2*)sm.t)[NW_TR / 8 + lane];
            qm[0] = xm.x; qm[1] = xm.y; tm[0] = ym.x; tm[1] = ym.y;
#pragma unroll
            for (int k = 0; k < 8; k++) { qw[k] = nw_undot4(qw[k]); tw[k] = nw_undot4(tw[k]); }
#pragma unroll
            for (int k = 0; k < 2; k++) { qm[k] = nw_undot4(qm[k]); tm[k] = nw_undot4(tm[k]); }
        }
        const u32 col0 = lane * 32;                       // first tile column of this lane (window-relative)
        const u32 nvalid = col0 < min(nwin, NW_TR) ? min(32u, min(nwin, NW_TR) - col0) : 0u;
        const u32 mcol0 = NW_TR + lane * 8;
        const u32 mvalid = mcol0 < nwin ? min(8u, nwin - mcol0) : 0u;
        // ---- base counts, ranks, compacted streams
        u32 nT = 0, nQ = 0;
#pragma unroll
        for (int k = 0; k < 8; k++) { nT += nw_count_bases(tw[k], nvalid, 4 * k); nQ += nw_count_bases(qw[k], nvalid, 4 * k); }
        u32 mT = 0, mQ = 0;
#pragma unroll
        for (int k = 0; k < 2; k++) { mT += nw_count_bases(tm[k], mvalid, 4 * k); mQ += nw_count_bases(qm[k], mvalid, 4 * k); }
        const u32 inc = warp_incl_scan(nT | (nQ << 16), lane);
        const u32 tot = __shfl_sync(0xFFFFFFFFu, inc, 31);
        const u32 ex = inc - (nT | (nQ << 16));
        const u32 minc = warp_incl_scan(mT | (mQ << 16), lane);
        const u32 mtot = __shfl_sync(0xFFFFFFFFu, minc, 31);
        const u32 mex = tot + minc - (mT | (mQ << 16));
        const u32 ot0 = ex & 0xFFFFu, oq0 = ex >> 16;
        {
            u32 rt = ot0, rq = oq0;
#pragma unroll
            for (int j = 0; j < 32; j++)
                if ((u32)j < nvalid) {
                    const u32 tb = nw_byte(tw, j), qb = nw_byte(qw, j);
                    if (tb != DASH) sm.ts[rt++] = (u8)tb;
                    if (qb != DASH) sm.qs[rq++] = (u8)qb;
                }
            rt = mex & 0xFFFFu; rq = mex >> 16;
#pragma unroll
            for (int j = 0; j < 8; j++)
                if ((u32)j < mvalid) {
                    const u32 tb = nw_byte(tm, j), qb = nw_byte(qm, j);
                    if (tb != DASH) sm.ts[rt++] = (u8)tb;
                    if (qb != DASH) sm.qs[rq++] = (u8)qb;
                }
        }
        const u32 availT = (tot & 0xFFFFu) + (mtot & 0xFFFFu), availQ = (tot >> 16) + (mtot >> 16);
        __syncwarp();
        // ---- speculative pass + repair rounds (counts only)
        const bool has_last = nvalid && (c0 + col0 + nvalid == n);
        u32 lead_in = lane == 0 ? carry : 0u, lead_out = lead_in;
        // every pass writes the lane's columns to its staging row: when the speculation holds (the common case) the
        // transducer has run once, not twice, and the tile goes out with coalesced stores
        u16* const stg = sm.stage + lane * NW_STG;
        u32 nout = nw_chunk<true>(qw, tw, nvalid, has_last, ot0, oq0, lead_out, sm.ts, sm.qs, availT, availQ, more, overflow, stg);
        for (;;) {
            u32 prev = __shfl_up_sync(0xFFFFFFFFu, lead_out, 1);
            if (lane == 0) prev = lead_in;
            const bool bad = nvalid != 0 && prev != lead_in;  // empty lanes past the end carry no state
            if (!__any_sync(0xFFFFFFFFu, bad)) break;
            if (bad) {
                lead_in = prev; lead_out = prev;
                nout = nw_chunk<true>(qw, tw, nvalid, has_last, ot0, oq0, lead_out, sm.ts, sm.qs, availT, availQ, more, overflow, stg);
            }
        }
        if (__any_sync(0xFFFFFFFFu, overflow != 0)) { overflow = 1; break; }
        // ---- emit
        const u32 oinc = warp_incl_scan(nout, lane);
        __syncwarp();
#pragma unroll 1
        for (u32 l4 = 0; l4 < 32; l4 += 4) {  // four rows of the staging area at a time, eight lanes each: a row holds 32-40
            const u32 l = l4 + (lane >> 3);     // columns, and consecutive rows are consecutive in the output
            const u32 nl = __shfl_sync(0xFFFFFFFFu, nout, l), dl = __shfl_sync(0xFFFFFFFFu, oinc - nout, l);
            for (u32 i = lane & 7u; i < nl; i += 8) out[m + dl + i] = sm.stage[l * NW_STG + i];
        }
        m += __shfl_sync(0xFFFFFFFFu, oinc, 31);
        carry = __shfl_sync(0xFFFFFFFFu, lead_out, 31);
    }
    if (overflow) {
        if (lane == 0) a.o_flags[ai] = AF_RETRY;
        return;
    }
    __syncwarp();
    // ---- trimAln (Alignment.cpp:219-242): lo = one past the column holding the trim-th target base from the
    // left, hi = the column holding the trim-th target base from the right (never left of lo)
    u32 lo = 0, hi = m, lb = 0, rb = 0;
    if (!(a.flags & F_NO_TRIM) && a.trim) {
        for (u32 c = 0; c < m && lb < a.trim; c += 32) {
            const bool tb = (c + lane < m) && (out[c + lane] >> 8) != DASH;
            const u32 bal = __ballot_sync(0xFFFFFFFFu, tb), cnt = __popc(bal);
            if (lb + cnt >= a.trim) { lo = c + __fns(bal, 0, a.trim - lb) + 1; lb = a.trim; break; }
            lb += cnt;
            lo = min(m, c + 32);
        }
        for (u32 e = m; e > lo && rb < a.trim;) {  // columns [e-32, e) clipped at lo, scanned from the right
            const u32 c = e - 1 - lane;            // lane 0 = rightmost column
            const bool ok = e >= lane + 1 && c >= lo;
            const bool tb = ok && (out[c] >> 8) != DASH;
            const u32 bal = __ballot_sync(0xFFFFFFFFu, tb), cnt = __popc(bal);
            if (rb + cnt >= a.trim) { hi = e - 1 - __fns(bal, 0, a.trim - rb); rb = a.trim; break; }
            rb += cnt;
            hi = e > lo + 32 ? e - 32 : lo;
            e = hi;
        }
    }
    // ---- what addAln will create from the kept columns (same classes as AlnGraphBoost.cpp:75-104)
    u32 nI = 0, nE = 0, tk = 0;
    for (u32 c = lo + lane; c < hi; c += 32) {
        const u32 col = out[c], q = col & 0xFFu, t = col >> 8;
        const bool match = q == t, del = !match && q == DASH, ins = !match && !del && t == DASH;
        bool prev_match = false;
        if (c > lo) { const u32 pc = out[c - 1]; prev_match = (pc & 0xFFu) == (pc >> 8); }
        nI += ins;
        tk += match || del;
        nE += ins || (match && !prev_match);
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) {
        nI += __shfl_xor_sync(0xFFFFFFFFu, nI, d);
        nE += __shfl_xor_sync(0xFFFFFFFFu, nE, d);
        tk += __shfl_xor_sync(0xFFFFFFFFu, tk, d);
    }
    if (lane == 0) {
        a.o_off[ai] = base + lo;
        a.o_len[ai] = hi - lo;
        a.o_start[ai] = a.aln_start[ai] + lb;
        a.o_flags[ai] = 0;
        a.o_nI[ai] = nI;
        a.o_nE[ai] = nE + 1;  // + the edge into the exit vertex
        a.o_tkept[ai] = tk;
    }
}
