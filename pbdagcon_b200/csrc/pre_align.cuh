// pre_align.cuh -- the `-a` step on the device: one 'pre' record (ungapped query and target segments) -> gapped strings.
//
// Stands where reference src/cpp/SimpleAligner.cpp:25-63 stands (called from main.cpp:127-129).  The reference hands the
// arithmetic to blasr_libcpp (SDPAlign + GuidedAlign), whose source is not part of the reference tree: PARITY WITH
// BLASR IS UNPINNED; this kernel is held bit for bit to the host aligner of this repository
// (pbdagcon_b200/host/simple_align.cpp), which in turn passes the reference's one known-answer vector
// (test/cpp/SimpleAlignerTest.cpp:8-21).  Same method, same tie-breaks: global DP in a band of 2*band+1 columns around
// the scaled diagonal (row i centred on column i*m/n), costs match -5 / mismatch +6 / insertion 4 / deletion 5
// (SimpleAligner.cpp:10-23), traceback preferring diagonal, then deletion, then insertion; the band doubles until the
// path keeps two cells clear of its edges; the result is cut to start and end on a matching column.
//
// Mapping: one WARP per record.  A DP row is 2*band+1 cells; the lanes take 32 cells at a time.  The only dependency
// inside a row is the deletion move cur[k-1] + 5, which is a prefix minimum with a linear penalty:
//   cur[k] = 5k + min_{j<=k} (h[j] - 5j),  h[j] = min(diagonal move, insertion move)
// i.e. one warp min-scan per 32 cells, exact in integers.  Move codes go to a byte matrix in global scratch; lane 0
// walks it back and writes the strings in reverse -- which is also the order the reverse complement of a '-' strand
// record wants (SimpleAligner.cpp:54-61).
#pragma once

#define PA_INF (1 << 29)
#define PA_G 4            // guard cells on both sides of a row (a row shifts by at most 3 columns)
#define PA_MAX_BAND 8192   // widest band the device tries (two rows of 2*band+1 ints in shared memory, one warp per block)

struct PreArgs {
    const u32* rec_ids;        // records of this launch (indices), or nullptr = all
    u32 n_launch;
    u32 band;
    u32 warps;                 // warps (= records) per block; the rows live in dynamic shared memory
    const u32* rec_start;      // 'pre' tstart: 0-based, on the record's strand (Alignment.cpp:95-111)
    const u32* rec_tlen;       // length of the whole target (dagcon::Alignment::tlen)
    u32* aln_start;            // out: dagcon::Alignment::start after SimpleAligner.cpp:49-62 (1-based, target orientation)
    const u8* arena_in;        // ungapped segments (ASCII)
    const u64* q_off; const u64* t_off; const u32* q_len; const u32* t_len;
    const u32* strand;         // 1 = '-': the output is reverse-complemented (SimpleAligner.cpp:54-61)
    u8* tb; const u64* tb_off; // move matrix slabs, one per launch slot
    // output
    u8* arena_out; const u64* out_q_off; const u64* out_t_off;
    u32* out_len;              // gapped columns (0 = nothing aligns)
    u32* t_begin; u32* t_end;  // target bases of the segment the result covers
    u32* state;                // 0 pending, 1 done, 2 needs a wider band, 3 nothing aligns
};

__device__ __forceinline__ long long pa_centre(long long i, long long n, long long m) { return n ? (i * m) / n : 0; }
__device__ __forceinline__ u8 pa_comp(u8 c) {   // revComp of Alignment.cpp:15-26: upper-case ACGT only
    return c == 'A' ? 'T' : c == 'T' ? 'A' : c == 'C' ? 'G' : c == 'G' ? 'C' : c;
}

__global__ void __launch_bounds__(128) k_pre_align(PreArgs a) {
    extern __shared__ int s_rows[];   // per warp: two rows of (2*a.band + 1 + 2*PA_G) ints
    const u32 lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const u32 slot = blockIdx.x * a.warps + wid;
    if (slot >= a.n_launch) return;
    const u32 r = a.rec_ids ? a.rec_ids[slot] : slot;
    const long long n = a.q_len[r], m = a.t_len[r];
    const long long full = (n > m ? n : m) + 1;
    const long long band = a.band < full ? a.band : full;
    const bool final_pass = a.band >= full;   // the whole matrix is inside the band: nothing wider to try
    if (n == 0 || m == 0) {
        if (lane == 0) { a.state[r] = 3; a.out_len[r] = 0; a.t_begin[r] = 0; a.t_end[r] = 0; a.aln_start[r] = a.rec_start[r] + 1; }
        return;
    }
    const int W = (int)(2 * band + 1);
    const u8* q = a.arena_in + a.q_off[r];
    const u8* t = a.arena_in + a.t_off[r];
    u8* tb = a.tb + a.tb_off[slot];
    const u32 stride = 2 * a.band + 1 + 2 * PA_G;
    int* prev = s_rows + (size_t)wid * 2 * stride + PA_G;
    int* cur = prev + stride;
    for (int k = -PA_G + (int)lane; k < W + PA_G; k += 32) { prev[k] = PA_INF; cur[k] = PA_INF; }
    __syncwarp();
    {   // row 0: leading deletions
        const long long c0 = pa_centre(0, n, m);
        for (int k = (int)lane; k < W; k += 32) {
            const long long j = c0 - band + k;
            const bool in = j >= 0 && j <= m;
            if (in) prev[k] = (int)(j * 5);
            tb[k] = in ? 2 : 0;
        }
    }
    __syncwarp();
    bool fail = false;
    for (long long i = 1; i <= n; i++) {
        const long long c1 = pa_centre(i, n, m), sh = c1 - pa_centre(i - 1, n, m);
        if (sh > PA_G - 1) { fail = true; break; }   // a target more than 3x the query: the host aligner gives up as well
        u8* tbr = tb + (size_t)i * W;
        const u32 qc = q[i - 1];
        const long long klo_l = band - c1 > 0 ? band - c1 : 0, khi_l = (W - 1) < (m - c1 + band) ? (W - 1) : (m - c1 + band);
        const int klo = (int)klo_l, khi = (int)khi_l;
        const long long tj = c1 - band - 1;           // t[j - 1] = t[tj + k]
        int carry = PA_INF;                           // cur[k-1] of the chunk's first cell, as (value - 5*(k-1)) once valid
        bool have_carry = false;
        int carry_val = PA_INF;
        for (int kb = 0; kb < W; kb += 32) {
            const int k = kb + (int)lane;
            const bool inside = k < W && k >= klo && k <= khi;
            const bool col0 = inside && (c1 - band + k == 0);
            int vd = PA_INF + 6, vu = PA_INF + 4;
            if (inside) {
                vu = prev[k + (int)sh] + 4;
                if (!col0) vd = prev[k + (int)sh - 1] + (qc == (u32)t[tj + k] ? -5 : 6);
            }
            const int h = inside ? (col0 ? vu : (vd < vu ? vd : vu)) : (PA_INF + 6);
            // prefix minimum of (h[j] - 5j) over the row so far, then + 5k; cells outside the row never win
            int v = inside ? h - 5 * k : (1 << 30);
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int x = __shfl_up_sync(0xFFFFFFFFu, v, d);
                if ((int)lane >= d && x < v) v = x;
            }
            if (have_carry && carry < v) v = carry;
            int best = v + 5 * k;
            // the move: diagonal first, deletion (from the left) if strictly better, insertion if strictly better still
            int left;   // cur[k-1] as the sequential recurrence sees it (clamped to INF when that cell is dead)
            {
                int pv = __shfl_up_sync(0xFFFFFFFFu, best, 1);
                const bool pin = __shfl_up_sync(0xFFFFFFFFu, (int)inside, 1) != 0;
                if (lane == 0) { pv = carry_val; left = have_carry ? pv : PA_INF; }
                else left = pin ? pv : PA_INF;
                if (left >= PA_INF) left = PA_INF;
            }
            u8 dir = 0;
            int val = PA_INF;
            if (inside) {
                if (col0) {
                    val = vu < PA_INF ? vu : PA_INF;
                    dir = vu < PA_INF ? 3 : 0;
                } else {
                    int b = vd;
                    dir = 1;
                    const int vl = left + 5;
                    if (vl < b) { b = vl; dir = 2; }
                    if (vu < b) { b = vu; dir = 3; }
                    if (b >= PA_INF) { b = PA_INF; dir = 0; }
                    val = b;
                }
            }
            if (k < W) { cur[k] = val; tbr[k] = dir; }
            // carry to the next chunk: the running minimum and the last cell's value
            const int lastv = __shfl_sync(0xFFFFFFFFu, v, 31);
            const int lastval = __shfl_sync(0xFFFFFFFFu, val, 31);
            carry = lastv; have_carry = true; carry_val = lastval;
            (void)best;
        }
        __syncwarp();
        int* x = prev; prev = cur; cur = x;
    }
    __syncwarp();
    // ---- traceback (lane 0), strings written back to front into the record's output slot
    u32 st = 1, len = 0, tbeg = 0, tend = (u32)m;
    if (lane == 0) {
        if (fail) st = final_pass ? 3 : 2;
        long long i = n, j = m;
        if (!fail) {
            const long long k = j - (pa_centre(i, n, m) - band);
            if (k < 0 || k >= W || prev[k] >= PA_INF) st = final_pass ? 3 : 2;
        }
        u8* oq = a.arena_out + a.out_q_off[r];
        u8* ot = a.arena_out + a.out_t_off[r];
        const u32 cap = (u32)(n + m);
        // pass 1: walk back, remember the path as moves in the (now free) slot tail: write the REVERSED strings at the end
        // of the slot, then cut and move (or complement) them into place
        u32 w = cap;   // next free position from the back
        if (st == 1) {
            while (i > 0 || j > 0) {
                const long long k = j - (pa_centre(i, n, m) - band);
                if (k < 0 || k >= W) { st = final_pass ? 3 : 2; break; }
                if (!final_pass && (k < 2 || k > W - 3) && j > 0 && j < m) { st = 2; break; }
                const u8 dir = i == 0 ? (u8)2 : tb[(size_t)i * W + k];
                u8 cq, ct;
                if (dir == 1) { cq = q[--i]; ct = t[--j]; }
                else if (dir == 2) { cq = '-'; ct = t[--j]; }
                else if (dir == 3) { cq = q[--i]; ct = '-'; }
                else { st = final_pass ? 3 : 2; break; }
                --w;
                oq[w] = cq; ot[w] = ct;      // forward order: column c of the alignment sits at w0 + c
            }
        }
        if (st == 1) {
            // local result: cut the columns before the first and after the last matching column
            const u32 total = cap - w;
            u32 lo = 0, hi = total;
            while (lo < hi && oq[w + lo] != ot[w + lo]) { tbeg += ot[w + lo] != '-'; lo++; }
            while (hi > lo && oq[w + hi - 1] != ot[w + hi - 1]) { tend -= ot[w + hi - 1] != '-'; hi--; }
            if (lo >= hi) st = 3;
            else {
                len = hi - lo;
                if (a.strand[r]) {   // reverse complement in place: output column c = comp(column len-1-c)
                    // the kept columns sit at [w+lo, w+hi) and the output goes to [0, len): w + lo >= len is not guaranteed,
                    // so reverse inside the source range first, then move down
                    for (u32 x = 0, y = len - 1; x < y; x++, y--) {
                        u8 c1 = oq[w + lo + x], c2 = oq[w + lo + y]; oq[w + lo + x] = pa_comp(c2); oq[w + lo + y] = pa_comp(c1);
                        c1 = ot[w + lo + x]; c2 = ot[w + lo + y]; ot[w + lo + x] = pa_comp(c2); ot[w + lo + y] = pa_comp(c1);
                    }
                    if (len & 1) { const u32 mid = w + lo + len / 2; oq[mid] = pa_comp(oq[mid]); ot[mid] = pa_comp(ot[mid]); }
                }
                if (w + lo) for (u32 c = 0; c < len; c++) { oq[c] = oq[w + lo + c]; ot[c] = ot[w + lo + c]; }
            }
        }
        if (st == 3) { len = 0; tbeg = 0; tend = 0; }
        a.state[r] = st;
        if (st != 2) {
            a.out_len[r] = len; a.t_begin[r] = tbeg; a.t_end[r] = tend;
            // SimpleAligner.cpp:49-62, to the letter: start moves by the cut, end = the MOVED start + the segment-relative
            // end; a '-' record is flipped into target orientation; then 1-based.  A record that does not align keeps its
            // place in the target's alignment count (main.cpp:66-72,118) as a zero-column alignment.
            u32 s0 = a.rec_start[r] + tbeg;
            if (st == 1 && a.strand[r]) { const u32 e0 = s0 + tend; s0 = a.rec_tlen[r] - e0; }
            a.aln_start[r] = s0 + 1;
        }
    }
}
