// daz_realign.cuh -- dazcon's trace-point realignment on the device.
//
// Replaces, for LAS files with byte-wide trace values (trace spacing <= 125, daligner's default 100), the work the
// reference's single Reader thread does for every selected overlap (reference src/cpp/DazAlnProvider.cpp:304-369):
//   Load_Subread x2 + Complement_Seq          DAZZ_DB/DB.c:1392-1440, DALIGNER/align.c:3190-3205
//   Compute_Trace_PTS(GREEDIEST)              DALIGNER/align.c:4799-4896: one iter_np per trace segment
//   iter_np                                   DALIGNER/align.c:4132-4459: O(NP) waves + greedy traceback
//   decodeAlignment                           DazAlnProvider.cpp:383-417: trace -> gapped ASCII strings
// Only the 2-bit reads and the trace points cross PCIe; the gapped strings are written straight into the batch
// arena the consensus kernels read.
//
// Mapping.  The trace points cut an overlap into independent segments (A advances to the next multiple of the trace
// spacing, B by the recorded amount), so k_daz_iter runs one THREAD per segment: both substrings are packed 2 bits per
// base into registers/local words and the furthest-point slide compares 16 bases per XOR; the waves (furthest point
// and move code per (wave, diagonal)) live in a per-thread slab of global scratch that stays L1/L2 resident.
// decodeAlignment is a transducer over the concatenated edit script whose state after entry n is a closed form of
// the entry's value and the counts of A-gaps / B-gaps before it, so k_daz_emit also runs one thread per segment.
#pragma once

#define DZ_AW 9    // packed A segment: up to 128 bases + one slack word
#define DZ_BW 18   // packed B segment: up to 272 bases + one slack word
#define DZ_MAX_M 128
#define DZ_MAX_N 256

struct DazDev {            // the trimmed DAZZ_DB, resident in HBM
    const u8* bases;       // the .bps file: 4 bases per byte, first base in the top bits (DB.c Compress_Read)
    const u64* boff;       // per read: byte offset of its first base
    const u32* rlen;
};

struct DazArgs {
    DazDev db;
    u32 n_ovl, n_seg, tspace;
    // per overlap
    const u32* ovl_aread; const u32* ovl_bread; const u32* ovl_comp;
    const int* ovl_abpos; const int* ovl_aepos; const int* ovl_bbpos; const int* ovl_bepos;
    const u64* ovl_trace_off; const u32* ovl_tlen;   // into `trace` (pairs: diffs, b advance)
    const u32* ovl_seg_begin;                        // CSR over segments
    const u64* ovl_entry_off;                        // first edit-script slot of the overlap
    const u8* trace;
    // per segment (written by k_daz_prep / k_daz_iter / k_daz_scan)
    int* seg_bb;          // B position where the segment starts
    u32* seg_eoff;        // first edit-script slot, relative to the overlap's
    u16* seg_cnt;         // edit-script entries
    u16* seg_neg;         // ... of which gaps in A (B base against '-')
    u16* seg_ppos; u16* seg_pneg;   // entries with gaps in B / in A before the segment (within the overlap)
    int* entries;
    // waves
    short* vf; signed char* hf; u32 rows, width, koff, n_slabs;
    // output
    u8* arena; const u64* aln_q_off; const u64* aln_t_off; u32* aln_len;
    u32* error;           // set when a segment needs more waves than its trace point promised
};

__device__ __forceinline__ u32 dz_base(const DazDev& db, u64 boff, int p) {
    return (db.bases[boff + (u32)(p >> 2)] >> (6 - 2 * (p & 3))) & 3u;
}
// base i of a read as the reference's buffers hold it: 4 (the terminator Load_Subread writes) at index len
__device__ __forceinline__ u32 dz_read_a(const DazDev& db, u64 boff, int len, int i) {
    return (u32)i >= (u32)len ? 4u : dz_base(db, boff, i);
}
// ... and of the B read, reverse-complemented when the overlap says so (Complement_Seq)
__device__ __forceinline__ u32 dz_read_b(const DazDev& db, u64 boff, int len, bool comp, int j) {
    if ((u32)j >= (u32)len) return 4u;
    return comp ? 3u - dz_base(db, boff, len - 1 - j) : dz_base(db, boff, j);
}

// ToU of DazAlnProvider.cpp:18 for the codes that can occur here: 0..3 = ACGT, 4 = the read terminator '.'
__device__ __forceinline__ u8 dz_tou(u32 c) { return c < 4u ? (u8)(0x54474341u >> (8 * c)) : (u8)'.'; }

// 16 bases starting at base `pos` of a packed string (base p in bits 2*(p%16) of word p/16)
__device__ __forceinline__ u32 dz_win(const u32* w, int pos) {
    const int wi = pos >> 4;
    return __funnelshift_r(w[wi], w[wi + 1], (pos & 15) * 2);
}
// number of leading positions (at most n) where A[ai..] equals B[bj..]
__device__ __forceinline__ int dz_slide(const u32* aw, const u32* bw, int ai, int bj, int n) {
    int done = 0;
    while (done < n) {
        const u32 x = dz_win(aw, ai + done) ^ dz_win(bw, bj + done);
        int m = x ? (__ffs((int)x) - 1) >> 1 : 16;
        if (m > n - done) m = n - done;
        done += m;
        if (x) break;
    }
    return done;
}

// ---- per overlap: where every segment starts in B and where its edit-script slots are
__global__ void __launch_bounds__(128) k_daz_prep(DazArgs a) {
    const u32 o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= a.n_ovl) return;
    const u8* pts = a.trace + a.ovl_trace_off[o];
    const u32 nseg = a.ovl_tlen[o] / 2, s0 = a.ovl_seg_begin[o];
    int bb = a.ovl_bbpos[o];
    u32 eoff = 0;
    for (u32 s = 0; s < nseg; s++) {
        a.seg_bb[s0 + s] = bb;
        a.seg_eoff[s0 + s] = eoff;
        eoff += pts[2 * s];
        bb += pts[2 * s + 1];
    }
}

// ---- iter_np (DALIGNER/align.c:4132-4459, GREEDIEST) for one segment per thread
// Move codes as in the reference: -1 / 2 = arrived from diagonal k-1 (two waves back / this wave), 1 / 4 = from k+1
// (two waves back / this wave), 0 = substitution on k (previous wave), 3 = end marker.
#define DZ_VF(d, k) vf[(size_t)((d) + 2) * W + (size_t)((k) + KO)]
#define DZ_HF(d, k) hf[(size_t)((d) + 2) * W + (size_t)((k) + KO)]
__global__ void __launch_bounds__(128) k_daz_iter(DazArgs a) {
    const u32 slab = blockIdx.x * blockDim.x + threadIdx.x;
    if (slab >= a.n_slabs) return;
    const int W = (int)a.width, KO = (int)a.koff, ROWS = (int)a.rows;
    short* vf = a.vf + (size_t)slab * ROWS * W;
    signed char* hf = a.hf + (size_t)slab * ROWS * W;
    u32 o = 0;   // overlap of the current segment (segments are visited in increasing order: search forwards only)
    for (u32 sg = slab; sg < a.n_seg; sg += a.n_slabs) {
        {
            u32 lo = o, hi = a.n_ovl;
            while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (a.ovl_seg_begin[mid] <= sg) lo = mid; else hi = mid; }
            o = lo;
        }
        const u32 si = sg - a.ovl_seg_begin[o], nseg = a.ovl_tlen[o] / 2;
        const int abpos = a.ovl_abpos[o], aepos = a.ovl_aepos[o], bepos = a.ovl_bepos[o];
        const int ts = (int)a.tspace;
        const int ab = si == 0 ? abpos : (abpos / ts + (int)si) * ts;
        const int ae = si + 1 == nseg ? aepos : (abpos / ts + (int)si + 1) * ts;
        const int bb = a.seg_bb[sg];
        const int be = si + 1 == nseg ? bepos : bb + (int)a.trace[a.ovl_trace_off[o] + 2 * si + 1];
        const int M = ae - ab, N = be - bb;
        const int dcap = (int)a.trace[a.ovl_trace_off[o] + 2 * si];   // the waves this segment may need
        int* out = a.entries + a.ovl_entry_off[o] + a.seg_eoff[sg];
        if (M < 0 || N < 0 || M > DZ_MAX_M || N > DZ_MAX_N) { atomicOr(a.error, 1u); a.seg_cnt[sg] = 0; a.seg_neg[sg] = 0; continue; }
        // ---- pack both substrings
        u32 aw[DZ_AW], bw[DZ_BW];
#pragma unroll
        for (int i = 0; i < DZ_AW; i++) aw[i] = 0;
#pragma unroll
        for (int i = 0; i < DZ_BW; i++) bw[i] = 0;
        {
            const u64 aoff = a.db.boff[a.ovl_aread[o]];
            for (int i = 0; i < M; i++) aw[i >> 4] |= dz_base(a.db, aoff, ab + i) << (2 * (i & 15));
            const u32 br = a.ovl_bread[o];
            const u64 boffb = a.db.boff[br];
            const int blen = (int)a.db.rlen[br];
            const bool comp = a.ovl_comp[o] != 0;
            for (int j = 0; j < N; j++) {
                const int p = bb + j;
                const u32 c = comp ? 3u - dz_base(a.db, boffb, blen - 1 - p) : dz_base(a.db, boffb, p);
                bw[j >> 4] |= c << (2 * (j & 15));
            }
        }
        // ---- forward waves
        const int del = M - N;
        int low = del < 0 ? del : 0, hgh = del > 0 ? del : 0;
        for (int k = low - 1; k <= hgh + 1; k++) { DZ_VF(-2, k) = -2; DZ_VF(-1, k) = -2; }
        DZ_VF(-1, 0) = -1;
        low += 1; hgh -= 1;
        int D = 0;
        bool failed = false;
        for (;; D++) {
            if (D + 2 >= ROWS) { failed = true; break; }
            if ((D & 1) == 0) { low -= 1; hgh += 1; }
            if (low - 1 + KO < 0 || hgh + 1 + KO >= W) { failed = true; break; }
            DZ_VF(D, hgh + 1) = -2; DZ_VF(D, low - 1) = -2;
            int j = -2;
            // the furthest point on diagonal k of this wave: best of "from k-1" (am), "on k" (ac), "from k+1" (ap), with
            // the reference's preference order; then slide along the diagonal while the bases agree
#define DZ_MOVE(mdir, pdir)                                                         \
            {                                                                       \
                const int ac = (int)DZ_VF(D - 1, k) + 1;                            \
                int code;                                                           \
                if (ac < am) { if (ap < am) { code = (mdir); j = am; } else { code = (pdir); j = ap; } } \
                else { if (ap < ac) { code = 0; j = ac; } else { code = (pdir); j = ap; } } \
                DZ_HF(D, k) = (signed char)code;                                    \
                const int lim = N < M - k ? N : M - k;                              \
                if (j < lim) j += dz_slide(aw, bw, k + j, j, lim - j);              \
                DZ_VF(D, k) = (short)j;                                             \
            }
            for (int k = hgh; k > del; k--) {
                const int ap = j + 1, am = (int)DZ_VF(D - 2, k - 1);
                DZ_MOVE(-1, 4)
            }
            j = -2;
            for (int k = low; k < del; k++) {
                const int ap = (int)DZ_VF(D - 2, k + 1) + 1, am = j;
                DZ_MOVE(2, 1)
            }
            {
                const int k = del;
                const int ap = (int)DZ_VF(D, del + 1) + 1, am = j;
                DZ_MOVE(2, 4)
            }
#undef DZ_MOVE
            if ((int)DZ_VF(D, del) >= N) break;
        }
        if (failed) { atomicOr(a.error, 2u); a.seg_cnt[sg] = 0; a.seg_neg[sg] = 0; continue; }
        // ---- reverse the move pointers from (D, del) back to (0, 0), then walk them forwards emitting the script
        {
            DZ_HF(0, 0) = 3;
            int k = del, e = DZ_HF(D, k);
            DZ_HF(D, k) = 3;
            while (e != 3) {
                int h = k + e;
                if (e > 1) h -= 3;
                else if (e == 0) D -= 1;
                else D -= 2;
                const int m = DZ_HF(D, h);
                DZ_HF(D, h) = (signed char)e;
                e = m;
                k = h;
            }
        }
        u32 cnt = 0, neg = 0;
        {
            int k = 0, e;
            D = 0;
            e = DZ_HF(0, 0);
            while (e != 3) {
                int h = k - e;
                const int c = DZ_VF(D, k);
                if (e > 1) h += 3;
                else if (e == 0) D += 1;
                else D += 2;
                if (h != k && cnt >= (u32)dcap) { failed = true; break; }   // more edits than the trace point recorded: no slot
                if (h > k) out[cnt++] = bb + 1 + c;                 // A base against a gap in B: the B position, 1-based
                else if (h < k) { out[cnt++] = -(ab + 1 + c + k); neg++; }   // B base against a gap in A: minus the A position
                k = h;
                e = DZ_HF(D, h);
            }
        }
        if (failed) { atomicOr(a.error, 4u); cnt = 0; neg = 0; }
        a.seg_cnt[sg] = (u16)cnt;
        a.seg_neg[sg] = (u16)neg;
    }
}
#undef DZ_VF
#undef DZ_HF

// ---- per overlap: entry counts before every segment, and the number of columns decodeAlignment will write
__global__ void __launch_bounds__(128) k_daz_scan(DazArgs a) {
    const u32 o = blockIdx.x * blockDim.x + threadIdx.x;
    if (o >= a.n_ovl) return;
    const u32 nseg = a.ovl_tlen[o] / 2, s0 = a.ovl_seg_begin[o];
    u32 ppos = 0, pneg = 0;
    for (u32 s = 0; s < nseg; s++) {
        a.seg_ppos[s0 + s] = (u16)ppos;
        a.seg_pneg[s0 + s] = (u16)pneg;
        const u32 c = a.seg_cnt[s0 + s], n = a.seg_neg[s0 + s];
        pneg += n;
        ppos += c - n;
    }
    // every column consumes an A base except the gaps in A; the tail runs to aepos inclusive (DazAlnProvider.cpp:412-416)
    a.aln_len[o] = (u32)(a.ovl_aepos[o] - a.ovl_abpos[o] + 1) + pneg;
}

// ---- decodeAlignment (DazAlnProvider.cpp:383-417), one thread per segment
// State after an entry: (i, j) = next A / B index.  With d = (abpos - bbpos) + (#gaps-in-B - #gaps-in-A so far):
// an entry -p leaves i = p, j = p - d; an entry +p leaves j = p, i = p + d.  Column index = (i - abpos) + #gaps-in-A.
__global__ void __launch_bounds__(128) k_daz_emit(DazArgs a) {
    const u32 sg = blockIdx.x * blockDim.x + threadIdx.x;
    if (sg >= a.n_seg) return;
    // the overlap this segment belongs to: binary search in the CSR
    u32 lo = 0, hi = a.n_ovl;
    while (hi - lo > 1) { const u32 mid = (lo + hi) >> 1; if (a.ovl_seg_begin[mid] <= sg) lo = mid; else hi = mid; }
    const u32 o = lo, s0 = a.ovl_seg_begin[o], nseg = a.ovl_tlen[o] / 2, si = sg - s0;
    const u32 cnt = a.seg_cnt[sg];
    const bool last = si + 1 == nseg;
    if (cnt == 0 && !last) return;
    const int abpos = a.ovl_abpos[o], bbpos = a.ovl_bbpos[o], aepos = a.ovl_aepos[o];
    const int* ent = a.entries + a.ovl_entry_off[o];
    // state after the last entry before this segment
    int i = abpos, j = bbpos;
    int npos = (int)a.seg_ppos[sg], nneg = (int)a.seg_pneg[sg];
    if (npos + nneg > 0) {
        u32 ps = si;
        do { ps--; } while (a.seg_cnt[s0 + ps] == 0);      // entries exist before us: terminates
        const int v = ent[a.seg_eoff[s0 + ps] + a.seg_cnt[s0 + ps] - 1];
        const int d = (abpos - bbpos) + (npos - nneg);
        if (v < 0) { i = -v; j = i - d; } else { j = v; i = j + d; }
    }
    const u32 ar = a.ovl_aread[o], br = a.ovl_bread[o];
    const u64 aoff = a.db.boff[ar], boffb = a.db.boff[br];
    const int alen = (int)a.db.rlen[ar], blen = (int)a.db.rlen[br];
    const bool comp = a.ovl_comp[o] != 0;
    u8* q = a.arena + a.aln_q_off[o];
    u8* t = a.arena + a.aln_t_off[o];
    u32 col = (u32)(i - abpos) + (u32)nneg;
#define DZ_CH(c) dz_tou(c)
    const int* e = ent + a.seg_eoff[sg];
    for (u32 n = 0; n < cnt; n++) {
        const int v = e[n];
        if (v < 0) {
            for (const int stop = -v; i != stop; i++, j++, col++) {
                t[col] = DZ_CH(dz_read_a(a.db, aoff, alen, i)); q[col] = DZ_CH(dz_read_b(a.db, boffb, blen, comp, j));
            }
            t[col] = '-'; q[col] = DZ_CH(dz_read_b(a.db, boffb, blen, comp, j));
            j++; col++;
        } else {
            for (; j != v; i++, j++, col++) {
                t[col] = DZ_CH(dz_read_a(a.db, aoff, alen, i)); q[col] = DZ_CH(dz_read_b(a.db, boffb, blen, comp, j));
            }
            t[col] = DZ_CH(dz_read_a(a.db, aoff, alen, i)); q[col] = '-';
            i++; col++;
        }
    }
    if (last)
        for (; i <= aepos; i++, j++, col++) {
            t[col] = DZ_CH(dz_read_a(a.db, aoff, alen, i)); q[col] = DZ_CH(dz_read_b(a.db, boffb, blen, comp, j));
        }
#undef DZ_CH
}

// ---- the A read of every target as ASCII (DazAlnProvider.cpp:119-131): the backbone of AlnGraphBoost(const std::string&)
__global__ void __launch_bounds__(256) k_daz_backbone(DazDev db, u32 n_targets, const u32* tgt_aread, const u64* tgt_bb_off, u8* arena) {
    const u32 t = blockIdx.x;
    if (t >= n_targets) return;
    const u32 r = tgt_aread[t];
    const u64 boff = db.boff[r];
    const int len = (int)db.rlen[r];
    u8* out = arena + tgt_bb_off[t];
    for (int i = threadIdx.x; i < len; i += blockDim.x) out[i] = dz_tou(dz_base(db, boff, i));
}
