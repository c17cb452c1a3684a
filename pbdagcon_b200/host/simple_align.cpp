// simple_align.cpp -- the `-a` step: turn one ungapped 'pre' record (query segment, target segment) into the gapped
// alignment strings the consensus path consumes.
//
// Stands where reference src/cpp/SimpleAligner.cpp:10-63 stands.  The reference delegates the arithmetic to
// blasr_libcpp (SDPAlign -> GuidedAlign -> CreateAlignmentStrings, SimpleAligner.cpp:33-48), a submodule whose source
// is not part of the reference tree: PARITY UNPINNED except for the one known-answer vector of
// reference test/cpp/SimpleAlignerTest.cpp:8-21, which tests/test_simple_align_cpu.py holds this file to.  What is
// kept from the reference: the cost model it configures (SimpleAligner.cpp:10-23: match/mismatch from
// SMRTDistanceMatrix = -5/+6, insertion 4, deletion 5), a LOCAL result (the alignment starts and ends on a matching
// column, as an alignment seeded from exact k-mer anchors does), and the coordinate/strand bookkeeping of
// SimpleAligner.cpp:49-61 to the letter.
//
// Method: global DP inside a band around the scaled diagonal (row i of the query is centred on column i*m/n of the
// target), widened by doubling until the optimal path keeps clear of the band's edges; the traceback prefers the
// diagonal, so gaps in homopolymer runs land leftmost.  O(n * band) time and bytes (at most 256 MB per thread).
#include <algorithm>
#include <cstring>
#include <vector>

#include "dagcon_host.hpp"

namespace dgh {

namespace {

constexpr int32_t kMatch = -5, kMismatch = 6, kIns = 4, kDel = 5;  // SimpleAligner.cpp:10-23
constexpr int32_t kInf = 1 << 29;
constexpr int64_t G0 = 4;  // guard cells on each side of a DP row

struct Scratch {
    std::vector<uint8_t> tb;
    std::vector<int32_t> prev, cur;
    std::vector<char> q, t;  // the gapped strings, built back to front
};

inline int64_t centre(int64_t i, int64_t n, int64_t m) { return n ? (i * m) / n : 0; }

// One banded pass.  Returns false when the band cannot connect the corners or the best path touches its edge
// (the caller widens the band); on success `s.q`/`s.t` hold the gapped strings REVERSED.
bool banded_pass(const char* q, int64_t n, const char* t, int64_t m, int64_t band, Scratch& s, bool final_pass) {
    const int64_t W = 2 * band + 1;
    constexpr int64_t G = G0;  // guard cells (kept at kInf) on both sides of a row: no index checks in the inner loop
    s.tb.assign((size_t)((n + 1) * W), 0);
    s.prev.assign((size_t)(W + 2 * G), kInf);
    s.cur.assign((size_t)(W + 2 * G), kInf);
    for (int64_t k = 0; k < W; k++) {
        const int64_t j = centre(0, n, m) - band + k;
        if (j >= 0 && j <= m) { s.prev[(size_t)(k + G)] = (int32_t)(j * kDel); s.tb[(size_t)k] = 2; }
    }
    for (int64_t i = 1; i <= n; i++) {
        const int64_t c1 = centre(i, n, m), sh = c1 - centre(i - 1, n, m);
        uint8_t* tbr = s.tb.data() + (size_t)(i * W);
        const char qc = q[i - 1];
        int32_t* cur = s.cur.data() + G;
        const int32_t* prev = s.prev.data() + G;
        // columns j = c1 - band + k inside [0, m]
        const int64_t klo = std::max<int64_t>(0, band - c1), khi = std::min<int64_t>(W - 1, m - c1 + band);
        for (int64_t k = 0; k < klo && k < W; k++) { cur[k] = kInf; tbr[k] = 0; }
        for (int64_t k = std::max<int64_t>(khi + 1, 0); k < W; k++) { cur[k] = kInf; tbr[k] = 0; }
        int64_t k = klo;
        if (k <= khi && c1 - band + k == 0) {  // column 0: only (i-1, 0) leads here
            const int64_t ku = k + sh;
            const int32_t v = (ku < W + G) ? prev[ku] + kIns : kInf;
            cur[k] = v < kInf ? v : kInf;
            tbr[k] = v < kInf ? 3 : 0;
            k++;
        }
        if (sh > G - 1) return false;  // a target more than 3x the query: not an alignment this band can hold
        const int64_t tj = c1 - band - 1;  // t[j - 1] = t[tj + k]
        for (; k <= khi; k++) {
            const int32_t vd = prev[k + sh - 1] + (qc == t[tj + k] ? kMatch : kMismatch);  // (i-1, j-1)
            const int32_t vl = cur[k - 1] + kDel;                                       // (i, j-1): target base against '-'
            const int32_t vu = prev[k + sh] + kIns;                                     // (i-1, j): query base against '-'
            int32_t best = vd;
            uint8_t dir = 1;
            if (vl < best) { best = vl; dir = 2; }
            if (vu < best) { best = vu; dir = 3; }
            const bool dead = best >= kInf;
            cur[k] = dead ? kInf : best;
            tbr[k] = dead ? (uint8_t)0 : dir;
        }
        s.prev.swap(s.cur);
    }
    int64_t i = n, j = m;
    {
        const int64_t k = j - (centre(i, n, m) - band);
        if (k < 0 || k >= W || s.prev[(size_t)(k + G0)] >= kInf) return false;
    }
    s.q.clear();
    s.t.clear();
    while (i > 0 || j > 0) {
        const int64_t k = j - (centre(i, n, m) - band);
        if (k < 0 || k >= W) return false;
        if (!final_pass && (k < 2 || k > W - 3) && j > 0 && j < m) return false;  // too close to the edge: widen
        const uint8_t dir = i == 0 ? 2 : s.tb[(size_t)(i * W + k)];
        if (dir == 1) { s.q.push_back(q[--i]); s.t.push_back(t[--j]); }
        else if (dir == 2) { s.q.push_back('-'); s.t.push_back(t[--j]); }
        else if (dir == 3) { s.q.push_back(q[--i]); s.t.push_back('-'); }
        else return false;
    }
    return true;
}

}  // namespace

bool simple_align(const char* q, size_t qn, const char* t, size_t tn, std::string& qout, std::string& tout,
                  uint32_t& t_begin, uint32_t& t_end) {
    static thread_local Scratch s;
    qout.clear();
    tout.clear();
    t_begin = t_end = 0;
    if (qn == 0 || tn == 0) return false;
    const int64_t n = (int64_t)qn, m = (int64_t)tn, full = std::max(n, m) + 1;
    bool ok = false;
    // the traceback matrix is held to 256 MB per packer thread: the widest band that fits is the last one tried, and
    // on it a path along the band's edge is accepted (unrelated sequences end there instead of in a full n x m matrix)
    constexpr int64_t kMaxCells = 256ll << 20;
    for (int64_t band = 64;; band *= 2) {
        const bool last = band >= full || (n + 1) * (4 * band + 1) > kMaxCells;
        ok = banded_pass(q, n, t, m, std::min(band, full), s, last);
        if (ok || last) break;
    }
    if (!ok) return false;
    // local result: cut the columns before the first and after the last matching column
    const size_t len = s.q.size();
    size_t lo = 0, hi = len;  // in forward order; s.q/s.t are reversed
    auto qa = [&](size_t c) { return s.q[len - 1 - c]; };
    auto ta = [&](size_t c) { return s.t[len - 1 - c]; };
    uint32_t tb = 0, te = (uint32_t)tn;
    while (lo < hi && qa(lo) != ta(lo)) { tb += ta(lo) != '-'; lo++; }
    while (hi > lo && qa(hi - 1) != ta(hi - 1)) { te -= ta(hi - 1) != '-'; hi--; }
    if (lo >= hi) return false;
    qout.resize(hi - lo);
    tout.resize(hi - lo);
    for (size_t c = lo; c < hi; c++) { qout[c - lo] = qa(c); tout[c - lo] = ta(c); }
    t_begin = tb;
    t_end = te;
    return true;
}

}  // namespace dgh

// ---- reference-shaped wrapper (SimpleAligner.hpp: align(Alignment&), operator()) ----
bool SimpleAligner::align(dagcon::Alignment& aln) {
    std::string qs, ts;
    uint32_t tb = 0, te = 0;
    if (!dgh::simple_align(aln.qstr.data(), aln.qstr.size(), aln.tstr.data(), aln.tstr.size(), qs, ts, tb, te)) {
        aln.qstr.clear();
        aln.tstr.clear();
        return false;
    }
    // SimpleAligner.cpp:49-61, kept to the letter (end = the moved start + the segment-relative end)
    aln.start += tb;
    aln.end = aln.start + te;
    if (aln.strand == '-') {
        aln.start = aln.tlen - aln.end;
        std::string rq(qs.size(), '\0'), rt(ts.size(), '\0');
        dgh::revcomp_into(qs.data(), qs.size(), (uint8_t*)&rq[0]);
        dgh::revcomp_into(ts.data(), ts.size(), (uint8_t*)&rt[0]);
        aln.qstr.swap(rq);
        aln.tstr.swap(rt);
    } else {
        aln.qstr.swap(qs);
        aln.tstr.swap(ts);
    }
    aln.start++;
    return true;
}
