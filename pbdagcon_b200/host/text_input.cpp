// text_input.cpp -- BLASR -m5 / m4topre 'pre' line parsing and packing into dgc_batch buffers.
// Field semantics follow reference src/cpp/Alignment.cpp:44-112 (parseM5, parsePre) and the grouping rule of
// reference src/cpp/BlasrM5AlnProvider.cpp:34-55 (consecutive lines with equal target id form one target).
#include <algorithm>
#include <cerrno>
#include <cstdlib>
#include <chrono>
#include <cstring>
#include <mutex>

#include "dagcon_host.hpp"

namespace dgh {

namespace {

struct Tok {
    const char* p;
    size_t n;
};

// Fields are separated by single spaces; empty fields (runs of spaces) are skipped.  Stops after `want` fields.
size_t tokenize(const char* p, size_t n, Tok* out, size_t want) {
    size_t k = 0;
    const char* e = p + n;
    while (p < e && k < want) {
        const char* sp = (const char*)memchr(p, ' ', (size_t)(e - p));
        const char* fe = sp ? sp : e;
        if (fe > p) { out[k].p = p; out[k].n = (size_t)(fe - p); k++; }
        p = fe + 1;
    }
    return k;
}

// `std::istringstream >> uint32_t` on a field: leading digits are taken, anything after is ignored,
// no digits reads as 0.
uint32_t leading_u32(const Tok& t) {
    uint64_t v = 0;
    size_t i = 0;
    while (i < t.n && (t.p[i] == ' ' || t.p[i] == '\t')) i++;
    if (i < t.n && t.p[i] == '+') i++;
    for (; i < t.n && t.p[i] >= '0' && t.p[i] <= '9'; i++) {
        v = v * 10 + (uint64_t)(t.p[i] - '0');
        if (v > 0xFFFFFFFFull) return 0xFFFFFFFFu;
    }
    return (uint32_t)v;
}

size_t trim_cr(const char* p, size_t n) { return (n && p[n - 1] == '\r') ? n - 1 : n; }

}  // namespace

bool split_line(const char* p, size_t n, TextFormat fmt, LineFields& o) {
    n = trim_cr(p, n);
    Tok f[19];
    if (fmt == TextFormat::M5) {
        const size_t k = tokenize(p, n, f, 19);
        if (k == 0) return false;
        if (k < 19) throw M5Exception::FormatError{"Expected 19 fields, found " + std::to_string(k)};
        o.sid = f[0].p; o.sid_len = f[0].n;
        o.id = f[5].p; o.id_len = f[5].n;
        o.tlen = leading_u32(f[6]);
        o.start = leading_u32(f[7]) + 1;  // conforming offsets are 1-based
        o.end = 0;
        o.strand = f[9].p[0];
        o.q = f[16].p; o.q_len = f[16].n;
        o.t = f[18].p; o.t_len = f[18].n;
    } else {
        const size_t k = tokenize(p, n, f, 8);
        if (k == 0) return false;
        if (k < 8) throw M5Exception::FormatError{"Expected 8 fields, found " + std::to_string(k)};
        o.sid = f[0].p; o.sid_len = f[0].n;
        o.id = f[1].p; o.id_len = f[1].n;
        o.strand = f[2].p[0];
        o.tlen = leading_u32(f[3]);
        o.start = leading_u32(f[4]);
        o.end = leading_u32(f[5]);
        o.q = f[6].p; o.q_len = f[6].n;
        o.t = f[7].p; o.t_len = f[7].n;
    }
    return true;
}

bool line_target_id(const char* p, size_t n, TextFormat fmt, const char*& id, size_t& id_len) {
    n = trim_cr(p, n);
    Tok f[6];
    const size_t want = fmt == TextFormat::M5 ? 6 : 2;
    const size_t k = tokenize(p, n, f, want);
    if (k == 0) return false;
    if (k < want) throw M5Exception::FormatError{"Expected a target id in field " + std::to_string(want) + ", line has " + std::to_string(k) + " fields"};
    id = f[want - 1].p;
    id_len = f[want - 1].n;
    return true;
}

void revcomp_into(const char* src, size_t n, uint8_t* dst) {
    static const struct Table {
        uint8_t t[256];
        Table() {
            for (int i = 0; i < 256; i++) t[i] = (uint8_t)i;
            t['A'] = 'T'; t['T'] = 'A'; t['C'] = 'G'; t['G'] = 'C';  // upper case only, as the reference
        }
    } tab;
    for (size_t i = 0; i < n; i++) dst[i] = tab.t[(uint8_t)src[n - 1 - i]];
}

// ------------------------------------------------------------------------------------------ HostBatch
double g_pin_seconds = 0;
static std::mutex g_pin_mu;

void* HostBatch::alloc(size_t bytes) {
    void* p = nullptr;
    if (pinned_) {
        const auto t0 = std::chrono::steady_clock::now();
        if (dgc_host_alloc(&p, bytes) != DGC_OK) throw std::bad_alloc();
        if (bytes > (1u << 20)) {
            const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            std::lock_guard<std::mutex> lk(g_pin_mu);
            g_pin_seconds += dt;
        }
    } else {
        p = malloc(bytes ? bytes : 1);
        if (!p) throw std::bad_alloc();
    }
    return p;
}
void HostBatch::release(void* p) {
    if (!p) return;
    if (pinned_) dgc_host_free(p);
    else free(p);
}

HostBatch::HostBatch(bool pinned, size_t arena_cap, size_t aln_cap, size_t tgt_cap) : pinned_(pinned) {
    arena_cap_ = std::max<size_t>(arena_cap, 4096);
    aln_cap_ = std::max<size_t>(aln_cap, 64);
    tgt_cap_ = std::max<size_t>(tgt_cap, 16);
    arena = (uint8_t*)alloc(arena_cap_);
    aln_start = (uint32_t*)alloc(4 * aln_cap_);
    aln_len = (uint32_t*)alloc(4 * aln_cap_);
    aln_q_off = (uint64_t*)alloc(8 * aln_cap_);
    aln_t_off = (uint64_t*)alloc(8 * aln_cap_);
    tgt_tlen = (uint32_t*)alloc(4 * tgt_cap_);
    tgt_aln_begin = (uint32_t*)alloc(4 * (tgt_cap_ + 1));
    tgt_bb_off = (uint64_t*)alloc(8 * tgt_cap_);
    tgt_aln_begin[0] = 0;
}

HostBatch::~HostBatch() {
    release(arena); release(aln_start); release(aln_len); release(aln_q_off); release(aln_t_off);
    release(tgt_tlen); release(tgt_aln_begin); release(tgt_bb_off);
}

void HostBatch::clear() {
    n_targets = n_alns = 0;
    arena_bytes = 0;
    has_backbone = false;
    ids.clear();
    tgt_aln_begin[0] = 0;
}

template <class T>
static void grow(HostBatch* hb, T*& p, size_t used, size_t new_cap, void* (HostBatch::*al)(size_t), void (HostBatch::*rel)(void*)) {
    T* np = (T*)(hb->*al)(sizeof(T) * new_cap);
    memcpy(np, p, sizeof(T) * used);
    (hb->*rel)(p);
    p = np;
}

void HostBatch::need_arena(size_t extra) {
    if (arena_bytes + extra <= arena_cap_) return;
    // an empty batch is sized to what its unit needs (page-locking is paid per byte); a filling one doubles
    size_t nc = arena_bytes ? std::max(arena_cap_ * 2, (size_t)arena_bytes + extra) : extra + (extra >> 4);
    uint8_t* np = (uint8_t*)alloc(nc);
    memcpy(np, arena, arena_bytes);
    release(arena);
    arena = np;
    arena_cap_ = nc;
}
void HostBatch::reserve(size_t arena_need, size_t alns, size_t targets) {
    need_arena(arena_need);
    need_alns(alns);
    need_targets(targets);
}
void HostBatch::need_alns(size_t extra) {
    if (n_alns + extra <= aln_cap_) return;
    const size_t nc = std::max(aln_cap_ * 2, (size_t)n_alns + extra);
    grow(this, aln_start, n_alns, nc, &HostBatch::alloc, &HostBatch::release);
    grow(this, aln_len, n_alns, nc, &HostBatch::alloc, &HostBatch::release);
    grow(this, aln_q_off, n_alns, nc, &HostBatch::alloc, &HostBatch::release);
    grow(this, aln_t_off, n_alns, nc, &HostBatch::alloc, &HostBatch::release);
    aln_cap_ = nc;
}
void HostBatch::need_targets(size_t extra) {
    if (n_targets + extra <= tgt_cap_) return;
    const size_t nc = std::max(tgt_cap_ * 2, (size_t)n_targets + extra);
    grow(this, tgt_tlen, n_targets, nc, &HostBatch::alloc, &HostBatch::release);
    grow(this, tgt_bb_off, n_targets, nc, &HostBatch::alloc, &HostBatch::release);
    uint32_t* nb = (uint32_t*)alloc(4 * (nc + 1));
    memcpy(nb, tgt_aln_begin, 4 * ((size_t)n_targets + 1));
    release(tgt_aln_begin);
    tgt_aln_begin = nb;
    tgt_cap_ = nc;
}

static inline uint64_t align16(uint64_t x) { return (x + 15) & ~15ull; }

// ---- 4-bit wire codes (include/dagcon_b200.h: DGC_PACK4_ALPHABET "-ACGTN.")
namespace {
struct Pack4Tables {
    uint8_t code[256], rc[256];   // code of a character / of its complement as revComp forms it (Alignment.cpp:15-26)
    Pack4Tables() {
        memset(code, 0xFF, sizeof code);
        const char* al = DGC_PACK4_ALPHABET;
        for (int i = 0; al[i]; i++) code[(uint8_t)al[i]] = (uint8_t)i;
        memcpy(rc, code, sizeof rc);
        rc[(uint8_t)'A'] = code[(uint8_t)'T']; rc[(uint8_t)'T'] = code[(uint8_t)'A'];
        rc[(uint8_t)'C'] = code[(uint8_t)'G']; rc[(uint8_t)'G'] = code[(uint8_t)'C'];
    }
};
const Pack4Tables kP4;

// columns src[0..n) -> n/2 (rounded up) bytes at dst, even columns in the low nibble; false on a foreign character
inline bool pack4_into(const char* src, size_t n, uint8_t* dst) {
    uint8_t bad = 0;
    size_t c = 0;
    for (; c + 1 < n; c += 2) {
        const uint8_t a = kP4.code[(uint8_t)src[c]], b = kP4.code[(uint8_t)src[c + 1]];
        bad |= a | b;
        dst[c >> 1] = (uint8_t)(a | (b << 4));
    }
    if (c < n) { const uint8_t a = kP4.code[(uint8_t)src[c]]; bad |= a; dst[c >> 1] = a; }
    return !(bad & 0x80);
}
// the same for the reverse complement of src (revcomp_into followed by pack4_into, in one pass)
inline bool pack4_revcomp_into(const char* src, size_t n, uint8_t* dst) {
    uint8_t bad = 0;
    size_t c = 0;
    for (; c + 1 < n; c += 2) {
        const uint8_t a = kP4.rc[(uint8_t)src[n - 1 - c]], b = kP4.rc[(uint8_t)src[n - 2 - c]];
        bad |= a | b;
        dst[c >> 1] = (uint8_t)(a | (b << 4));
    }
    if (c < n) { const uint8_t a = kP4.rc[(uint8_t)src[n - 1 - c]]; bad |= a; dst[c >> 1] = a; }
    return !(bad & 0x80);
}
}  // namespace

// One 'pre' line through the -a step: ungapped segments in, gapped strings in target orientation out
// (reference main.cpp:127-129 -> SimpleAligner.cpp:25-63).
bool align_pre_line(const char* p, size_t n, PreAligned& out) {
    LineFields f;
    out.ok = false;
    if (!split_line(p, n, TextFormat::PRE, f)) return false;
    uint32_t tb = 0, te = 0;
    if (!simple_align(f.q, f.q_len, f.t, f.t_len, out.q, out.t, tb, te)) return true;  // parsed, nothing aligns
    // SimpleAligner.cpp:49-61
    f.start += tb;
    f.end = f.start + te;
    if (f.strand == '-') {
        f.start = f.tlen - f.end;
        std::string rq(out.q.size(), '\0'), rt(out.t.size(), '\0');
        revcomp_into(out.q.data(), out.q.size(), (uint8_t*)&rq[0]);
        revcomp_into(out.t.data(), out.t.size(), (uint8_t*)&rt[0]);
        out.q.swap(rq);
        out.t.swap(rt);
    }
    out.start = f.start + 1;
    out.ok = true;
    return true;
}

void HostBatch::add_target_lines(const std::vector<std::pair<const char*, size_t>>& lines, TextFormat fmt, const PreAligned* pre) {
    need_targets(1);
    need_alns(lines.size());
    bool first = true;
    PreAligned local;
    size_t li = 0;
    for (const auto& ln : lines) {
        const size_t idx = li++;
        LineFields f;
        if (!split_line(ln.first, ln.second, fmt, f)) continue;
        if (fmt == TextFormat::PRE) {  // 'pre' records are ungapped segments: aligned first (by the caller, or here)
            const PreAligned* a = pre ? &pre[idx] : &local;
            if (!pre) align_pre_line(ln.first, ln.second, local);
            if (a->ok) {
                f.start = a->start;
                f.q = a->q.data(); f.q_len = a->q.size();
                f.t = a->t.data(); f.t_len = a->t.size();
            } else {
                // Nothing aligns: the record still counts towards the target's alignment count (the reference tests
                // alns.size() before it aligns, main.cpp:66-72,118) and is then dropped by the raw minLen filter
                // (main.cpp:132): packed as a zero-column alignment.
                f.start += 1;
                f.q_len = f.t_len = 0;
            }
        }
        if (f.q_len != f.t_len) throw M5Exception::FormatError{"query and target alignment strings differ in length"};
        if (f.q_len > 0x7FFFFFFFull) throw M5Exception::FormatError{"alignment too long"};
        if (first) {
            tgt_tlen[n_targets] = f.tlen;
            ids.emplace_back(f.id, f.id_len);
            first = false;
        }
        const size_t n = f.q_len, nb = packed4 ? (n + 1) / 2 : n;
        need_arena(2 * (nb + 16));
        const uint64_t qo = align16(arena_bytes), to = align16(qo + nb);
        const bool rc = fmt == TextFormat::M5 && f.strand == '-';  // Alignment.cpp:72-76
        if (packed4) {
            const bool ok = rc ? (pack4_revcomp_into(f.q, n, arena + qo) && pack4_revcomp_into(f.t, n, arena + to))
                               : (pack4_into(f.q, n, arena + qo) && pack4_into(f.t, n, arena + to));
            if (!ok) throw NeedAscii{};
        } else if (rc) {
            revcomp_into(f.q, n, arena + qo);
            revcomp_into(f.t, n, arena + to);
        } else {
            memcpy(arena + qo, f.q, n);
            memcpy(arena + to, f.t, n);
        }
        arena_bytes = to + nb;
        aln_start[n_alns] = f.start;
        aln_len[n_alns] = (uint32_t)n;
        aln_q_off[n_alns] = qo;
        aln_t_off[n_alns] = to;
        n_alns++;
    }
    if (first) return;  // nothing but blank lines
    tgt_bb_off[n_targets] = 0;
    n_targets++;
    tgt_aln_begin[n_targets] = n_alns;
}

void HostBatch::add_target(const std::vector<dagcon::Alignment>& alns, const std::string* backbone) {
    if (alns.empty()) return;
    if (n_targets && has_backbone != (backbone != nullptr))
        throw PacBio::DagCon::IOException("a batch cannot mix targets with and without a backbone sequence");
    need_targets(1);
    need_alns(alns.size());
    tgt_tlen[n_targets] = alns[0].tlen;
    ids.push_back(alns[0].id);
    has_backbone = backbone != nullptr;
    if (backbone) {
        if (backbone->size() != alns[0].tlen) throw PacBio::DagCon::IOException("backbone length differs from tlen");
        need_arena(backbone->size() + 16);
        const uint64_t bo = align16(arena_bytes);
        memcpy(arena + bo, backbone->data(), backbone->size());
        arena_bytes = bo + backbone->size();
        tgt_bb_off[n_targets] = bo;
    } else {
        tgt_bb_off[n_targets] = 0;
    }
    for (const auto& a : alns) {
        if (a.qstr.size() != a.tstr.size()) throw M5Exception::FormatError{"query and target alignment strings differ in length"};
        const size_t n = a.qstr.size(), nb = packed4 ? (n + 1) / 2 : n;
        need_arena(2 * (nb + 16));
        const uint64_t qo = align16(arena_bytes), to = align16(qo + nb);
        if (packed4) {
            if (!pack4_into(a.qstr.data(), n, arena + qo) || !pack4_into(a.tstr.data(), n, arena + to)) throw NeedAscii{};
        } else {
            memcpy(arena + qo, a.qstr.data(), n);
            memcpy(arena + to, a.tstr.data(), n);
        }
        arena_bytes = to + nb;
        aln_start[n_alns] = a.start;
        aln_len[n_alns] = (uint32_t)n;
        aln_q_off[n_alns] = qo;
        aln_t_off[n_alns] = to;
        n_alns++;
    }
    n_targets++;
    tgt_aln_begin[n_targets] = n_alns;
}

dgc_batch HostBatch::view() const {
    dgc_batch b;
    b.n_targets = n_targets;
    b.n_alns = n_alns;
    b.tgt_tlen = tgt_tlen;
    b.tgt_aln_begin = tgt_aln_begin;
    b.tgt_bb_off = has_backbone ? tgt_bb_off : nullptr;
    b.aln_start = aln_start;
    b.aln_len = aln_len;
    b.aln_q_off = aln_q_off;
    b.aln_t_off = aln_t_off;
    b.arena = arena;
    b.arena_bytes = arena_bytes;
    b.flags = packed4 ? DGC_BATCH_PACKED4 : 0u;
    b.reserved = 0;
    return b;
}

dgc_pre_batch PrePayload::view() const {
    dgc_pre_batch b{};
    b.n_targets = (uint32_t)tgt_tlen.size();
    b.n_recs = (uint32_t)rec_start.size();
    b.tgt_tlen = tgt_tlen.data(); b.tgt_rec_begin = tgt_rec_begin.data();
    b.rec_start = rec_start.data(); b.rec_strand = rec_strand.data(); b.rec_qlen = rec_qlen.data(); b.rec_tseglen = rec_tseglen.data();
    b.rec_q_off = rec_q_off.data(); b.rec_t_off = rec_t_off.data();
    b.arena = arena.data(); b.arena_bytes = arena.size();
    return b;
}

dgc_daz_batch DazPayload::view() const {
    dgc_daz_batch b{};
    b.n_targets = (uint32_t)tgt_aread.size();
    b.n_ovls = (uint32_t)ovl_bread.size();
    b.tspace = tspace;
    b.tgt_aread = tgt_aread.data(); b.tgt_ovl_begin = tgt_ovl_begin.data();
    b.ovl_bread = ovl_bread.data(); b.ovl_comp = ovl_comp.data();
    b.ovl_abpos = ovl_abpos.data(); b.ovl_aepos = ovl_aepos.data(); b.ovl_bbpos = ovl_bbpos.data(); b.ovl_bepos = ovl_bepos.data();
    b.ovl_trace_off = ovl_trace_off.data(); b.ovl_tlen = ovl_tlen.data();
    b.trace = trace.data(); b.trace_bytes = trace.size();
    return b;
}

}  // namespace dgh

// ------------------------------------------------------------------------------------------ provider
BlasrM5AlnProvider::BlasrM5AlnProvider(const std::string& fpath, dgh::TextFormat fmt) : fmt_(fmt) {
    fp_ = fopen(fpath.c_str(), "rb");
    if (!fp_) throw M5Exception::FileOpenError();
    owns_ = true;
}
BlasrM5AlnProvider::BlasrM5AlnProvider(FILE* stream, dgh::TextFormat fmt) : fp_(stream), owns_(false), fmt_(fmt) {}
BlasrM5AlnProvider::~BlasrM5AlnProvider() {
    if (owns_ && fp_) fclose(fp_);
    free(line_);
}

// false at end of input; `blank` marks a line without fields
bool BlasrM5AlnProvider::read_line(dagcon::Alignment& aln, bool& blank) {
    const ssize_t n = getline(&line_, &line_cap_, fp_);
    if (n < 0) return false;
    size_t len = (size_t)n;
    if (len && line_[len - 1] == '\n') len--;
    dgh::LineFields f;
    blank = !dgh::split_line(line_, len, fmt_, f);
    if (blank) return true;
    aln.tlen = f.tlen; aln.start = f.start; aln.end = f.end; aln.strand = f.strand;
    aln.id.assign(f.id, f.id_len);
    aln.sid.assign(f.sid, f.sid_len);
    if (fmt_ == dgh::TextFormat::M5 && f.strand == '-') {
        aln.qstr.resize(f.q_len); aln.tstr.resize(f.t_len);
        dgh::revcomp_into(f.q, f.q_len, (uint8_t*)&aln.qstr[0]);
        dgh::revcomp_into(f.t, f.t_len, (uint8_t*)&aln.tstr[0]);
    } else {
        aln.qstr.assign(f.q, f.q_len);
        aln.tstr.assign(f.t, f.t_len);
    }
    return true;
}

bool BlasrM5AlnProvider::nextTarget(std::vector<dagcon::Alignment>& dest) {
    dest.clear();
    if (!firstAln_) dest.push_back(prevAln_);
    dagcon::Alignment aln;
    bool blank = false;
    while (read_line(aln, blank)) {
        // The reference mishandles blank lines (it re-pushes the previous record or splits the group,
        // SURVEY App. C.1); a trailing blank line is harmless there and is skipped here, one in the
        // middle of the stream is a format error.
        if (blank) {
            int c;
            while ((c = fgetc(fp_)) == '\n' || c == '\r') {}
            if (c == EOF) return false;
            ungetc(c, fp_);
            throw M5Exception::FormatError{"blank line inside the alignment stream"};
        }
        if (aln.id != currId_) {
            firstAln_ = false;
            prevAln_ = aln;
            currId_ = aln.id;
            return true;
        }
        dest.push_back(aln);
    }
    return false;
}

bool BlasrM5AlnProvider::nextTarget(std::string&, std::vector<dagcon::Alignment>&) { return false; }  // as the reference: NOOP
