// host_capi.cpp -- extern "C" hooks over the host-side reader/packer so that tests (ctypes) can check the
// text boundary without a GPU: the line parser against reference Alignment.cpp:44-112 and the scanner/packer
// against BlasrM5AlnProvider grouping (reference src/cpp/BlasrM5AlnProvider.cpp:34-55).  No compute in here.
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "dagcon_host.hpp"

namespace {

struct Packed {  // every batch the scanner produced, concatenated
    std::vector<uint32_t> tgt_tlen, tgt_aln_begin{0}, aln_start, aln_len, batch_targets;
    std::vector<uint64_t> aln_q_off, aln_t_off, tgt_bb_off;
    std::vector<uint8_t> arena;
    std::string ids;  // '\n'-separated
    dgh::RunStats st;
    std::string err;
};

void sink(const dgh::HostBatch& b, void* user) {
    Packed& p = *(Packed*)user;
    p.arena.resize((p.arena.size() + 15) & ~(size_t)15);  // keep every string 16-byte aligned across batches
    const uint64_t base = p.arena.size();
    const uint32_t a0 = (uint32_t)p.aln_len.size();
    p.arena.insert(p.arena.end(), b.arena, b.arena + b.arena_bytes);
    for (uint32_t k = 0; k < b.n_alns; k++) {
        p.aln_start.push_back(b.aln_start[k]);
        p.aln_len.push_back(b.aln_len[k]);
        p.aln_q_off.push_back(base + b.aln_q_off[k]);
        p.aln_t_off.push_back(base + b.aln_t_off[k]);
    }
    for (uint32_t t = 0; t < b.n_targets; t++) {
        p.tgt_tlen.push_back(b.tgt_tlen[t]);
        if (b.has_backbone) p.tgt_bb_off.push_back(base + b.tgt_bb_off[t]);
        p.tgt_aln_begin.push_back(a0 + b.tgt_aln_begin[t + 1]);
        p.ids += b.ids[t];
        p.ids.push_back('\n');
    }
    p.batch_targets.push_back(b.n_targets);
}

}  // namespace

extern "C" {

// Scan + pack a whole text file (fmt 0 = m5, 1 = pre).  Returns a handle or NULL (error text via dgh_pack_error).
void* dgh_pack_file(const char* path, int fmt, uint32_t min_cov, int threads, uint32_t batch_targets, uint64_t batch_bytes) {
    Packed* p = new Packed();
    dgh::ProgramOpts o;
    o.input = path; o.minCov = min_cov; o.threads = threads; o.align = fmt == 1;
    o.asciiWire = true;   // the tests read the packed strings back as text
    if (batch_targets) o.batchTargets = batch_targets;
    if (batch_bytes) o.batchBytes = batch_bytes;
    const int rc = dgh::pack_text(o, false, sink, p, p->st, p->err);
    if (rc && p->err.empty()) p->err = "pack_text failed";
    return p;
}
#ifdef DGH_HAVE_DAZ
// The dazcon front over a DB + LAS pair: every target with its backbone and the alignments rebuilt from the trace
// points, packed as the C ABI takes them (same handle type as dgh_pack_file; array 9 = tgt_bb_off).
void* dgh_pack_daz(const char* db, const char* las, uint32_t min_cov, uint32_t max_hits, int sort_cov, int proper, int threads,
                   uint32_t batch_targets, double* seconds, const int* targets, int n_targets) {
    Packed* p = new Packed();
    dgh::ProgramOpts o;
    o.seqFile = db; o.alnFile = las; o.minCov = min_cov; o.maxHits = max_hits; o.sortCov = sort_cov != 0; o.properOvls = proper != 0;
    o.threads = threads;
    for (int i = 0; i < n_targets; i++) o.targets.insert(targets[i]);   // 1-based ids, as dazcon's positional arguments
    if (batch_targets) o.batchTargets = batch_targets;
    const int rc = dgh::pack_daz(o, false, sink, p, p->st, p->err);
    if (rc && p->err.empty()) p->err = "pack_daz failed";
    if (seconds) *seconds = p->st.seconds;
    return p;
}
#endif
const char* dgh_pack_error(void* h) { return ((Packed*)h)->err.c_str(); }
uint32_t dgh_pack_counts(void* h, int which) {
    Packed* p = (Packed*)h;
    switch (which) {
        case 0: return (uint32_t)p->tgt_tlen.size();
        case 1: return (uint32_t)p->aln_len.size();
        case 2: return (uint32_t)p->batch_targets.size();
        case 3: return (uint32_t)p->st.targets_dropped;
        case 4: return (uint32_t)p->st.targets_in;
        default: return 0;
    }
}
uint64_t dgh_pack_arena_bytes(void* h) { return ((Packed*)h)->arena.size(); }
const void* dgh_pack_array(void* h, int which) {
    Packed* p = (Packed*)h;
    switch (which) {
        case 0: return p->tgt_tlen.data();
        case 1: return p->tgt_aln_begin.data();
        case 2: return p->aln_start.data();
        case 3: return p->aln_len.data();
        case 4: return p->aln_q_off.data();
        case 5: return p->aln_t_off.data();
        case 6: return p->arena.data();
        case 7: return p->ids.c_str();
        case 8: return p->batch_targets.data();
        case 9: return p->tgt_bb_off.empty() ? nullptr : p->tgt_bb_off.data();
        default: return nullptr;
    }
}
void dgh_pack_free(void* h) { delete (Packed*)h; }

// One line through the parser: returns 1 (parsed), 0 (blank) or -1 (format error).  Strings are malloc'd
// copies after strand handling (the caller frees them with dgh_free).
int dgh_parse_line(const char* line, uint64_t n, int fmt, uint32_t* tlen, uint32_t* start, uint32_t* end, char* strand,
                   char** id, char** sid, char** q, char** t) {
    dgh::LineFields f;
    try {
        if (!dgh::split_line(line, n, fmt == 1 ? dgh::TextFormat::PRE : dgh::TextFormat::M5, f)) return 0;
    } catch (const M5Exception::FormatError&) {
        return -1;
    }
    *tlen = f.tlen; *start = f.start; *end = f.end; *strand = f.strand;
    auto dup = [](const char* p, size_t k) { char* r = (char*)malloc(k + 1); memcpy(r, p, k); r[k] = 0; return r; };
    *id = dup(f.id, f.id_len);
    *sid = dup(f.sid, f.sid_len);
    if (fmt == 0 && f.strand == '-') {
        *q = (char*)malloc(f.q_len + 1); *t = (char*)malloc(f.t_len + 1);
        dgh::revcomp_into(f.q, f.q_len, (uint8_t*)*q); (*q)[f.q_len] = 0;
        dgh::revcomp_into(f.t, f.t_len, (uint8_t*)*t); (*t)[f.t_len] = 0;
    } else {
        *q = dup(f.q, f.q_len);
        *t = dup(f.t, f.t_len);
    }
    return 1;
}
void dgh_free(void* p) { free(p); }

// SimpleAligner::align on one 'pre' record (reference test/cpp/SimpleAlignerTest.cpp:8-21 drives it this way).
// Returns 1 and malloc'd gapped strings (dgh_free), or 0 when nothing aligns.
int dgh_simple_align(const char* q, const char* t, char strand, uint32_t tlen, uint32_t* start, uint32_t* end, char** qout,
                     char** tout) {
    dagcon::Alignment a;
    a.tlen = tlen; a.start = *start; a.end = *end; a.strand = strand; a.qstr = q; a.tstr = t;
    SimpleAligner sa;
    if (!sa.align(a)) return 0;
    *start = a.start; *end = a.end;
    *qout = strdup(a.qstr.c_str());
    *tout = strdup(a.tstr.c_str());
    return 1;
}

// The reference-shaped provider: number of alignments per nextTarget() call, in call order (first call yields 0).
// Returns the number of calls made (<= cap) or -1 on error.
int dgh_provider_groups(const char* path, int fmt, uint32_t* sizes, int cap) {
    try {
        BlasrM5AlnProvider ap(path, fmt == 1 ? dgh::TextFormat::PRE : dgh::TextFormat::M5);
        std::vector<dagcon::Alignment> alns;
        int n = 0;
        bool more = true;
        while (more && n < cap) {
            more = ap.nextTarget(alns);
            sizes[n++] = (uint32_t)alns.size();
        }
        return n;
    } catch (...) {
        return -1;
    }
}

}  // extern "C"
