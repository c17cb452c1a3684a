// TEST INFRASTRUCTURE (oracle/): C-ABI wrapper around the UNMODIFIED reference objects
// (normalizeGaps/trimAln from /root/reference/src/cpp/Alignment.cpp and AlnGraphBoost from
// /root/reference/src/cpp/AlnGraphBoost.cpp), compiled where they lie against oracle/shim.
// Built into oracle/_ref/libdagcon_ref.so by oracle/Makefile.  Used only by tests/ to pin
// the C restatement (oracle/dagcon_oracle.c) and to generate golden vectors.  Never linked
// into, loaded by or shipped with the product library.
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include <map>
#include <boost/graph/adjacency_list.hpp>
#include "Alignment.hpp"
#define private public
#include "AlnGraphBoost.hpp"
#undef private

extern "C" {

// normalizeGaps (+ optional trimAln).  out buffers need 2*len+1 bytes.
// returns the new length; *start_io is updated like Alignment::start.
int64_t ref_normalize_trim(const char* q, const char* t, int64_t len, uint32_t* start_io,
                           int do_trim, int trim, char* qout, char* tout) {
    dagcon::Alignment a;
    a.qstr.assign(q, (size_t)len);
    a.tstr.assign(t, (size_t)len);
    a.start = *start_io;
    dagcon::Alignment n = normalizeGaps(a);
    if (do_trim) trimAln(n, trim);
    memcpy(qout, n.qstr.data(), n.qstr.size());
    memcpy(tout, n.tstr.data(), n.tstr.size());
    *start_io = n.start;
    return (int64_t)n.qstr.size();
}

struct RefResult {
    std::vector<CnsResult> seqs;
    std::string single;
    std::string dump;
};

static void dump_graph(AlnGraphBoost& ag, std::string& out) {
    // one line per vertex: id base weight coverage backbone deleted | out: dst:count:visited ... | in: src:count:visited ...
    G& g = ag._g;
    size_t nv = g.m_vertices.size();
    char buf[128];
    for (size_t v = 0; v < nv; v++) {
        AlnNode& n = g[v];
        snprintf(buf, sizeof buf, "%zu %c %d %d %d %d |", v, n.base, n.weight, n.coverage,
                 (int)n.backbone, (int)n.deleted);
        out += buf;
        for (size_t i = 0; i < g.m_vertices[v].out.size(); i++) {
            auto r = g.m_vertices[v].out[i];
            snprintf(buf, sizeof buf, " %zu:%d:%d", r->dst, r->prop.count, (int)r->prop.visited);
            out += buf;
        }
        out += " |";
        for (size_t i = 0; i < g.m_vertices[v].in.size(); i++) {
            auto r = g.m_vertices[v].in[i];
            snprintf(buf, sizeof buf, " %zu:%d:%d", r->src, r->prop.count, (int)r->prop.visited);
            out += buf;
        }
        out += "\n";
    }
}

// mode bits: 1 = normalizeGaps each alignment, 2 = trimAln, 4 = mergeNodes,
//            8 = use consensus(minWeight) (single longest string) instead of the vector form,
//           16 = dump graph after build (before merge), 32 = dump graph after merge
// backbone == NULL -> AlnGraphBoost(size_t tlen) ctor, else string ctor.
// Mirrors the Consensus functor, /root/reference/src/cpp/main.cpp:130-138.
void* ref_consensus(uint32_t tlen, const char* backbone, int n_aln, const char* const* q,
                    const char* const* t, const int64_t* len, const uint32_t* start, int mode,
                    int trim, int min_weight, int64_t min_len) {
    RefResult* r = new RefResult();
    AlnGraphBoost* ag = backbone ? new AlnGraphBoost(std::string(backbone, tlen))
                                 : new AlnGraphBoost((size_t)tlen);
    for (int i = 0; i < n_aln; i++) {
        dagcon::Alignment a;
        a.qstr.assign(q[i], (size_t)len[i]);
        a.tstr.assign(t[i], (size_t)len[i]);
        a.start = start[i];
        a.tlen = tlen;
        if ((int64_t)a.qstr.length() < min_len) continue;
        if (mode & 1) {
            dagcon::Alignment n = normalizeGaps(a);
            if (mode & 2) trimAln(n, trim);
            ag->addAln(n);
        } else {
            if (mode & 2) trimAln(a, trim);
            ag->addAln(a);
        }
    }
    if (mode & 16) dump_graph(*ag, r->dump);
    if (mode & 4) ag->mergeNodes();
    if (mode & 32) dump_graph(*ag, r->dump);
    if (mode & 8) r->single = ag->consensus(min_weight);
    else ag->consensus(r->seqs, min_weight, (size_t)min_len);
    delete ag;
    return r;
}

int ref_result_count(void* h) { return (int)((RefResult*)h)->seqs.size(); }
int ref_result_range(void* h, int i, int which) { return ((RefResult*)h)->seqs[i].range[which]; }
int64_t ref_result_seq(void* h, int i, const char** p) {
    RefResult* r = (RefResult*)h;
    *p = r->seqs[i].seq.data();
    return (int64_t)r->seqs[i].seq.size();
}
int64_t ref_result_single(void* h, const char** p) {
    RefResult* r = (RefResult*)h;
    *p = r->single.data();
    return (int64_t)r->single.size();
}
int64_t ref_result_dump(void* h, const char** p) {
    RefResult* r = (RefResult*)h;
    *p = r->dump.data();
    return (int64_t)r->dump.size();
}
void ref_result_free(void* h) { delete (RefResult*)h; }

// parseM5 / parsePre through the reference's operator>> on one line
// (/root/reference/src/cpp/Alignment.cpp:44-120).  which: 0 = m5, 1 = pre.
struct RefAln { dagcon::Alignment a; };
void* ref_parse_line(const char* line, int which) {
    RefAln* r = new RefAln();
    std::istringstream is(line);
    if (which == 0) parseM5(is, &r->a); else parsePre(is, &r->a);
    return r;
}
const char* ref_aln_str(void* h, int which, int64_t* n) {
    dagcon::Alignment& a = ((RefAln*)h)->a;
    const std::string& s = which == 0 ? a.id : which == 1 ? a.sid : which == 2 ? a.qstr : a.tstr;
    *n = (int64_t)s.size();
    return s.data();
}
uint32_t ref_aln_u32(void* h, int which) {
    dagcon::Alignment& a = ((RefAln*)h)->a;
    return which == 0 ? a.tlen : which == 1 ? a.start : which == 2 ? a.end : (uint32_t)a.strand;
}
void ref_aln_free(void* h) { delete (RefAln*)h; }

}  // extern "C"
