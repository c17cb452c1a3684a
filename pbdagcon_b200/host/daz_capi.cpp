// daz_capi.cpp -- extern "C" hooks over the dazcon front so that tests (ctypes) can pin the host provider against the
// reference's DazAlnProvider (built by the test infrastructure) and its TargetHit scoring (reference test/cpp/TargetHitTest.cpp).
// No compute on the consensus path in here.
#include <cstring>
#include <string>
#include <vector>

#include "dagcon_host.hpp"
#include "daz_realign.hpp"

namespace {
struct Dump { std::string text; };
}

extern "C" {

// Every (targSeq, alns) the provider yields, as text: "T id tlen nalns targSeq" then "A sid start qstr tstr" lines
// (the checker built from the reference prints the same format).
void* dgh_daz_dump(const char* db, const char* las, unsigned min_cov, unsigned max_hits, int sort_cov, int proper,
                   const int* targets, int n_targets) {
    Dump* d = new Dump();
    dgh::ProgramOpts po;
    po.minCov = min_cov; po.alnFile = las; po.seqFile = db; po.maxHits = max_hits; po.sortCov = sort_cov != 0; po.properOvls = proper != 0;
    for (int i = 0; i < n_targets; i++) po.targets.insert(targets[i]);
    try {
        DazAlnProvider ap(po);
        std::string targSeq;
        std::vector<dagcon::Alignment> alns;
        bool more = true;
        do {
            more = ap.nextTarget(targSeq, alns);
            if (!alns.empty()) {
                d->text += "T " + alns[0].id + " " + std::to_string(alns[0].tlen) + " " + std::to_string(alns.size()) + " " + targSeq + "\n";
                for (auto& a : alns) d->text += "A " + a.sid + " " + std::to_string(a.start) + " " + a.qstr + " " + a.tstr + "\n";
            }
        } while (more);
    } catch (const std::exception& e) {
        d->text = std::string("ERROR ") + e.what() + "\n";
    }
    return d;
}
// The same overlaps through the DEVICE realignment (needs a GPU): "T id tlen n" / "A sid start q t" lines.
void* dgh_daz_dump_device(const char* db, const char* las, unsigned min_cov, unsigned max_hits, int sort_cov, int proper,
                          const int* targets, int n_targets) {
    Dump* d = new Dump();
    dgh::ProgramOpts po;
    po.minCov = min_cov; po.alnFile = las; po.seqFile = db; po.maxHits = max_hits; po.sortCov = sort_cov != 0; po.properOvls = proper != 0;
    for (int i = 0; i < n_targets; i++) po.targets.insert(targets[i]);
    std::string err;
    try {
        if (dgh::daz_device_strings(po, d->text, err)) d->text = "ERROR " + err + "\n";
    } catch (const std::exception& e) {
        d->text = std::string("ERROR ") + e.what() + "\n";
    }
    return d;
}
const char* dgh_daz_text(void* h, uint64_t* n) { Dump* d = (Dump*)h; *n = d->text.size(); return d->text.data(); }
void dgh_daz_free(void* h) { delete (Dump*)h; }

// Hit::add + Hit::score on bare paths: n x {abpos, aepos, bbpos, bepos, diffs} (reference TargetHitTest.cpp:4-77).
float dgh_daz_ovl_score(const int* paths, int n) {
    dgh::daz::Hit h;
    for (int i = 0; i < n; i++) {
        dgh::daz::LasRec r;
        r.abpos = paths[5 * i]; r.aepos = paths[5 * i + 1]; r.bbpos = paths[5 * i + 2]; r.bepos = paths[5 * i + 3]; r.diffs = paths[5 * i + 4];
        h.add(std::move(r));
        h.score(false);
    }
    return h.ovlScore;
}

// Pile grouping alone (reference test/cpp/TargetTest.cpp:4-22): records as {aread, bread, flags}; returns the number of hits
// of the first pile.
int dgh_daz_group_hits(const int* recs, int n) {
    dgh::daz::Pile p;
    for (int i = 0; i < n; i++) {
        dgh::daz::LasRec r;
        r.aread = recs[3 * i]; r.bread = recs[3 * i + 1]; r.flags = (uint32_t)recs[3 * i + 2];
        if (!p.hits.empty()) {
            dgh::daz::Hit& prev = p.hits.back();
            if (prev.aread == r.aread && prev.bread == r.bread && prev.flags == r.flags) { prev.add(std::move(r)); continue; }
        }
        dgh::daz::Hit h;
        h.aread = r.aread; h.bread = r.bread; h.flags = r.flags;
        h.add(std::move(r));
        p.hits.push_back(std::move(h));
    }
    return (int)p.hits.size();
}

}  // extern "C"
