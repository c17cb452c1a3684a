// TEST INFRASTRUCTURE (oracle/): C-ABI wrapper around the UNMODIFIED reference DazAlnProvider
// (/root/reference/src/cpp/DazAlnProvider.cpp) and its TargetHit scoring, compiled where they lie and linked
// with the vendored DALIGNER/DAZZ_DB C sources.  Built into oracle/_ref/libdaz_ref.so by oracle/Makefile.
// Used only by tests/ to pin the product's host-side dazcon provider (pbdagcon_b200/host/daz_provider.cpp).
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include "DazAlnProvider.hpp"

namespace {
struct Dump {
    std::string text;  // one line per target: "T id tlen nalns targSeq", then per alignment "A sid start qstr tstr"
};
}  // namespace

extern "C" {

// Runs the reference provider over a whole DB + LAS pair exactly as dazcon's Reader does
// (/root/reference/src/cpp/dazcon.cpp:38-48) and returns every (targSeq, alns) it would push.
void* dazref_dump(const char* db, const char* las, unsigned min_cov, unsigned max_hits, int sort_cov, int proper,
                  const int* targets, int n_targets) {
    Dump* d = new Dump();
    ProgramOpts po;
    po.minCov = min_cov; po.minLen = 500; po.trim = 10; po.threads = 1; po.align = false;
    po.alnFile = las; po.seqFile = db; po.maxHits = max_hits; po.sortCov = sort_cov != 0; po.properOvls = proper != 0;
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
                for (auto& a : alns)
                    d->text += "A " + a.sid + " " + std::to_string(a.start) + " " + a.qstr + " " + a.tstr + "\n";
            }
        } while (more);
    } catch (PacBio::DagCon::IOException& e) {
        d->text = std::string("ERROR ") + e.what() + "\n";
    }
    return d;
}
const char* dazref_text(void* h, uint64_t* n) { Dump* d = (Dump*)h; *n = d->text.size(); return d->text.data(); }
void dazref_free(void* h) { delete (Dump*)h; }

// TargetHit::add + computeOvlScore on bare paths (reference test/cpp/TargetHitTest.cpp:4-77 drives it this way).
// paths: n x {abpos, aepos, bbpos, bepos, diffs}; returns ovlScore after adding all of them.
float dazref_ovl_score(const int* paths, int n) {
    TargetHit th;
    for (int i = 0; i < n; i++) {
        Record rec;
        Overlap ovl;
        memset(&ovl, 0, sizeof ovl);
        rec.ovl = ovl;
        Path p;
        memset(&p, 0, sizeof p);
        p.abpos = paths[5 * i]; p.aepos = paths[5 * i + 1]; p.bbpos = paths[5 * i + 2]; p.bepos = paths[5 * i + 3]; p.diffs = paths[5 * i + 4];
        rec.ovl.path = p;
        th.add(rec);
        th.computeOvlScore();
    }
    return th.ovlScore;
}

}  // extern "C"
