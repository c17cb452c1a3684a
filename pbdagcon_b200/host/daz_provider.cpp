// daz_provider.cpp -- the dazcon front: daligner LAS + DAZZ_DB in, packed consensus targets out.
//
// Behaviour follows the reference's DazAlnProvider (reference src/cpp/DazAlnProvider.cpp):
//   :23-70    open the DB (Open_DB, Trim_DB) and the LAS header (novl, tspace -> 1- or 2-byte trace values)
//   :79-132   targets = maximal runs of records with the same A read; provider-level filters (target list, minCov)
//   :171-211  TargetHit: the records of one (aread, bread, orientation) chain; add() keeps non-overlapping records
//             or the longer of two overlapping ones; computeOvlScore()
//   :244-302  addRecord (chain only with the PREVIOUS hit), sortHits (ovlScore, optionally coverage re-scoring)
//   :304-369  getAlignments: first maxHits hits, every record rebuilt with Compute_Trace_PTS(GREEDIEST)
//   :383-417  decodeAlignment: trace -> gapped strings (one column past aepos included)
// The structure is not the reference's: the LAS file is mapped and cut into piles by one scanner; hit selection
// happens on the pile before any sequence is touched; the selected overlaps are rebuilt by a pool of workers (one
// dazzler Work_Data and DB handle each) straight into pinned dgc_batch buffers; the reference does all of this in its
// single Reader thread, one std::string append per column.  The dazzler C library (DALIGNER/align.c, DAZZ_DB/DB.c,
// QV.c) is linked unmodified from the path given at build time (DAZ_SRC in the Makefile).
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <chrono>
#include <cstring>
#include <mutex>
#include <numeric>
#include <thread>

#include "dagcon_host.hpp"
#include "channels.hpp"
#include "daz_realign.hpp"

extern "C" {
#include "DB.h"
#include "align.h"
}

namespace dgh {
namespace daz {

static const char kToU[8] = {'A', 'C', 'G', 'T', '.', '[', ']', '-'};   // DazAlnProvider.cpp:18
static const int kBorder = 10;                                            // DazAlnProvider.cpp:19

// ------------------------------------------------------------------------------------------ LAS file
LasFile::LasFile(const std::string& path) {
    fd_ = open(path.c_str(), O_RDONLY);
    if (fd_ < 0) throw PacBio::DagCon::IOException("Open failed: " + path);
    struct stat sb;
    if (fstat(fd_, &sb) != 0 || sb.st_size < 12) { close(fd_); throw PacBio::DagCon::IOException("Failed to read novl"); }
    len_ = (size_t)sb.st_size;
    map_ = (const uint8_t*)mmap(nullptr, len_, PROT_READ, MAP_PRIVATE, fd_, 0);
    if (map_ == MAP_FAILED) { close(fd_); throw PacBio::DagCon::IOException("Open failed: " + path); }
    madvise((void*)map_, len_, MADV_SEQUENTIAL);
    memcpy(&novl, map_, 8);
    memcpy(&tspace, map_ + 8, 4);
    tbytes = tspace <= TRACE_XOVR ? 1 : 2;   // DazAlnProvider.cpp:56-62
    pos_ = 12;
}
LasFile::~LasFile() {
    if (map_ && map_ != MAP_FAILED) munmap((void*)map_, len_);
    if (fd_ >= 0) close(fd_);
}
// One record: the 40 bytes behind the path pointer of dazzler's Overlap (align.c:3005-3026), then tlen trace values.
bool LasFile::next(LasRec& r) {
    if (read_ >= novl || pos_ + 40 > len_) return false;
    int32_t f[10];
    memcpy(f, map_ + pos_, 40);
    pos_ += 40;
    r.tlen = f[0]; r.diffs = f[1]; r.abpos = f[2]; r.bbpos = f[3]; r.aepos = f[4]; r.bepos = f[5];
    r.flags = (uint32_t)f[6]; r.aread = f[7]; r.bread = f[8];
    const size_t nb = (size_t)(r.tlen > 0 ? r.tlen : 0) * (size_t)tbytes;
    if (pos_ + nb > len_) return false;
    r.trace.resize((size_t)(r.tlen > 0 ? r.tlen : 0));
    if (tbytes == 1) for (int i = 0; i < r.tlen; i++) r.trace[(size_t)i] = map_[pos_ + (size_t)i];   // Decompress_TraceTo16
    else if (nb) memcpy(r.trace.data(), map_ + pos_, nb);
    pos_ += nb;
    read_++;
    return true;
}

// ------------------------------------------------------------------------------------------ hits
// TargetHit::add (DazAlnProvider.cpp:171-189): a record joins the chain when it starts behind the last one's end;
// when it overlaps the last one, the longer of the two stays.
void Hit::add(LasRec&& rec) {
    if (records.empty()) { records.push_back(std::move(rec)); return; }
    const LasRec& prev = records.back();
    const int prevLen = prev.aepos - prev.abpos, currLen = rec.aepos - rec.abpos;
    if (rec.abpos > prev.aepos) records.push_back(std::move(rec));
    else if (currLen > prevLen) records.back() = std::move(rec);
}
// TargetHit::computeOvlScore (:190-211).  The running lengths enter the difference term cumulatively, record by
// record, exactly as the reference accumulates them; all arithmetic in float as there.
void Hit::score(bool proper) {
    int ahlen = 0, bhlen = 0, diff = 0;
    for (const LasRec& r : records) {
        ahlen += r.aepos - r.abpos;
        bhlen += r.bepos - r.bbpos;
        diff += std::abs(ahlen - bhlen) + r.diffs;
    }
    ovlScore = (1 - diff / (float)ahlen) * ahlen;
    if (proper) {
        const LasRec& f = records.front();
        const LasRec& b = records.back();
        if (f.abpos != 0 && b.bbpos != 0) ovlScore = 0.0f;
        if (f.aepos != alen && b.bepos != blen) ovlScore = 0.0f;
    }
}

static bool by_ovl_score(const Hit& l, const Hit& r) { return l.ovlScore > r.ovlScore; }
static bool by_cov_score(const Hit& l, const Hit& r) { return l.covScore > r.covScore; }
static float inverted_sum(float x, unsigned int y) { return x + 1 / (float)y; }

// Target::sortHits (:285-302).  std::sort is not stable: ties keep the order libstdc++'s introsort leaves them in,
// which depends only on the sequence of comparisons -- same library, same comparator, same input order here.
void Pile::sort_hits(bool sort_cov) {
    std::sort(hits.begin(), hits.end(), by_ovl_score);
    if (!sort_cov) return;
    std::vector<unsigned int> coverage((size_t)std::max(length, 0), 0u);
    for (Hit& h : hits)
        for (const LasRec& r : h.records) {
            auto beg = coverage.begin() + r.abpos, end = coverage.begin() + r.aepos;
            for (auto it = beg; it != end; ++it) ++*it;
            h.covScore = std::accumulate(beg, end, 0.0, inverted_sum);   // the last record of a hit decides (:294-297)
        }
    std::sort(hits.begin(), hits.end(), by_cov_score);
}

// ------------------------------------------------------------------------------------------ piles
PileScanner::PileScanner(const std::string& las, const ReadLengths& rl, bool proper) : las_(las), rl_(rl), proper_(proper) {
    have_pending_ = las_.next(pending_);
}

Hit PileScanner::new_hit(LasRec&& rec, int alen) {
    Hit h;
    h.aread = rec.aread; h.bread = rec.bread; h.flags = rec.flags;
    h.alen = alen; h.blen = rl_.rlen(rec.bread);
    h.add(std::move(rec));
    h.score(proper_);
    return h;
}

// One pile = the maximal run of records with the same A read (DazAlnProvider.cpp:79-117: a target ends when
// rec.ovl.aread changes or the file does).  Target::firstRecord / addRecord (:225-283): a record extends the
// PREVIOUS hit when A read, B read and flags agree, else it opens a new hit.
bool PileScanner::next(Pile& p) {
    if (!have_pending_) return false;
    p.hits.clear();
    p.aread = pending_.aread;
    p.length = rl_.rlen(p.aread);
    p.hits.push_back(new_hit(std::move(pending_), p.length));
    LasRec rec;
    for (;;) {
        have_pending_ = las_.next(rec);
        if (!have_pending_) break;
        if (rec.aread != p.aread) { pending_ = std::move(rec); break; }
        Hit& prev = p.hits.back();
        if (prev.aread == rec.aread && prev.bread == rec.bread && prev.flags == rec.flags) {
            prev.add(std::move(rec));
            prev.score(proper_);
        } else {
            p.hits.push_back(new_hit(std::move(rec), p.length));
        }
    }
    return true;
}

// Target::getAlignments' selection (:304-309): the records of the first `max_hits` hits after sorting.
size_t Pile::select(unsigned max_hits, bool sort_cov, std::vector<const LasRec*>& recs, std::vector<int>& blens) {
    sort_hits(sort_cov);
    recs.clear(); blens.clear();
    const size_t nh = hits.size() > max_hits ? max_hits : hits.size();
    for (size_t i = 0; i < nh; i++)
        for (const LasRec& r : hits[i].records) { recs.push_back(&r); blens.push_back(hits[i].blen); }
    return recs.size();
}

// ------------------------------------------------------------------------------------------ read store
struct DazDb::Impl {
    HITS_DB db;
    bool open = false;
};
// The dazzler library builds file names in static buffers (DAZZ_DB/DB.c: Catenate, PathTo): everything that opens a
// file -- Open_DB, and the first Load_Subread of a handle, which opens the .bps -- runs under this lock.  Reads
// after that are per handle and need no lock.
static std::mutex g_daz_open_mu;

DazDb::DazDb(const std::string& path) : impl_(new Impl) {
    std::vector<char> p(path.begin(), path.end());
    p.push_back('\0');                       // Open_DB wants a mutable string
    std::lock_guard<std::mutex> lk(g_daz_open_mu);
    if (Open_DB(p.data(), &impl_->db) < 0) throw PacBio::DagCon::IOException("Failed to open DB");
    impl_->open = true;
    Trim_DB(&impl_->db);
}
DazDb::~DazDb() { if (impl_->open) Close_DB(&impl_->db); }
int DazDb::rlen(int read) const { return impl_->db.reads[read].rlen; }
int DazDb::nreads() const { return impl_->db.nreads; }
void* DazDb::handle() { return &impl_->db; }
void DazDb::export_packed(std::vector<uint64_t>& boff, std::vector<uint32_t>& rlen, std::vector<uint8_t>& bases) {
    HITS_DB& db = impl_->db;
    boff.resize((size_t)db.nreads); rlen.resize((size_t)db.nreads);
    for (int i = 0; i < db.nreads; i++) { boff[(size_t)i] = (uint64_t)db.reads[i].boff; rlen[(size_t)i] = (uint32_t)db.reads[i].rlen; }
    const std::string path = std::string(db.path) + ".bps";     // HITS_DB.path = "<dir>/.<root>" (DB.c Open_DB)
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) throw PacBio::DagCon::IOException("cannot open " + path);
    fseeko(f, 0, SEEK_END);
    const off_t n = ftello(f);
    fseeko(f, 0, SEEK_SET);
    bases.resize((size_t)n);
    const size_t got = n ? fread(bases.data(), 1, (size_t)n, f) : 0;
    fclose(f);
    if (got != (size_t)n) throw PacBio::DagCon::IOException("short read on " + path);
}

// ------------------------------------------------------------------------------------------ realignment (host form)
struct Realizer::Impl {
    DazDb db;
    Work_Data* work = nullptr;
    char* abuf = nullptr;
    char* bbuf = nullptr;
    char* tbuf = nullptr;
    int tspace;
    Impl(const std::string& path, int ts) : db(path), tspace(ts) {
        HITS_DB* h = (HITS_DB*)db.handle();
        work = New_Work_Data();
        abuf = New_Read_Buffer(h);
        bbuf = New_Read_Buffer(h);
        tbuf = New_Read_Buffer(h);
        if (!work || !abuf || !bbuf || !tbuf) throw PacBio::DagCon::IOException("Failed to allocate dazzler work buffers");
        if (h->nreads > 0 && h->reads[0].rlen > 0) {   // opens the .bps now, under the lock (see g_daz_open_mu)
            std::lock_guard<std::mutex> lk(g_daz_open_mu);
            if (!Load_Subread(h, 0, 0, 1, tbuf, 0)) throw PacBio::DagCon::IOException("Failed to open the DB's base file");
        }
    }
    ~Impl() {
        if (abuf) free(abuf - 1);
        if (bbuf) free(bbuf - 1);
        if (tbuf) free(tbuf - 1);
        if (work) Free_Work_Data(work);
    }
};
Realizer::Realizer(const std::string& db_path, int tspace) : impl_(new Impl(db_path, tspace)) {}
Realizer::~Realizer() {}
const ReadLengths& Realizer::lengths() const { return impl_->db; }

// The A read of a pile as ASCII (DazAlnProvider.cpp:119-131).
void Realizer::target_seq(int aread, int length, std::string& out) {
    out.resize((size_t)length);
    const char* s = Load_Subread((HITS_DB*)impl_->db.handle(), aread, 0, length, impl_->tbuf, 0);
    for (int i = 0; i < length; i++) out[(size_t)i] = kToU[(int)s[i]];
}

// One overlap -> dagcon::Alignment (DazAlnProvider.cpp:311-366): load both reads with a 10-base border, complement B
// when the overlap says so, rebuild the exact alignment segment by segment from the trace points
// (Compute_Trace_PTS, GREEDIEST: DALIGNER/align.c:4799-4896 -> iter_np :4132-4459), then write the gapped strings.
void Realizer::realize(const LasRec& rec, int alen, int blen, dagcon::Alignment& dest) {
    HITS_DB* db = (HITS_DB*)impl_->db.handle();
    Path path;
    std::vector<uint16_t> points(rec.trace);            // the dazzler routine is handed a mutable trace
    points.resize(points.size() + 2);
    path.trace = points.data();
    path.tlen = rec.tlen; path.diffs = rec.diffs;
    path.abpos = rec.abpos; path.bbpos = rec.bbpos; path.aepos = rec.aepos; path.bepos = rec.bepos;
    ::Alignment aln;
    aln.path = &path;
    aln.alen = alen; aln.blen = blen; aln.flags = rec.flags;

    int amin = std::max(rec.abpos - kBorder, 0), amax = std::min(rec.aepos + kBorder, alen);
    int bmin, bmax;
    if (COMP(rec.flags)) {
        bmin = std::max((blen - rec.bepos) - kBorder, 0);
        bmax = std::min((blen - rec.bbpos) + kBorder, blen);
    } else {
        bmin = std::max(rec.bbpos - kBorder, 0);
        bmax = std::min(rec.bepos + kBorder, blen);
    }
    char* aseq = Load_Subread(db, rec.aread, amin, amax, impl_->abuf, 0);
    char* bseq = Load_Subread(db, rec.bread, bmin, bmax, impl_->bbuf, 0);
    aln.aseq = aseq - amin;
    if (COMP(rec.flags)) {
        Complement_Seq(bseq, bmax - bmin);
        aln.bseq = bseq - (blen - bmax);
    } else {
        aln.bseq = bseq - bmin;
    }
    Compute_Trace_PTS(&aln, impl_->work, impl_->tspace, GREEDIEST);

    dest.id = std::to_string(rec.aread + 1);
    dest.sid = std::to_string(rec.bread + 1);
    dest.tlen = (uint32_t)alen;
    dest.start = (uint32_t)(rec.abpos + 1);
    dest.end = (uint32_t)(rec.aepos + 1);
    dest.strand = '+';
    // trace -> columns.  A negative value -p: copy up to A position p, then a B base against a gap in A; a positive
    // value p: copy up to B position p, then an A base against a gap in B.  The tail runs to aepos INCLUSIVE (the
    // reference's `while (i <= p)`, :412-416): one column past the overlap, possibly the '.' terminator of a read.
    const int* tr = (const int*)path.trace;
    const char* a = aln.aseq;
    const char* b = aln.bseq;
    int i = path.abpos, j = path.bbpos;
    std::string& ts = dest.tstr;
    std::string& qs = dest.qstr;
    ts.clear(); qs.clear();
    const size_t guess = (size_t)(rec.aepos - rec.abpos) + (size_t)path.tlen + 2;
    ts.reserve(guess); qs.reserve(guess);
    for (int c = 0; c < path.tlen; c++) {
        const int p = tr[c];
        if (p < 0) {
            for (const int stop = -p; i != stop; ) { ts.push_back(kToU[(int)a[i++]]); qs.push_back(kToU[(int)b[j++]]); }
            ts.push_back('-'); qs.push_back(kToU[(int)b[j++]]);
        } else {
            for (; j != p; ) { ts.push_back(kToU[(int)a[i++]]); qs.push_back(kToU[(int)b[j++]]); }
            ts.push_back(kToU[(int)a[i++]]); qs.push_back('-');
        }
    }
    for (; i <= path.aepos; ) { ts.push_back(kToU[(int)a[i++]]); qs.push_back(kToU[(int)b[j++]]); }
}

}  // namespace daz

}  // namespace dgh

// ------------------------------------------------------------------------------------------ the reference-shaped provider
using namespace dgh::daz;

struct DazAlnProvider::Impl {
    dgh::ProgramOpts opts;
    std::unique_ptr<Realizer> rz;
    std::unique_ptr<PileScanner> scan;
    Pile pile;
    int last_id = -1, last_len = 0;
    bool done = false;
};

DazAlnProvider::DazAlnProvider(const dgh::ProgramOpts& popts) : impl_(new Impl) {
    impl_->opts = popts;
    LasFile probe(popts.alnFile);       // header first: tspace sizes the work data (and a bad LAS fails before the DB is touched)
    const int tspace = probe.tspace;
    impl_->rz.reset(new Realizer(popts.seqFile, tspace));
    impl_->scan.reset(new PileScanner(popts.alnFile, impl_->rz->lengths(), popts.properOvls));
}
DazAlnProvider::~DazAlnProvider() {}

// Reference :79-117.  Piles that the target list excludes or that keep fewer than minCov alignments are passed over
// inside the call; the call that consumes the last pile returns false (dest may still hold that pile's alignments).
bool DazAlnProvider::nextTarget(std::vector<dagcon::Alignment>& dest) {
    dest.clear();
    Impl& m = *impl_;
    if (m.done) return false;
    std::vector<const LasRec*> recs;
    std::vector<int> blens;
    for (;;) {
        if (!m.scan->next(m.pile)) { m.done = true; return false; }
        m.last_id = m.pile.aread; m.last_len = m.pile.length;
        const bool last = m.scan->at_end();
        const int tid = m.pile.aread + 1;
        bool keep = m.opts.targets.empty() || m.opts.targets.count(tid);
        if (keep) {
            m.pile.select(m.opts.maxHits, m.opts.sortCov, recs, blens);
            if (recs.size() < m.opts.minCov) keep = false;
        }
        if (keep) {
            dest.resize(recs.size());
            for (size_t k = 0; k < recs.size(); k++) m.rz->realize(*recs[k], m.pile.length, blens[k], dest[k]);
        }
        if (last) { m.done = true; return false; }
        if (keep) return true;
    }
}

bool DazAlnProvider::nextTarget(std::string& targSeq, std::vector<dagcon::Alignment>& dest) {
    const bool more = nextTarget(dest);
    Impl& m = *impl_;
    if (m.last_id >= 0) m.rz->target_seq(m.last_id, m.last_len, targSeq);   // reference :119-131 loads it whatever dest holds
    return more;
}

// ------------------------------------------------------------------------------------------ the dazcon pipeline
namespace dgh {

namespace {

struct SelPile {                 // a pile after hit selection: what a worker rebuilds
    int aread = 0, length = 0;
    std::vector<LasRec> recs;
    std::vector<int> blens;
};
struct DazUnit {
    uint64_t seq = 0;
    std::vector<SelPile> piles;
    size_t cols = 0;             // estimate of the alignment columns (sizes the host batch)
};

// scanner + realignment workers, shared by run_daz_pipeline (dispatcher behind it) and pack_daz (a sink behind it)
struct DazFront {
    const ProgramOpts& opts;
    bool pinned;
    RunStats& stats;
    std::chrono::steady_clock::time_point t0 = std::chrono::steady_clock::now();
    Channel<std::unique_ptr<DazUnit>> units{8};
    Channel<std::unique_ptr<HostBatch>> free_batches{1024};
    OrderedOutboxT<HostBatch> outbox;
    std::mutex err_mu;
    std::string ferr;
    int frc = 0;
    int tspace = 0;
    std::unique_ptr<DazDb> lens;
    std::thread scanner;
    std::vector<std::thread> workers;
    std::atomic<size_t> pool_made{0};
    std::atomic<int> live{0};
    size_t pool_max = 0;

    // device realignment (dgc_submit_daz): the batches carry overlaps + trace points instead of strings
    bool device_realign = false;
    int tbytes = 1;
    std::vector<uint64_t> db_boff;
    std::vector<uint32_t> db_rlen;
    std::vector<uint8_t> db_bases;
    dgc_daz_db db_view{};

    DazFront(const ProgramOpts& o, bool pin, RunStats& st) : opts(o), pinned(pin), stats(st) {}

    // Can the device take this unit?  Same conditions as the library's own validation: byte-wide trace values, a
    // trace of one pair per trace-spacing block of A, the last block's B stretch within the kernel's limit.
    bool unit_device_ok(const DazUnit& u) const {
        if (!device_realign || tbytes != 1 || tspace <= 0 || tspace > 125) return false;
        for (const auto& p : u.piles)
            for (const auto& r : p.recs) {
                if (r.tlen < 2 || (r.tlen & 1)) return false;
                const int want = r.aepos / tspace - r.abpos / tspace + ((r.aepos % tspace) ? 1 : 0);
                if (r.tlen / 2 != (want ? want : 1)) return false;
                long sb = r.bbpos;
                for (int i = 0; i + 1 < r.tlen / 2; i++) sb += r.trace[(size_t)(2 * i + 1)];
                if (sb > r.bepos || r.bepos - sb > 256) return false;
                if (r.abpos < 0 || r.aepos < r.abpos || r.aepos > p.length || r.bbpos < 0 || r.bepos < r.bbpos) return false;
            }
        return true;
    }
    void pack_device(const DazUnit& u, HostBatch& hb) {
        if (!hb.daz) hb.daz.reset(new DazPayload());
        DazPayload& z = *hb.daz;
        z = DazPayload();
        z.tspace = (uint32_t)tspace;
        for (const auto& p : u.piles) {
            z.tgt_aread.push_back((uint32_t)p.aread);
            hb.ids.push_back(std::to_string(p.aread + 1));
            for (const auto& r : p.recs) {
                z.ovl_bread.push_back((uint32_t)r.bread);
                z.ovl_comp.push_back(r.flags & 1u);
                z.ovl_abpos.push_back(r.abpos); z.ovl_aepos.push_back(r.aepos); z.ovl_bbpos.push_back(r.bbpos); z.ovl_bepos.push_back(r.bepos);
                z.ovl_trace_off.push_back(z.trace.size());
                z.ovl_tlen.push_back((uint32_t)r.tlen);
                for (int i = 0; i < r.tlen; i++) z.trace.push_back((uint8_t)r.trace[(size_t)i]);
            }
            z.tgt_ovl_begin.push_back((uint32_t)z.ovl_bread.size());
        }
        hb.n_targets = (uint32_t)u.piles.size();
        hb.n_alns = (uint32_t)z.ovl_bread.size();
    }

    void fail(int code, const std::string& msg) {
        std::lock_guard<std::mutex> lk(err_mu);
        if (!frc) { frc = code; ferr = msg; }
    }
    bool failed() {
        std::lock_guard<std::mutex> lk(err_mu);
        return frc != 0;
    }

    // opens the inputs (errors come back here, before any thread runs)
    int open(std::string& err) {
        try {
            LasFile probe(opts.alnFile);
            tspace = probe.tspace;
            tbytes = probe.tbytes;
            lens.reset(new DazDb(opts.seqFile));
            device_realign = !opts.hostRealign && pinned && tbytes == 1;   // pack_daz (tests, bench) asks for strings
            if (device_realign) {
                lens->export_packed(db_boff, db_rlen, db_bases);
                db_view.n_reads = (uint32_t)db_rlen.size();
                db_view.read_boff = db_boff.data(); db_view.read_len = db_rlen.data();
                db_view.bases = db_bases.data(); db_view.bases_bytes = db_bases.size();
            }
        } catch (const std::exception& e) { err = e.what(); return DGC_E_INVALID; }
        return 0;
    }

    void start(size_t pool) {
        pool_max = pool;
        // ---- scanner: LAS -> piles -> hit selection -> units of whole targets (reference :79-117 filters applied here)
        scanner = std::thread([this] {
            try {
                PileScanner scan(opts.alnFile, *lens, opts.properOvls);
                Pile pile;
                std::vector<const LasRec*> recs;
                std::vector<int> blens;
                std::unique_ptr<DazUnit> u(new DazUnit());
                uint64_t seq = 0;
                // Unit sizes.  A batch costs the latency of the longest dependent chain of the graph sweeps however few
                // targets it holds, so batches want to be large.  With the realignment on the device the scanner is the
                // only host work per target and a large unit is cut in milliseconds: start at 1024 targets.  With the
                // dazzler library rebuilding the alignments on the host threads, small early units start the GPUs sooner.
                size_t want = std::min<size_t>(opts.batchTargets, device_realign ? 1024 : 32);
                auto flush = [&] {
                    if (u->piles.empty()) return;
                    u->seq = seq++;
                    units.push(std::move(u));
                    u.reset(new DazUnit());
                    want = std::min(opts.batchTargets, want * (device_realign ? 4 : 2));
                };
                while (scan.next(pile)) {
                    stats.targets_in++;
                    const int tid = pile.aread + 1;
                    if (!opts.targets.empty() && !opts.targets.count(tid)) { stats.targets_dropped++; continue; }
                    pile.select(opts.maxHits, opts.sortCov, recs, blens);
                    if (recs.size() < opts.minCov) { stats.targets_dropped++; continue; }
                    SelPile sp;
                    sp.aread = pile.aread; sp.length = pile.length;
                    sp.blens = blens;
                    sp.recs.reserve(recs.size());
                    size_t cols = 0;
                    for (const LasRec* r : recs) { sp.recs.push_back(*r); cols += (size_t)(r->aepos - r->abpos); }
                    stats.alignments += recs.size();
                    u->cols += cols;
                    u->piles.push_back(std::move(sp));
                    if (u->piles.size() >= want || 3 * u->cols > opts.batchBytes) flush();
                }
                flush();
            } catch (const std::exception& e) { fail(DGC_E_INVALID, e.what()); }
            stats.scan_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            units.close();
        });
        // ---- workers: rebuild the alignments of a unit straight into a (page-locked) batch
        const int nw = std::max(1, opts.threads);
        live = nw;
        for (int w = 0; w < nw; w++) {
            workers.emplace_back([this] {
                std::unique_ptr<Realizer> rz;   // the host form: one dazzler DB handle + Work_Data per worker, made on first use
                auto need_rz = [&] {
                    if (!rz) try { rz.reset(new Realizer(opts.seqFile, tspace)); } catch (const std::exception& e) { fail(DGC_E_INVALID, e.what()); }
                };
                std::vector<dagcon::Alignment> alns;
                std::string tseq;
                for (;;) {
                    std::unique_ptr<HostBatch> hb;
                    if (!free_batches.try_pop(hb)) {
                        if (pool_made.fetch_add(1) < pool_max) {
                            try { hb.reset(new HostBatch(pinned, 0, 0, 0)); } catch (const std::exception& e) { fail(DGC_E_NOMEM, e.what()); break; }
                        } else if (!free_batches.pop(hb)) break;
                    }
                    std::unique_ptr<DazUnit> u;
                    if (!units.pop(u)) { free_batches.push(std::move(hb)); break; }
                    try {
                        hb->clear();
                        hb->seq = u->seq;
                        hb->daz.reset();
                        hb->packed4 = !opts.asciiWire && pinned;   // host-rebuilt strings are over "ACGT-.": always packable
                        if (unit_device_ok(*u)) {
                            pack_device(*u, *hb);
                            u.reset();
                            outbox.put(std::move(hb));
                            continue;
                        }
                        size_t nal = 0, bb = 0;
                        for (auto& p : u->piles) { nal += p.recs.size(); bb += (size_t)p.length + 16; }
                        hb->reserve((hb->packed4 ? 1 : 2) * (u->cols + u->cols / 6) + 64 * nal + bb, nal, u->piles.size());
                        need_rz();
                        if (rz) for (auto& p : u->piles) {
                            alns.resize(p.recs.size());
                            for (size_t k = 0; k < p.recs.size(); k++) rz->realize(p.recs[k], p.length, p.blens[k], alns[k]);
                            rz->target_seq(p.aread, p.length, tseq);
                            hb->add_target(alns, &tseq);
                        }
                    } catch (const std::exception& e) { fail(DGC_E_INVALID, e.what()); }
                    u.reset();
                    outbox.put(std::move(hb));
                }
                if (--live == 0) outbox.finish();
            });
        }
    }

    void stop() {
        units.close();
        free_batches.close();
        if (scanner.joinable()) scanner.join();
        for (auto& t : workers) t.join();
        workers.clear();
    }
};

}  // namespace

// Test hook: what the DEVICE writes for every selected overlap (dgc_debug_daz_strings), as "A sid start q t" lines
// under "T id tlen n" headers -- the host provider's dump format without the target sequence.
int daz_device_strings(const ProgramOpts& opts, std::string& text, std::string& err) {
    RunStats st;
    DazFront fr(opts, true, st);
    int rc = fr.open(err);
    if (rc) return rc;
    if (!fr.device_realign) { err = "this LAS cannot take the device path (two-byte trace values)"; return DGC_E_DAZ; }
    dgc_opts o{};
    o.min_cov = (uint32_t)opts.minCov; o.min_len = (uint32_t)opts.minLen; o.trim = opts.trim; o.device = 0;
    dgc_handle* h = nullptr;
    if ((rc = dgc_create(&o, &h)) != 0) { err = dgc_last_error(nullptr); return rc; }
    if ((rc = dgc_daz_set_db(h, &fr.db_view)) != 0) { err = dgc_last_error(h); dgc_destroy(h); return rc; }
    PileScanner scan(opts.alnFile, *fr.lens, opts.properOvls);
    Pile pile;
    std::vector<const LasRec*> recs;
    std::vector<int> blens;
    std::vector<uint8_t> q, t;
    std::vector<uint64_t> off;
    std::vector<uint32_t> len;
    while (scan.next(pile)) {
        if (!opts.targets.empty() && !opts.targets.count(pile.aread + 1)) continue;
        pile.select(opts.maxHits, opts.sortCov, recs, blens);
        if (recs.size() < opts.minCov) continue;
        DazUnit u;
        SelPile sp;
        sp.aread = pile.aread; sp.length = pile.length; sp.blens = blens;
        for (const LasRec* r : recs) sp.recs.push_back(*r);
        u.piles.push_back(std::move(sp));
        if (!fr.unit_device_ok(u)) { err = "a pile the device path cannot take"; dgc_destroy(h); return DGC_E_DAZ; }
        HostBatch hb(false, 0, 0, 0);
        fr.pack_device(u, hb);
        const dgc_daz_batch zb = hb.daz->view();
        size_t cap = 0;
        for (const LasRec* r : recs) { cap += (size_t)(r->aepos - r->abpos + 1); for (int i = 0; i < r->tlen; i += 2) cap += r->trace[(size_t)i]; }
        q.resize(cap + 16); t.resize(cap + 16); off.resize(recs.size() + 1); len.resize(recs.size() + 1);
        if ((rc = dgc_debug_daz_strings(h, &zb, q.data(), t.data(), cap + 16, off.data(), len.data())) != 0) { err = dgc_last_error(h); dgc_destroy(h); return rc; }
        text += "T " + std::to_string(pile.aread + 1) + " " + std::to_string(pile.length) + " " + std::to_string(recs.size()) + "\n";
        for (size_t k = 0; k < recs.size(); k++) {
            text += "A " + std::to_string(recs[k]->bread + 1) + " " + std::to_string(recs[k]->abpos + 1) + " ";
            text.append((const char*)q.data() + off[k], len[k]);
            text.push_back(' ');
            text.append((const char*)t.data() + off[k], len[k]);
            text.push_back('\n');
        }
    }
    dgc_destroy(h);
    return 0;
}

int pack_daz(const ProgramOpts& opts, bool pinned, BatchSink sink, void* user, RunStats& stats, std::string& err) {
    DazFront fr(opts, pinned, stats);
    int rc = fr.open(err);
    if (rc) return rc;
    fr.start((size_t)std::max(2, opts.threads) + 2);
    for (uint64_t seq = 0;; seq++) {
        auto hb = fr.outbox.take(seq);
        if (!hb) break;
        if (!fr.failed()) { sink(*hb, user); stats.batches++; }
        fr.free_batches.push(std::move(hb));
    }
    fr.stop();
    stats.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - fr.t0).count();
    err = fr.ferr;
    return fr.frc;
}

int run_daz_pipeline(const ProgramOpts& opts, FILE* out, RunStats& stats, std::string& err) {
    if (dgc_device_count() <= 0) {
        err = "no CUDA device: the consensus path runs on the GPU only (there is no CPU fallback)";
        return DGC_E_NODEVICE;
    }
    DazFront fr(opts, true, stats);
    int rc = fr.open(err);
    if (rc) return rc;
    const int G = opts.gpus > 0 ? opts.gpus : std::max(1, dgc_device_count());
    fr.start((size_t)G * DGC_MAX_INFLIGHT + (size_t)std::max(1, opts.threads) + 2);
    BatchFeed feed;
    feed.take = [&](uint64_t seq) { return fr.outbox.take(seq); };
    feed.recycle = [&](std::unique_ptr<HostBatch> hb) { fr.free_batches.push(std::move(hb)); };
    feed.failed = [&] { return fr.failed(); };
    if (fr.device_realign) feed.daz_db = &fr.db_view;
    rc = dispatch_batches(opts, feed, out, stats, err, true);
    fr.stop();
    if (!rc && fr.frc) { rc = fr.frc; err = fr.ferr; }
    stats.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - fr.t0).count();
    return rc;
}

}  // namespace dgh
