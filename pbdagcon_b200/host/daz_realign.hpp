// daz_realign.hpp -- internal pieces of the dazcon front (daz_provider.cpp): the mapped LAS reader, the pile / hit
// bookkeeping of reference src/cpp/DazAlnProvider.{hpp,cpp} and the per-overlap realignment.  No dazzler types here:
// DB.h / align.h stay inside the .cpp so that the rest of the host code does not depend on the vendored headers.
#pragma once
#include <cstdint>
#include <memory>
#include <string>
#include <vector>

#include "dagcon_host.hpp"

namespace dgh {
namespace daz {

// One LAS record (dazzler Overlap + Path, DALIGNER/align.h:89-95,295-300); trace values always widened to 16 bits.
struct LasRec {
    int tlen = 0, diffs = 0, abpos = 0, bbpos = 0, aepos = 0, bepos = 0;
    uint32_t flags = 0;
    int aread = 0, bread = 0;
    std::vector<uint16_t> trace;
};

class LasFile {
public:
    explicit LasFile(const std::string& path);   // throws PacBio::DagCon::IOException
    ~LasFile();
    LasFile(const LasFile&) = delete;
    LasFile& operator=(const LasFile&) = delete;
    bool next(LasRec& r);                        // false after novl records (or at a truncated tail)
    int64_t novl = 0;
    int tspace = 0, tbytes = 1;

private:
    int fd_ = -1;
    const uint8_t* map_ = nullptr;
    size_t len_ = 0, pos_ = 0;
    int64_t read_ = 0;
};

struct ReadLengths {
    virtual int rlen(int read) const = 0;
    virtual ~ReadLengths() {}
};

// The trimmed DAZZ_DB (Open_DB + Trim_DB, reference DazAlnProvider.cpp:29-41).
class DazDb : public ReadLengths {
public:
    explicit DazDb(const std::string& path);
    ~DazDb() override;
    int rlen(int read) const override;
    int nreads() const;
    void* handle();   // HITS_DB*
    // the whole trimmed DB as the device wants it (dgc_daz_db): per read its byte offset in the .bps file and its
    // length, plus the .bps file itself
    void export_packed(std::vector<uint64_t>& boff, std::vector<uint32_t>& rlen, std::vector<uint8_t>& bases);

private:
    struct Impl;
    std::unique_ptr<Impl> impl_;
};

// TargetHit (reference DazAlnProvider.hpp:55-96)
struct Hit {
    int aread = -1, bread = 0;
    uint32_t flags = 0;
    int alen = 0, blen = 0;
    float ovlScore = 0.0f, covScore = 0.0f;
    std::vector<LasRec> records;
    void add(LasRec&& rec);
    void score(bool proper);
};

// Target (reference DazAlnProvider.hpp:101-139) without the sequence buffers
struct Pile {
    int aread = 0, length = 0;
    std::vector<Hit> hits;
    void sort_hits(bool sort_cov);
    // sorts, then lists the records of the first max_hits hits (and each one's B-read length); returns their number
    size_t select(unsigned max_hits, bool sort_cov, std::vector<const LasRec*>& recs, std::vector<int>& blens);
};

class PileScanner {
public:
    PileScanner(const std::string& las, const ReadLengths& rl, bool proper);
    bool next(Pile& p);
    bool at_end() const { return !have_pending_; }
    int tspace() const { return las_.tspace; }

private:
    Hit new_hit(LasRec&& rec, int alen);
    LasFile las_;
    const ReadLengths& rl_;
    bool proper_;
    LasRec pending_;
    bool have_pending_ = false;
};

// One dazzler DB handle + Work_Data: rebuilds alignments on the host (the reference's arithmetic, linked as it is).
class Realizer {
public:
    Realizer(const std::string& db_path, int tspace);
    ~Realizer();
    const ReadLengths& lengths() const;
    void realize(const LasRec& rec, int alen, int blen, dagcon::Alignment& dest);
    void target_seq(int aread, int length, std::string& out);

private:
    struct Impl;
    std::unique_ptr<Impl> impl_;
};

}  // namespace daz
}  // namespace dgh
