// dagcon_host.hpp -- C++ host side above the C ABI (include/dagcon_b200.h).
//
// Mirrors the reference's interfaces for this path so that its call sites read the same:
//   dagcon::Alignment          reference src/cpp/Alignment.hpp:12-42
//   AlnProvider                reference src/cpp/AlnProvider.hpp:32-46
//   BlasrM5AlnProvider         reference src/cpp/BlasrM5AlnProvider.hpp (m5 and 'pre' text, grouped by target)
//   ProgramOpts                reference src/cpp/ProgramOpts.hpp:8-36
//   SimpleAligner              reference src/cpp/SimpleAligner.hpp (the -a step; arithmetic builder-written, see simple_align.cpp)
// and replaces the Reader -> BoundedBuffer -> N x Consensus -> BoundedBuffer -> Writer thread pipeline
// (reference src/cpp/main.cpp:37-176,248-274) with: one scanner thread that cuts the input text into
// target groups, N packer threads that parse those lines straight into pinned dgc_batch buffers, and one
// dispatcher thread that deals batches round-robin to the GPUs and gathers the CnsResults in input order.
#pragma once
#include <cstdint>
#include <cstdio>
#include <exception>
#include <functional>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include "../../include/dagcon_b200.h"

namespace dagcon {
struct Alignment {
    uint32_t tlen = 0;   // length of the sequence being corrected
    uint32_t start = 0;  // 1-based target offset of the first target base
    uint32_t end = 0;
    std::string id;      // target id
    std::string sid;     // supporting read id
    char strand = '+';
    std::string qstr, tstr;  // equal-length gapped strings
};
}  // namespace dagcon

namespace PacBio {
namespace DagCon {
class IOException : public std::exception {
public:
    explicit IOException(std::string desc) : desc_(std::move(desc)) {}
    const char* what() const noexcept override { return desc_.c_str(); }

private:
    std::string desc_;
};
}  // namespace DagCon
}  // namespace PacBio

namespace M5Exception {
struct FormatError {
    std::string msg;
};
struct FileOpenError {};
struct SortError {};
}  // namespace M5Exception

class AlnProvider {
public:
    // Clears dest, fills it with the next target's alignments; false when the input is exhausted
    // (the last call may still have filled dest).
    virtual bool nextTarget(std::vector<dagcon::Alignment>& dest) = 0;
    virtual bool nextTarget(std::string& targSeq, std::vector<dagcon::Alignment>& dest) = 0;
    virtual ~AlnProvider() {}
};

namespace dgh {

enum class TextFormat { M5, PRE };

// One input line, already split: views into the line's bytes (no copies).
struct LineFields {
    const char* sid = nullptr; size_t sid_len = 0;
    const char* id = nullptr; size_t id_len = 0;
    uint32_t tlen = 0, start = 0, end = 0;
    char strand = '+';
    const char* q = nullptr; size_t q_len = 0;
    const char* t = nullptr; size_t t_len = 0;
};

// Splits on single spaces, skipping empty fields (reference Alignment.cpp:44-55,82-93).
// m5: id=f[5], tlen=f[6], start=f[7]+1, strand=f[9][0], q=f[16], t=f[18]  (Alignment.cpp:57-79)
// pre: sid=f[0], id=f[1], strand=f[2][0], tlen=f[3], start=f[4], end=f[5], q=f[6], t=f[7] (:95-111)
// Returns false on an empty line; throws M5Exception::FormatError on a short or non-numeric line.
bool split_line(const char* p, size_t n, TextFormat fmt, LineFields& out);
// Only the target-id field (what the scanner needs to group lines).
bool line_target_id(const char* p, size_t n, TextFormat fmt, const char*& id, size_t& id_len);
// revComp of reference Alignment.cpp:15-26: complements upper-case ACGT only, then reverses.
void revcomp_into(const char* src, size_t n, uint8_t* dst);
// The `-a` arithmetic (simple_align.cpp): aligns the ungapped query segment q[0..qn) to the ungapped target segment
// t[0..tn) and writes the gapped strings.  The result is local: it starts and ends on a matching column and covers the
// target bases [t_begin, t_end) of the segment.  False when nothing aligns.
bool simple_align(const char* q, size_t qn, const char* t, size_t tn, std::string& qout, std::string& tout,
                  uint32_t& t_begin, uint32_t& t_end);

// One 'pre' record after the -a step (text_input.cpp: align_pre_line).
struct PreAligned {
    bool ok = false;      // false: blank line, or nothing aligns (the record is dropped)
    uint32_t start = 0;   // 1-based, target orientation
    std::string q, t;     // gapped strings, target orientation
};
// False on a blank line; throws M5Exception::FormatError like split_line.
bool align_pre_line(const char* p, size_t n, PreAligned& out);

// The dazcon form of a batch (include/dagcon_b200.h: dgc_daz_batch): overlaps with their trace points; the gapped
// strings are built on the device.
struct DazPayload {
    uint32_t tspace = 0;
    std::vector<uint32_t> tgt_aread, tgt_ovl_begin{0}, ovl_bread, ovl_comp, ovl_tlen;
    std::vector<int32_t> ovl_abpos, ovl_aepos, ovl_bbpos, ovl_bepos;
    std::vector<uint64_t> ovl_trace_off;
    std::vector<uint8_t> trace;
    dgc_daz_batch view() const;
};

// The `-a` form of a batch (include/dagcon_b200.h: dgc_pre_batch): ungapped 'pre' records; aligned on the device.
struct PrePayload {
    std::vector<uint32_t> tgt_tlen, tgt_rec_begin{0}, rec_start, rec_strand, rec_qlen, rec_tseglen;
    std::vector<uint64_t> rec_q_off, rec_t_off;
    std::vector<uint8_t> arena;
    dgc_pre_batch view() const;
};

// A packed batch in host memory (pinned when `pinned`), laid out as include/dagcon_b200.h: dgc_batch.
class HostBatch {
public:
    HostBatch(bool pinned, size_t arena_cap, size_t aln_cap, size_t tgt_cap);
    ~HostBatch();
    HostBatch(const HostBatch&) = delete;
    HostBatch& operator=(const HostBatch&) = delete;

    void clear();
    // Capacity for what is about to be added (one allocation instead of doubling).
    void reserve(size_t arena_need, size_t alns, size_t targets);
    // Appends one target given its lines.  All lines must carry the same target id.
    // `pre` (optional, 'pre' format only): the lines' alignments, computed by the caller (in parallel).
    void add_target_lines(const std::vector<std::pair<const char*, size_t>>& lines, TextFormat fmt, const PreAligned* pre = nullptr);
    // Appends one target from provider output (the reference-facing form).
    void add_target(const std::vector<dagcon::Alignment>& alns, const std::string* backbone = nullptr);
    dgc_batch view() const;

    uint32_t n_targets = 0, n_alns = 0;
    uint64_t arena_bytes = 0;
    bool has_backbone = false;
    // The strings are stored as 4-bit codes, two columns per byte (DGC_BATCH_PACKED4): half the bytes to page-lock and to
    // send.  Set before the first add_*; an add_* that meets a character outside DGC_PACK4_ALPHABET throws NeedAscii
    // and the caller packs the unit again with packed4 = false.
    bool packed4 = false;
    struct NeedAscii {};
    std::vector<std::string> ids;   // Alignment::id of alns[0] per target
    uint64_t seq = 0;               // position in the input stream (batch number)
    std::unique_ptr<DazPayload> daz;   // set: submit with dgc_submit_daz (n_targets and ids are filled, the arrays below unused)
    std::unique_ptr<PrePayload> pre;   // set: submit with dgc_submit_pre (likewise)

    uint32_t* tgt_tlen = nullptr;
    uint32_t* tgt_aln_begin = nullptr;
    uint64_t* tgt_bb_off = nullptr;
    uint32_t* aln_start = nullptr;
    uint32_t* aln_len = nullptr;
    uint64_t* aln_q_off = nullptr;
    uint64_t* aln_t_off = nullptr;
    uint8_t* arena = nullptr;

private:
    void* alloc(size_t bytes);
    void release(void* p);
    void need_arena(size_t extra);
    void need_alns(size_t extra);
    void need_targets(size_t extra);
    bool pinned_;
    size_t arena_cap_ = 0, aln_cap_ = 0, tgt_cap_ = 0;
};

// Options shared by the CLI and the C test hooks (mirror of ProgramOpts + dispatcher knobs).
struct ProgramOpts {
    size_t minCov = 6;
    size_t minLen = 500;
    unsigned int trim = 50;
    int threads = 4;        // -j: packer threads here (the reference's consensus threads)
    bool align = false;     // -a: input is 'pre' and must be aligned first
    bool verbose = false;
    std::string input = "-";
    // dazcon (reference src/cpp/ProgramOpts.hpp:24-35)
    std::string alnFile;            // -a: the daligner .las file
    std::string seqFile;            // -s: the DAZZ_DB .db
    unsigned int maxHits = 85;      // -m
    bool sortCov = false;           // -x
    bool properOvls = false;        // -o
    std::set<int> targets;          // positional: limit consensus to these (1-based) target ids
    bool hostRealign = false;       // --host-realign: rebuild the alignments on the host threads (dazzler's Compute_Trace_PTS)
                                    // instead of on the device
    // dispatcher
    int gpus = 0;           // 0 = all visible devices
    size_t batchTargets = 1024;         // a batch closes at this many targets or this much input text (the first four
    size_t batchBytes = 768ull << 20;   // units at a quarter of it): several are packed at once, the page-locked ring stays small
    bool stats = false;
    bool hostAlign = false;             // -a with --host-align: align 'pre' records on the host threads instead of on the device
    bool asciiWire = false;             // --ascii-wire: send the alignment strings as ASCII instead of 4-bit codes
};

struct RunStats {
    uint64_t targets_in = 0, targets_dropped = 0, alignments = 0, input_bytes = 0;
    uint64_t batches = 0, records = 0, bases = 0;
    double seconds = 0, scan_seconds = 0;
    // where the host front spends its time (--stats): the scanner cutting lines (without its waits for free slots),
    // the packers' parsing + packing summed over threads, and page-locking batch buffers
    double scan_cut_seconds = 0, pack_seconds = 0, pin_seconds = 0;
};
extern double g_pin_seconds;   // accumulated by HostBatch::alloc (process-wide; read by the pipelines)

// Reads `opts.input` ("-" = stdin), runs every target through the GPUs, writes FASTA records to `out`
// in input-target order.  Returns 0 or a DGC_E_* code (message in `err`).
int run_text_pipeline(const ProgramOpts& opts, FILE* out, RunStats& stats, std::string& err);

// Where the dispatcher gets its packed batches from (the text front or the dazcon front), in input order.
struct BatchFeed {
    std::function<std::unique_ptr<HostBatch>(uint64_t seq)> take;   // nullptr = stream finished
    std::function<void(std::unique_ptr<HostBatch>)> recycle;
    std::function<bool()> failed;
    const dgc_daz_db* daz_db = nullptr;   // dazcon with device realignment: uploaded to every handle before the first batch
};
int dispatch_batches(const ProgramOpts& opts, BatchFeed& feed, FILE* out, RunStats& stats, std::string& err, bool dazcon_header);

// dazcon: reads the DAZZ_DB `opts.seqFile` and the LAS `opts.alnFile`, rebuilds every selected overlap's alignment
// from its trace points, runs every target through the GPUs and writes ">id/n/b_e\nSEQ\n" records to `out`
// (reference src/cpp/dazcon.cpp:38-120).
int run_daz_pipeline(const ProgramOpts& opts, FILE* out, RunStats& stats, std::string& err);
// Test hook: the gapped strings the device builds for every selected overlap (text, see daz_provider.cpp).
int daz_device_strings(const ProgramOpts& opts, std::string& text, std::string& err);
// The dazcon front alone (scanner + realignment workers): every packed batch goes to `sink` in input order.
typedef void (*BatchSink)(const HostBatch& batch, void* user);
int pack_daz(const ProgramOpts& opts, bool pinned, BatchSink sink, void* user, RunStats& stats, std::string& err);

// The scanner stage alone, for tests and for callers that bring their own dispatcher: cuts the text into
// batches of whole targets and hands each packed batch to `sink` in input order (sink owns nothing; the
// batch is recycled after it returns).
int pack_text(const ProgramOpts& opts, bool pinned, BatchSink sink, void* user, RunStats& stats, std::string& err);

}  // namespace dgh

// Reference src/cpp/SimpleAligner.hpp: re-aligns one 'pre' record in place (gapped strings, start/end moved to the
// aligned span, '-' strand records reverse-complemented into target orientation, start made 1-based).  Returns false
// (and empties the strings) when nothing aligns; the reference has no such case because blasr always returns blocks.
class SimpleAligner {
public:
    bool align(dagcon::Alignment& aln);
    void operator()(dagcon::Alignment& aln) { align(aln); }
};

// Reference src/cpp/DazAlnProvider.hpp:144-184: the targets of a daligner LAS file (the A reads, in file order) with
// the alignments of their best hits, rebuilt from the trace points (daz_provider.cpp).  Dazzler types stay behind
// the implementation pointer; the dazzler C library (DALIGNER align.c, DAZZ_DB DB.c/QV.c) is linked as it is.
class DazAlnProvider : public AlnProvider {
public:
    explicit DazAlnProvider(const dgh::ProgramOpts& popts);   // throws PacBio::DagCon::IOException
    ~DazAlnProvider() override;
    bool nextTarget(std::vector<dagcon::Alignment>& dest) override;
    bool nextTarget(std::string& targSeq, std::vector<dagcon::Alignment>& dest) override;

private:
    struct Impl;
    std::unique_ptr<Impl> impl_;
};

// Reference-shaped provider over the same line parser (m5 by default, 'pre' when asked).
class BlasrM5AlnProvider : public AlnProvider {
public:
    explicit BlasrM5AlnProvider(const std::string& fpath, dgh::TextFormat fmt = dgh::TextFormat::M5);
    explicit BlasrM5AlnProvider(FILE* stream, dgh::TextFormat fmt = dgh::TextFormat::M5);
    ~BlasrM5AlnProvider() override;
    bool nextTarget(std::vector<dagcon::Alignment>& dest) override;
    bool nextTarget(std::string& targSeq, std::vector<dagcon::Alignment>& dest) override;

private:
    bool read_line(dagcon::Alignment& aln, bool& blank);
    FILE* fp_ = nullptr;
    bool owns_ = false;
    dgh::TextFormat fmt_;
    std::string currId_;
    bool firstAln_ = true;
    dagcon::Alignment prevAln_;
    char* line_ = nullptr;
    size_t line_cap_ = 0;
};
