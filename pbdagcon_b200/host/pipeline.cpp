// pipeline.cpp -- the batched GPU dispatcher that replaces the reference's thread pipeline
// (reference src/cpp/main.cpp:37-176 Reader / Consensus / Writer, :248-274 thread set-up;
//  src/cpp/BoundedBuffer.hpp:18-50).
//
//   scanner (1 thread)   reads the text in 64 MiB blocks, finds line ends and each line's target id, applies the
//                        Reader's coverage filter (main.cpp:66-72), cuts the stream into work units of whole targets
//   packers (-j threads) parse a unit's lines straight into a pinned HostBatch (no per-alignment std::string)
//   dispatcher (caller)  deals batches round-robin over the GPUs (dgc_submit, <= DGC_MAX_INFLIGHT per GPU),
//                        collects in input order (dgc_collect) and writes ">id/b_e\nSEQ\n" (main.cpp:139-145)
#include <unistd.h>
#include <fcntl.h>

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <mutex>
#include <thread>

#include <sys/mman.h>
#include <sys/stat.h>

#include "dagcon_host.hpp"
#include "channels.hpp"

// NVTX ranges (header-only; no cost without a tool attached): a timeline capture of the CLI shows where its wall time goes
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
#define DGH_RANGE_PUSH(name) nvtxRangePushA(name)
#define DGH_RANGE_POP() nvtxRangePop()
#else
#define DGH_RANGE_PUSH(name) ((void)0)
#define DGH_RANGE_POP() ((void)0)
#endif

namespace dgh {
namespace {

struct Block {
    std::unique_ptr<char[]> buf;
    size_t cap = 0, len = 0;
    void* map = nullptr;  // a regular input file is mapped whole instead of being read into blocks
    size_t map_len = 0;
    ~Block() { if (map) munmap(map, map_len); }
};

struct WorkUnit {
    uint64_t seq = 0;
    std::vector<std::shared_ptr<Block>> blocks;
    std::vector<std::pair<const char*, size_t>> lines;
    std::vector<uint32_t> tgt_begin{0};  // CSR over lines
    size_t bytes = 0;
    size_t targets() const { return tgt_begin.size() - 1; }
};

constexpr size_t BLOCK_BYTES = 64u << 20;

class Scanner {
public:
    Scanner(const ProgramOpts& o, TextFormat fmt, Channel<std::unique_ptr<WorkUnit>>& out, RunStats& st)
        : o_(o), fmt_(fmt), out_(out), st_(st) {}

    void run() {
        int fd = 0;
        if (o_.input != "-") {
            fd = open(o_.input.c_str(), O_RDONLY);
            if (fd < 0) throw M5Exception::FileOpenError();
        }
        unit_.reset(new WorkUnit());
        if (fd > 0 && run_mapped(fd)) return;
        auto blk = new_block(BLOCK_BYTES);
        bool eof = false;
        while (!eof) {
            while (blk->len < blk->cap) {  // fill the block
                const ssize_t r = read(fd, blk->buf.get() + blk->len, blk->cap - blk->len);
                if (r < 0) { if (errno == EINTR) continue; throw PacBio::DagCon::IOException("read error on the alignment input"); }
                if (r == 0) { eof = true; break; }
                blk->len += (size_t)r;
                st_.input_bytes += (uint64_t)r;
            }
            const char* p = blk->buf.get();
            const char* e = p + blk->len;
            const char* line = p + scanned_;
            for (;;) {
                const char* nl = (const char*)memchr(line, '\n', (size_t)(e - line));
                if (!nl) break;
                take_line(blk, line, (size_t)(nl - line));
                line = nl + 1;
            }
            if (eof) {
                if (line < e) take_line(blk, line, (size_t)(e - line));  // last line without '\n'
                break;
            }
            // carry the partial last line into a fresh block (a line is always contiguous in one block)
            const size_t tail = (size_t)(e - line);
            auto nb = new_block(std::max(BLOCK_BYTES, 2 * tail + 4096));
            memcpy(nb->buf.get(), line, tail);
            nb->len = tail;
            blk->len -= tail;
            blk = nb;
            scanned_ = 0;
        }
        close_target();
        flush_unit();
        if (fd > 0) close(fd);
    }

private:
    // A regular file: one read-only mapping, scanned in place (no read() copy, no fresh pages to fault in); the
    // work units keep the mapping alive through their block lists.
    bool run_mapped(int fd) {
        struct stat sb;
        if (fstat(fd, &sb) != 0 || !S_ISREG(sb.st_mode) || sb.st_size <= 0) return false;
        const size_t len = (size_t)sb.st_size;
        void* m = mmap(nullptr, len, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m == MAP_FAILED) return false;
        madvise(m, len, MADV_SEQUENTIAL);
        auto blk = std::make_shared<Block>();
        blk->map = m;
        blk->map_len = blk->len = blk->cap = len;
        close(fd);
        st_.input_bytes += len;
        // The text is cut into lines (and every line's target id located) by several threads at once, a gigabyte at a
        // time: one thread finds line ends at ~5 GB/s and faults the mapping in page by page, which for 8 GB of m5
        // text is most of the run.  Grouping the lines into targets and units stays sequential (a few ns per line).
        const char* base = (const char*)m;
        const char* e = base + len;
        const size_t SUPER = 1ull << 30;
        const int nscan = std::max(1, std::min(8, o_.threads / 2));
        struct LineRec { const char* p; uint32_t n; uint32_t id_off; uint32_t idn; };   // id_off = 0xFFFFFFFF: no fields (blank)
        std::vector<std::vector<LineRec>> parts((size_t)nscan);
        auto next_line_start = [&](const char* p) -> const char* {   // first byte after the next newline at or after p
            if (p >= e) return e;
            const char* nl = (const char*)memchr(p, '\n', (size_t)(e - p));
            return nl ? nl + 1 : e;
        };
        const char* sbeg = base;
        while (sbeg < e) {
            const char* se = (size_t)(e - sbeg) > SUPER ? next_line_start(sbeg + SUPER) : e;
            std::vector<const char*> cut((size_t)nscan + 1);
            cut[0] = sbeg; cut[(size_t)nscan] = se;
            for (int i = 1; i < nscan; i++) {
                const char* guess = sbeg + (size_t)(se - sbeg) / (size_t)nscan * (size_t)i;
                cut[(size_t)i] = guess <= cut[(size_t)i - 1] ? cut[(size_t)i - 1] : std::min(se, next_line_start(guess));
            }
            std::string scan_err;
            std::mutex scan_mu;
            auto work = [&](int i) {
                auto& out = parts[(size_t)i];
                out.clear();
                const char* line = cut[(size_t)i];
                const char* end = cut[(size_t)i + 1];
                try {
                    while (line < end) {
                        const char* nl = (const char*)memchr(line, '\n', (size_t)(end - line));
                        const size_t n = nl ? (size_t)(nl - line) : (size_t)(end - line);
                        const char* id; size_t idn;
                        if (n > 0xFFFFFFFFull) throw M5Exception::FormatError{"line too long"};
                        if (line_target_id(line, n, fmt_, id, idn)) out.push_back({line, (uint32_t)n, (uint32_t)(id - line), (uint32_t)idn});
                        else out.push_back({line, (uint32_t)n, 0xFFFFFFFFu, 0});
                        line = nl ? nl + 1 : end;
                    }
                } catch (const M5Exception::FormatError& fe) {
                    std::lock_guard<std::mutex> lk(scan_mu);
                    if (scan_err.empty()) scan_err = fe.msg;
                }
            };
            const auto tc0 = std::chrono::steady_clock::now();
            DGH_RANGE_PUSH("cut lines");
            std::vector<std::thread> th;
            for (int i = 1; i < nscan; i++) th.emplace_back(work, i);
            work(0);
            for (auto& t : th) t.join();
            DGH_RANGE_POP();
            st_.scan_cut_seconds += std::chrono::duration<double>(std::chrono::steady_clock::now() - tc0).count();
            if (!scan_err.empty()) throw M5Exception::FormatError{scan_err};
            for (auto& part : parts)
                for (const LineRec& r : part) take_scanned(blk, r.p, r.n, r.id_off == 0xFFFFFFFFu ? nullptr : r.p + r.id_off, r.idn);
            sbeg = se;
        }
        close_target();
        flush_unit();
        return true;
    }

    std::shared_ptr<Block> new_block(size_t cap) {
        auto b = std::make_shared<Block>();
        b->buf.reset(new char[cap]);
        b->cap = cap;
        return b;
    }

    void take_line(const std::shared_ptr<Block>& blk, const char* p, size_t n) {
        const char* id; size_t idn;
        if (!line_target_id(p, n, fmt_, id, idn)) { take_scanned(blk, p, n, nullptr, 0); return; }
        take_scanned(blk, p, n, id, idn);
    }

    // one line whose target id has been located (id == nullptr: a line without fields)
    void take_scanned(const std::shared_ptr<Block>& blk, const char* p, size_t n, const char* id, size_t idn) {
        if (!id) { blank_pending_ = true; return; }
        if (blank_pending_) throw M5Exception::FormatError{"blank line inside the alignment stream"};
        if (!have_id_ || idn != cur_id_.size() || memcmp(id, cur_id_.data(), idn) != 0) {
            close_target();
            cur_id_.assign(id, idn);
            have_id_ = true;
        }
        if (cur_blocks_.empty() || cur_blocks_.back() != blk) cur_blocks_.push_back(blk);
        cur_lines_.emplace_back(p, n);
        cur_bytes_ += n;
    }

    // Reader's filter (main.cpp:66-72): groups with fewer than minCov alignments never reach a consumer.
    void close_target() {
        if (cur_lines_.empty()) return;
        st_.targets_in++;
        st_.alignments += cur_lines_.size();
        if (cur_lines_.size() < o_.minCov) {
            st_.targets_dropped++;
        } else {
            // the first units are small (the GPUs start while the bulk of the text is still being scanned and packed), the
            // later ones large: a batch costs the latency of its longest graph sweep however few targets it holds
            const size_t div = next_seq_ < 4 ? 4 : 1;
            if (unit_->targets() && (unit_->targets() >= std::max<size_t>(1, o_.batchTargets / div) || unit_->bytes + cur_bytes_ > o_.batchBytes / div)) flush_unit();
            for (auto& b : cur_blocks_)
                if (unit_->blocks.empty() || unit_->blocks.back() != b) unit_->blocks.push_back(b);
            unit_->lines.insert(unit_->lines.end(), cur_lines_.begin(), cur_lines_.end());
            unit_->tgt_begin.push_back((uint32_t)unit_->lines.size());
            unit_->bytes += cur_bytes_;
        }
        cur_lines_.clear();
        cur_blocks_.clear();
        cur_bytes_ = 0;
    }

    void flush_unit() {
        if (unit_->targets() == 0) return;
        unit_->seq = next_seq_++;
        out_.push(std::move(unit_));
        unit_.reset(new WorkUnit());
    }

    const ProgramOpts& o_;
    TextFormat fmt_;
    Channel<std::unique_ptr<WorkUnit>>& out_;
    RunStats& st_;
    std::unique_ptr<WorkUnit> unit_;
    std::string cur_id_;
    bool have_id_ = false, blank_pending_ = false;
    std::vector<std::pair<const char*, size_t>> cur_lines_;
    std::vector<std::shared_ptr<Block>> cur_blocks_;
    size_t cur_bytes_ = 0, scanned_ = 0;
    uint64_t next_seq_ = 0;
};

typedef OrderedOutboxT<HostBatch> OrderedOutbox;

struct Front {  // scanner + packers, shared by pack_text and run_text_pipeline
    const ProgramOpts& o;
    TextFormat fmt;
    bool pinned;
    RunStats& st;
    Channel<std::unique_ptr<WorkUnit>> units{4};
    Channel<std::unique_ptr<HostBatch>> free_batches{1024};
    OrderedOutbox outbox;
    std::vector<std::thread> threads;
    std::mutex err_mu;
    std::string err;
    int rc = 0;
    int live_packers = 0;
    bool wire4 = true;   // 4-bit wire format (ProgramOpts::asciiWire switches it off)

    Front(const ProgramOpts& o_, bool pinned_, RunStats& st_) : o(o_), fmt(o_.align ? TextFormat::PRE : TextFormat::M5), pinned(pinned_), st(st_) { wire4 = !o_.asciiWire; }

    void fail(int code, const std::string& msg) {
        std::lock_guard<std::mutex> lk(err_mu);
        if (!rc) { rc = code; err = msg; }
    }

    // Host batches are page-locked, which costs about a second per few GB: they are made when a packer first needs
    // one (at most `pool` of them), not up front, so a small input never pays for the whole ring.
    size_t pool_max = 0, pool_made = 0;
    std::mutex pool_mu;
    bool take_batch(std::unique_ptr<HostBatch>& hb) {
        if (free_batches.try_pop(hb)) return true;
        {
            std::lock_guard<std::mutex> lk(pool_mu);
            if (pool_made < pool_max) {
                pool_made++;
                hb.reset(new HostBatch(pinned, 0, 0, 0));  // sized by the packer once it holds its unit
                return true;
            }
        }
        return free_batches.pop(hb);
    }

    // -a: the alignments of one unit, by `-j` threads at once (a unit may be the whole input; the DP is the cost)
    void align_unit(const WorkUnit& u, std::vector<PreAligned>& out) {
        out.resize(u.lines.size());
        std::atomic<size_t> next{0};
        std::mutex mu;
        std::string first_err;
        auto work = [&] {
            for (size_t i; (i = next.fetch_add(1)) < u.lines.size();) {
                try {
                    align_pre_line(u.lines[i].first, u.lines[i].second, out[i]);
                } catch (const M5Exception::FormatError& e) {
                    std::lock_guard<std::mutex> lk(mu);
                    if (first_err.empty()) first_err = e.msg;
                    next.store(u.lines.size());
                }
            }
        };
        std::vector<std::thread> pool;
        for (int i = 1; i < std::max(1, o.threads); i++) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
        if (!first_err.empty()) throw M5Exception::FormatError{first_err};
    }

    void start(size_t pool) {
        pool_max = pool;
        // with -a a unit's alignments already use every thread: two packers, so that the next unit is aligned while
        // this one is packed and submitted
        const int np = fmt == TextFormat::PRE ? 2 : std::max(1, o.threads);
        live_packers = np;
        threads.emplace_back([this] {
            const auto t0 = std::chrono::steady_clock::now();
            try {
                Scanner(o, fmt, units, st).run();
            } catch (const M5Exception::FileOpenError&) { fail(DGC_E_INVALID, "Error opening file: " + o.input);
            } catch (const M5Exception::FormatError& e) { fail(DGC_E_INVALID, "Format error. Input: " + o.input + ", Error: " + e.msg);
            } catch (const std::exception& e) { fail(DGC_E_INVALID, e.what()); }
            st.scan_seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
            units.close();
        });
        for (int i = 0; i < np; i++) {
            threads.emplace_back([this] {
                std::unique_ptr<WorkUnit> u;
                for (;;) {
                    // A packer takes its host batch BEFORE it takes a unit: units are handed out in input order, so the
                    // oldest unit not yet packed is always in the hands of a packer that can finish it.  (Unit first,
                    // batch second, lets younger units take the last free batches while the consumer waits for the
                    // oldest one and therefore never collects and recycles: a deadlock, seen with many tiny batches.)
                    std::unique_ptr<HostBatch> hb;
                    try {
                        if (!take_batch(hb)) break;
                    } catch (const std::exception& e) { fail(DGC_E_NOMEM, e.what()); break; }
                    if (!units.pop(u)) { free_batches.push(std::move(hb)); break; }  // hand the batch on: another packer may be waiting for one just to find the input exhausted
                    const auto tp0 = std::chrono::steady_clock::now();
                    DGH_RANGE_PUSH("pack unit");
                    struct PackTimer {
                        Front* f; std::chrono::steady_clock::time_point t0;
                        ~PackTimer() { DGH_RANGE_POP(); const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); std::lock_guard<std::mutex> lk(f->err_mu); f->st.pack_seconds += dt; }
                    } pack_timer{this, tp0};
                    try {
                        hb->pre.reset();
                        if (fmt == TextFormat::PRE && !o.hostAlign && pinned) {
                            // -a: the records go to the device as they are (ungapped segments); dgc_submit_pre aligns them
                            hb->clear();
                            hb->seq = u->seq;
                            hb->pre.reset(new PrePayload());
                            PrePayload& z = *hb->pre;
                            z.arena.reserve(u->bytes + 64 * u->lines.size());
                            for (size_t t = 0; t + 1 < u->tgt_begin.size(); t++) {
                                bool first = true;
                                for (uint32_t li = u->tgt_begin[t]; li < u->tgt_begin[t + 1]; li++) {
                                    LineFields f;
                                    if (!split_line(u->lines[li].first, u->lines[li].second, fmt, f)) continue;
                                    if (first) { z.tgt_tlen.push_back(f.tlen); hb->ids.emplace_back(f.id, f.id_len); first = false; }
                                    z.rec_start.push_back(f.start);
                                    z.rec_strand.push_back(f.strand == '-' ? 1u : 0u);
                                    z.rec_qlen.push_back((uint32_t)f.q_len); z.rec_tseglen.push_back((uint32_t)f.t_len);
                                    z.rec_q_off.push_back(z.arena.size()); z.arena.insert(z.arena.end(), f.q, f.q + f.q_len);
                                    z.rec_t_off.push_back(z.arena.size()); z.arena.insert(z.arena.end(), f.t, f.t + f.t_len);
                                }
                                if (!first) z.tgt_rec_begin.push_back((uint32_t)z.rec_start.size());
                            }
                            hb->n_targets = (uint32_t)z.tgt_tlen.size();
                            hb->n_alns = (uint32_t)z.rec_start.size();
                            u.reset();
                            outbox.put(std::move(hb));
                            continue;
                        }
                        std::vector<PreAligned> aligned;
                        if (fmt == TextFormat::PRE) align_unit(*u, aligned);
                        std::vector<std::pair<const char*, size_t>> lines;
                        // the strings go out as 4-bit codes (half the bytes to page-lock and to send); a unit with a
                        // character outside "-ACGTN." is packed again as ASCII
                        for (int pass = wire4 ? 0 : 1; pass < 2; pass++) {
                            hb->clear();
                            hb->seq = u->seq;
                            hb->packed4 = pass == 0;
                            // m5 lines carry the two strings plus a match pattern of the same length; aligned 'pre'
                            // strings grow by their gaps (a low guess only costs one doubling)
                            const size_t est = fmt == TextFormat::M5 ? u->bytes * 2 / 3 : u->bytes + u->bytes / 4;
                            hb->reserve((hb->packed4 ? est / 2 : est) + 64 * u->lines.size(), u->lines.size(), u->targets());
                            try {
                                for (size_t t = 0; t + 1 < u->tgt_begin.size(); t++) {
                                    lines.assign(u->lines.begin() + u->tgt_begin[t], u->lines.begin() + u->tgt_begin[t + 1]);
                                    hb->add_target_lines(lines, fmt, aligned.empty() ? nullptr : &aligned[u->tgt_begin[t]]);
                                }
                                break;
                            } catch (const HostBatch::NeedAscii&) {
                                if (pass == 1) throw M5Exception::FormatError{"unpackable character in an ASCII batch"};
                            }
                        }
                    } catch (const M5Exception::FormatError& e) { fail(DGC_E_INVALID, "Format error. Input: " + o.input + ", Error: " + e.msg);
                    } catch (const std::exception& e) { fail(DGC_E_NOMEM, e.what()); }
                    u.reset();
                    outbox.put(std::move(hb));
                }
                std::lock_guard<std::mutex> lk(err_mu);
                if (--live_packers == 0) outbox.finish();
            });
        }
    }

    void recycle(std::unique_ptr<HostBatch> hb) { free_batches.push(std::move(hb)); }

    void stop() {
        units.close();
        free_batches.close();
        for (auto& t : threads) t.join();
        threads.clear();
    }
};

void append_int(std::string& s, long long v) {
    char buf[24];
    const int n = snprintf(buf, sizeof buf, "%lld", v);
    s.append(buf, (size_t)n);
}

}  // namespace

int pack_text(const ProgramOpts& opts, bool pinned, BatchSink sink, void* user, RunStats& stats, std::string& err) {
    Front fr(opts, pinned, stats);
    fr.start(std::max(2, opts.threads) + 1);
    for (uint64_t seq = 0;; seq++) {
        auto hb = fr.outbox.take(seq);
        if (!hb) break;
        if (!fr.rc) { sink(*hb, user); stats.batches++; }
        fr.recycle(std::move(hb));
    }
    fr.stop();
    err = fr.err;
    return fr.rc;
}

static double since(std::chrono::steady_clock::time_point t0) {
    return std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
}

// The dispatcher proper, shared by the text front (pbdagcon) and the dazcon front: creates one handle per GPU, deals
// the feed's batches round-robin (dgc_submit, <= DGC_MAX_INFLIGHT per GPU), collects in input order (dgc_collect) and
// writes the FASTA records: ">id/b_e\nSEQ\n" (reference main.cpp:139-145) or, for dazcon, ">id/n/b_e\nSEQ\n" with
// n the running record index in input order (= `dazcon -j 1`, reference dazcon.cpp:62,92-99).
int dispatch_batches(const ProgramOpts& opts, BatchFeed& feed, FILE* out, RunStats& stats, std::string& err, bool dazcon_header) {
    const auto t0 = std::chrono::steady_clock::now();
    int ndev = dgc_device_count();
    if (ndev <= 0) {
        err = "no CUDA device: the consensus path runs on the GPU only (there is no CPU fallback)";
        return DGC_E_NODEVICE;
    }
    // DGC_SHARE_GPUS=1: --gpus N may exceed the visible devices; handle g then drives device g % ndev.  Lets a one-GPU
    // box exercise the round-robin dealing and the ordered gather of the multi-GPU dispatcher (tests).
    const bool share = getenv("DGC_SHARE_GPUS") && getenv("DGC_SHARE_GPUS")[0] == '1';
    const int G = opts.gpus > 0 ? (share ? opts.gpus : std::min(opts.gpus, ndev)) : ndev;
    std::vector<dgc_handle*> hs((size_t)G, nullptr);
    auto destroy_all = [&] { for (auto h : hs) if (h) dgc_destroy(h); };
    {   // one CUDA context per GPU costs 0.3-0.5 s: the handles are created side by side
        std::vector<std::thread> mk;
        std::vector<int> rcs((size_t)G, 0);
        std::vector<std::string> errs((size_t)G);
        for (int g = 0; g < G; g++)
            mk.emplace_back([&, g] {
                dgc_opts o{};
                o.min_cov = (uint32_t)opts.minCov; o.min_len = (uint32_t)opts.minLen; o.trim = opts.trim; o.device = g % ndev;
                rcs[(size_t)g] = dgc_create(&o, &hs[(size_t)g]);
                if (rcs[(size_t)g]) { errs[(size_t)g] = std::string("dgc_create: ") + dgc_last_error(nullptr); return; }
                if (feed.daz_db) {
                    rcs[(size_t)g] = dgc_daz_set_db(hs[(size_t)g], feed.daz_db);
                    if (rcs[(size_t)g]) errs[(size_t)g] = std::string("dgc_daz_set_db: ") + dgc_last_error(hs[(size_t)g]);
                }
            });
        for (auto& t : mk) t.join();
        for (int g = 0; g < G; g++)
            if (rcs[(size_t)g]) { err = errs[(size_t)g]; destroy_all(); return rcs[(size_t)g]; }
    }
    if (opts.verbose) fprintf(stderr, "%s-b200: %.3f s %d GPU handle(s) ready\n", dazcon_header ? "dazcon" : "pbdagcon", since(t0), G);

    std::deque<std::unique_ptr<HostBatch>> inflight;  // in submit (= input) order; batch k runs on GPU k % G
    std::vector<int> on_gpu((size_t)G, 0);
    std::string text;
    long long well = 0;
    int rc = 0;
    const char* prog = dazcon_header ? "dazcon" : "pbdagcon";

    auto collect_oldest = [&]() {
        auto hb = std::move(inflight.front());
        inflight.pop_front();
        const int g = (int)(hb->seq % (uint64_t)G);
        dgc_results r{};
        DGH_RANGE_PUSH("dgc_collect");
        const int c = dgc_collect(hs[(size_t)g], &r);
        DGH_RANGE_POP();
        on_gpu[(size_t)g]--;
        if (opts.verbose) fprintf(stderr, "%s-b200: %.3f s collected batch %llu\n", prog, since(t0), (unsigned long long)hb->seq);
        if (c) {
            if (!rc) { rc = c; err = std::string("dgc_collect: ") + dgc_last_error(hs[(size_t)g]); }
        } else if (!rc) {
            text.clear();
            for (uint32_t t = 0; t < r.n_targets; t++) {
                if (r.tgt_status[t] > DGC_T_SKIPPED)  // rejected, never silently: the reference would walk off its vertex array here
                    fprintf(stderr, "%s: target %s rejected (status %u: %s)\n", prog, hb->ids[t].c_str(), r.tgt_status[t],
                            r.tgt_status[t] == DGC_T_BAD_ALIGNMENT ? "an alignment runs off the backbone" :
                            r.tgt_status[t] == DGC_T_TOO_MANY_ALNS ? "too many alignments" : "internal");
                for (uint32_t f = r.tgt_frag_begin[t]; f < r.tgt_frag_begin[t + 1]; f++) {
                    text.push_back('>');
                    text += hb->ids[t];
                    text.push_back('/');
                    if (dazcon_header) { append_int(text, well++); text.push_back('/'); }
                    append_int(text, r.frag_range0[f]);
                    text.push_back('_');
                    append_int(text, r.frag_range1[f]);
                    text.push_back('\n');
                    text.append((const char*)r.seq_arena + r.frag_seq_off[f], r.frag_seq_len[f]);
                    text.push_back('\n');
                    stats.records++;
                    stats.bases += r.frag_seq_len[f];
                }
            }
            if (!text.empty() && fwrite(text.data(), 1, text.size(), out) != text.size()) { rc = DGC_E_INVALID; err = "write error on the FASTA output"; }
        }
        feed.recycle(std::move(hb));
    };

    for (uint64_t seq = 0;; seq++) {
        auto hb = feed.take(seq);
        if (!hb) break;
        if (rc || feed.failed()) { feed.recycle(std::move(hb)); continue; }
        const int g = (int)(seq % (uint64_t)G);
        while (on_gpu[(size_t)g] >= DGC_MAX_INFLIGHT) collect_oldest();
        // the batch struct is read during submit only; the arrays it points to stay alive in `inflight`
        if (opts.verbose) fprintf(stderr, "%s-b200: %.3f s submit batch %llu (%u targets, %.1f MB) to GPU %d\n", prog, since(t0), (unsigned long long)seq, hb->n_targets, hb->arena_bytes / 1e6, g);
        int c;
        DGH_RANGE_PUSH("dgc_submit");
        if (hb->daz) { const dgc_daz_batch zb = hb->daz->view(); c = dgc_submit_daz(hs[(size_t)g], &zb); }
        else if (hb->pre) { const dgc_pre_batch pb = hb->pre->view(); c = dgc_submit_pre(hs[(size_t)g], &pb); }
        else { const dgc_batch b = hb->view(); c = dgc_submit(hs[(size_t)g], &b); }
        DGH_RANGE_POP();
        if (c) { rc = c; err = std::string("dgc_submit: ") + dgc_last_error(hs[(size_t)g]); feed.recycle(std::move(hb)); continue; }
        on_gpu[(size_t)g]++;
        stats.batches++;
        inflight.push_back(std::move(hb));
    }
    while (!inflight.empty()) collect_oldest();
    fflush(out);
    destroy_all();
    return rc;
}

int run_text_pipeline(const ProgramOpts& opts, FILE* out, RunStats& stats, std::string& err) {
    const auto t0 = std::chrono::steady_clock::now();
    if (dgc_device_count() <= 0) {
        err = "no CUDA device: the consensus path runs on the GPU only (there is no CPU fallback)";
        return DGC_E_NODEVICE;
    }
    // the reader and the packers start at once: scanning the text overlaps the CUDA context and handle creation
    Front fr(opts, true, stats);
    const int G = opts.gpus > 0 ? opts.gpus : std::max(1, dgc_device_count());
    fr.start((size_t)G * DGC_MAX_INFLIGHT + 4);
    BatchFeed feed;
    feed.take = [&](uint64_t seq) { return fr.outbox.take(seq); };
    feed.recycle = [&](std::unique_ptr<HostBatch> hb) { fr.recycle(std::move(hb)); };
    feed.failed = [&] { return fr.rc != 0; };
    int rc = dispatch_batches(opts, feed, out, stats, err, false);
    fr.stop();
    if (!rc && fr.rc) { rc = fr.rc; err = fr.err; }
    stats.pin_seconds = g_pin_seconds;
    stats.seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    return rc;
}

}  // namespace dgh
