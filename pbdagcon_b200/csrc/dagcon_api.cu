// dagcon_api.cu -- host side of libdagcon_b200.so: the C ABI of include/dagcon_b200.h and the
// batched dispatcher that replaces the reference's Reader -> BoundedBuffer -> N x Consensus ->
// BoundedBuffer -> Writer thread pipeline (src/cpp/main.cpp:37-176, 248-274;
// src/cpp/dazcon.cpp:38-120, 211-230; src/cpp/BoundedBuffer.hpp:18-50).
//
// One handle drives one GPU.  It owns NSLOT pipeline slots, each with its own CUDA stream,
// device buffers, pinned result buffers and a worker thread, so that the H2D copy of batch k+1
// overlaps the kernels of batch k.  There is no CPU implementation of any stage in here.
#include "../../include/dagcon_b200.h"
#include "dagcon_kernels.cuh"
#include "daz_realign.cuh"
#include "pre_align.cuh"
#include <nvtx3/nvToolsExt.h>

#include <algorithm>
#include <chrono>
#include <cstdarg>
#include <condition_variable>
#include <deque>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace {

thread_local std::string g_create_error;

struct DevBuf {
    void* p = nullptr;
    size_t cap = 0;
    // (A stream-ordered variant -- cudaMallocAsync / cudaFreeAsync on the slot's stream, to keep a growing buffer's
    // cudaFree from waiting behind the other slots' sweeps -- was measured in round 2 and rejected: the pool's own
    // growth made the CLI's first batches slower, 2000 targets in 2.1-3.6 s against 1.0 s; profiles/r02_cli_alloc.txt.)
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { cudaGetLastError(); e = cudaMalloc(&p, bytes + 256); want = bytes + 256; }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    // Grow without losing the first `live` bytes.  `st` is the stream whose kernels wrote those bytes: the copy is
    // enqueued behind them on that stream and has finished before the old buffer is freed (the slot streams are
    // cudaStreamNonBlocking, so a copy on the legacy default stream would NOT be ordered after them).
    cudaError_t ensure_keep(size_t bytes, size_t live, cudaStream_t st) {
        if (bytes <= cap) return cudaSuccess;
        void* np = nullptr;
        const size_t want = bytes + bytes / 4 + 256;
        cudaError_t e = cudaMalloc(&np, want);
        if (e != cudaSuccess) return e;
        if (p && live) e = cudaMemcpyAsync(np, p, std::min(live, cap), cudaMemcpyDeviceToDevice, st);
        if (e == cudaSuccess) e = cudaStreamSynchronize(st);
        if (p) cudaFree(p);
        p = np; cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return (T*)p; }
};
struct HostBuf {  // pinned
    void* p = nullptr;
    size_t cap = 0;
    std::vector<void*> old;   // outgrown blocks: cudaFreeHost waits for the whole device, so they are kept until release()
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) old.push_back(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 2 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() {
        if (p) cudaFreeHost(p);
        for (void* q : old) cudaFreeHost(q);
        old.clear();
        p = nullptr; cap = 0;
    }
    template <class T> T* as() const { return (T*)p; }
};

}  // namespace

// A batch resident in HBM: the dgc_batch arrays plus the derived normalisation slots.
struct dgc_dbatch {
    uint32_t T = 0, A = 0;
    bool has_bb = false;
    bool packed4 = false;  // DGC_BATCH_PACKED4: the strings are 4-bit codes
    uint64_t arena_bytes = 0, norm_cols = 0, raw_cols = 0;
    DevBuf tgt_tlen, tgt_aln_begin, tgt_bb_off, aln_start, aln_len, aln_q_off, aln_t_off, aln_norm_off, arena;
    std::vector<uint32_t> h_tlen, h_aln_begin;  // host copies the planner needs
    void release() {
        tgt_tlen.release(); tgt_aln_begin.release(); tgt_bb_off.release(); aln_start.release(); aln_len.release();
        aln_q_off.release(); aln_t_off.release(); aln_norm_off.release(); arena.release();
    }
};

namespace {

struct EventPool {
    std::vector<cudaEvent_t> ev;
    size_t used = 0;
    cudaEvent_t get() {
        if (used == ev.size()) { cudaEvent_t e; cudaEventCreate(&e); ev.push_back(e); }
        return ev[used++];
    }
    void reset() { used = 0; }
    void destroy() { for (auto e : ev) cudaEventDestroy(e); ev.clear(); }
};

enum Phase { PH_H2D, PH_NORM, PH_COUNT, PH_BUILD, PH_MERGE, PH_PATH, PH_COMPACT, PH_D2H, PH_DAZ, PH_N };

struct Slot {
    dgc_handle* h = nullptr;
    int index = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;  // staged uploads: the arena pieces travel here while s.stream computes
    dgc_dbatch db;  // device copy of a submitted batch
    HostBuf h_norm_off;
    // per-alignment work arrays
    DevBuf o_off, o_len, o_start, o_flags, o_nI, o_nE, o_tkept, norm;
    // per-target arrays
    DevBuf t_status, t_V, t_E, t_cols, t_ne, t_nv, t_cons_len, t_nfrag, t_cons_coff, t_frag_coff, totals;
    DevBuf rec_off, pos_off, pool_cap, t_pool_top, cons_off, frag_off, frag_cap, gs_off, gs_cap, order;
    HostBuf hp_order;     // sweep order per chunk: longest graphs first
    HostBuf hp_plan_in;   // status, V, E, cols read back after the plan kernel
    HostBuf hp_plan_out;  // the offset arrays
    // chunk arenas
    DevBuf rec, cons, frags, gscratch, gacc;
    // compacted outputs
    DevBuf out_cons, out_frags;
    HostBuf r_status, r_cons_coff, r_frag_coff, r_totals, r_cons, r_frags, r_nfrag;
    std::vector<uint32_t> res_frag_begin, res_seq_len;
    std::vector<uint64_t> res_seq_off;
    std::vector<int32_t> res_r0, res_r1;
    dgc_results results{};
    dgc_timings tm{};
    EventPool events;
    std::vector<std::pair<int, std::pair<cudaEvent_t, cudaEvent_t>>> spans;
    // worker
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    enum State { IDLE, QUEUED, RUNNING, DONE, STOP } state = IDLE;
    dgc_batch pending{};  // copied at submit: only the arrays it points to must outlive the call
    dgc_daz_batch pending_daz{};
    bool is_daz = false;
    dgc_pre_batch pending_pre{};
    bool is_pre = false;
    // `-a` alignment (pre_align.cuh)
    DevBuf p_in, p_qoff, p_toff, p_qlen, p_tlen, p_strand, p_start, p_tglen, p_ids, p_tboff, p_tb, p_tbeg, p_tend, p_state;
    HostBuf ph_plan;
    // dazcon realignment (daz_realign.cuh): per-overlap inputs, per-segment work arrays, edit scripts, wave slabs
    DevBuf z_aread, z_bread, z_comp, z_abpos, z_aepos, z_bbpos, z_bepos, z_troff, z_tlen, z_segb, z_eoff, z_trace, z_tgt_aread;
    DevBuf z_seg_bb, z_seg_eoff, z_seg_cnt, z_seg_neg, z_seg_ppos, z_seg_pneg, z_entries, z_vf, z_hf, z_err;
    HostBuf zh_plan;
    uint64_t seq = 0;     // submit sequence number of `pending`
    uint64_t hint[5] = {0, 0, 0, 0, 0};  // arena needs (units, positions, cons, frags, scratch) of the last staged batch that fit one chunk
    int rc = 0;
    std::string err;
};

}  // namespace

#define NSLOT (DGC_MAX_INFLIGHT + 1)
struct dgc_handle {
    dgc_opts opts{};
    int device = 0;
    int sm_count = 148;
    size_t build_smem_max = 0;
    // staged uploads (run_staged): piece size and the smallest arena worth staging; DGC_STAGE_PIECE_BYTES /
    // DGC_STAGE_MIN_BYTES override them (tests stage tiny batches)
    uint64_t stage_piece_bytes = 384ull << 20, stage_min_bytes = 64ull << 20;
    std::string err;
    Slot slots[NSLOT];
    uint64_t submitted = 0, collected = 0;
    dgc_timings last_tm{};
    // H2D copies of submitted batches run one at a time, in submit order: batch n's kernels start as soon as its own
    // bytes have landed while batch n+1 is on the wire (concurrent copies would share the link and all land late)
    std::mutex h2d_mu;
    std::condition_variable h2d_cv;
    uint64_t h2d_turn = 0;
    // slots in flight, in submit order; a submit takes the lowest free slot so that a caller with one batch in flight
    // keeps reusing one slot's arenas instead of growing all of them
    std::mutex q_mu;
    std::deque<int> inflight;
    bool busy[NSLOT] = {};
    // dazcon: the read database, resident (dgc_daz_set_db)
    DevBuf daz_bases, daz_boff, daz_rlen;
    std::vector<uint32_t> daz_h_rlen;
    uint32_t daz_n_reads = 0;
};

namespace {

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess) {                                                                  \
            s.err = std::string(#call) + ": " + cudaGetErrorString(e__);                           \
            return DGC_E_CUDA;                                                                     \
        }                                                                                          \
    } while (0)

// DGC_TRACE=1: host-side timestamps of a slot's work on stderr (where a batch's wall time goes besides its kernels)
static const bool g_trace = getenv("DGC_TRACE") && getenv("DGC_TRACE")[0] == '1';
static const auto g_t0 = std::chrono::steady_clock::now();
static void trace(const Slot& s, const char* what) {
    nvtxMarkA(what);   // an attached timeline tool sees the same milestones
    if (!g_trace) return;
    fprintf(stderr, "dgc-trace %.4f s slot %d seq %llu: %s\n", std::chrono::duration<double>(std::chrono::steady_clock::now() - g_t0).count(),
            s.index, (unsigned long long)s.seq, what);
}

void span_begin(Slot& s, int ph) {
    cudaEvent_t a = s.events.get(), b = s.events.get();
    cudaEventRecord(a, s.stream);
    s.spans.push_back({ph, {a, b}});
}
void span_end(Slot& s) { cudaEventRecord(s.spans.back().second.second, s.stream); }

int validate_batch(const dgc_batch* b, std::string& err) {
    if (!b) { err = "null batch"; return DGC_E_INVALID; }
    if (b->n_targets && (!b->tgt_tlen || !b->tgt_aln_begin)) { err = "null target arrays"; return DGC_E_INVALID; }
    if (b->n_alns && (!b->aln_start || !b->aln_len || !b->aln_q_off || !b->aln_t_off || !b->arena)) {
        err = "null alignment arrays"; return DGC_E_INVALID;
    }
    if (b->n_targets) {
        if (b->tgt_aln_begin[0] != 0 || b->tgt_aln_begin[b->n_targets] != b->n_alns) {
            err = "tgt_aln_begin is not a CSR over n_alns"; return DGC_E_INVALID;
        }
        for (uint32_t t = 0; t < b->n_targets; t++) {
            if (b->tgt_aln_begin[t + 1] < b->tgt_aln_begin[t]) { err = "tgt_aln_begin not monotone"; return DGC_E_INVALID; }
            if (b->tgt_tlen[t] > 0x3FFFFFFFu) { err = "tlen too large"; return DGC_E_INVALID; }
            if (b->tgt_bb_off && (b->tgt_bb_off[t] > b->arena_bytes || b->tgt_tlen[t] > b->arena_bytes - b->tgt_bb_off[t])) { err = "backbone outside arena"; return DGC_E_INVALID; }
        }
    } else if (b->n_alns) { err = "alignments without targets"; return DGC_E_INVALID; }
    if (b->flags & ~DGC_BATCH_PACKED4) { err = "unknown dgc_batch.flags"; return DGC_E_INVALID; }
    for (uint32_t k = 0; k < b->n_alns; k++) {
        uint64_t n = (b->flags & DGC_BATCH_PACKED4) ? ((uint64_t)b->aln_len[k] + 1) / 2 : b->aln_len[k];
        if (b->aln_q_off[k] > b->arena_bytes || n > b->arena_bytes - b->aln_q_off[k] || b->aln_t_off[k] > b->arena_bytes ||
            n > b->arena_bytes - b->aln_t_off[k]) {
            err = "alignment string outside arena"; return DGC_E_INVALID;
        }
    }
    return DGC_OK;
}

// H2D of a batch into db (async on s.stream; host arrays must stay valid until the stream syncs)
int upload_batch(Slot& s, const dgc_batch* b, dgc_dbatch& db, bool timed, bool with_arena = true) {
    db.T = b->n_targets; db.A = b->n_alns; db.has_bb = b->tgt_bb_off != nullptr; db.arena_bytes = b->arena_bytes;
    db.packed4 = (b->flags & DGC_BATCH_PACKED4) != 0;
    const size_t T = db.T, A = db.A;
    CU(db.tgt_tlen.ensure(4 * (T + 1))); CU(db.tgt_aln_begin.ensure(4 * (T + 1)));
    if (db.has_bb) CU(db.tgt_bb_off.ensure(8 * (T + 1)));
    CU(db.aln_start.ensure(4 * (A + 1))); CU(db.aln_len.ensure(4 * (A + 1)));
    CU(db.aln_q_off.ensure(8 * (A + 1))); CU(db.aln_t_off.ensure(8 * (A + 1))); CU(db.aln_norm_off.ensure(8 * (A + 1)));
    CU(db.arena.ensure(b->arena_bytes + 64));
    CU(s.h_norm_off.ensure(8 * (A + 1)));
    uint64_t* no = s.h_norm_off.as<uint64_t>();
    uint64_t acc = 0, raw = 0;
    for (size_t k = 0; k < A; k++) {  // a normalised alignment has at most 2*len columns; keep slots 8-column aligned
        no[k] = acc;
        raw += b->aln_len[k];
        acc += (2ull * b->aln_len[k] + 7) & ~7ull;
    }
    db.norm_cols = acc; db.raw_cols = raw;
    db.h_tlen.assign(b->tgt_tlen, b->tgt_tlen + T);
    db.h_aln_begin.assign(b->tgt_aln_begin, b->tgt_aln_begin + T + 1);
    if (timed) span_begin(s, PH_H2D);
    cudaStream_t st = s.stream;
    if (T) {
        CU(cudaMemcpyAsync(db.tgt_tlen.p, b->tgt_tlen, 4 * T, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.tgt_aln_begin.p, b->tgt_aln_begin, 4 * (T + 1), cudaMemcpyHostToDevice, st));
        if (db.has_bb) CU(cudaMemcpyAsync(db.tgt_bb_off.p, b->tgt_bb_off, 8 * T, cudaMemcpyHostToDevice, st));
    }
    if (A) {
        CU(cudaMemcpyAsync(db.aln_start.p, b->aln_start, 4 * A, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_len.p, b->aln_len, 4 * A, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_q_off.p, b->aln_q_off, 8 * A, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_t_off.p, b->aln_t_off, 8 * A, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_norm_off.p, no, 8 * A, cudaMemcpyHostToDevice, st));
    }
    if (b->arena_bytes && with_arena) CU(cudaMemcpyAsync(db.arena.p, b->arena, b->arena_bytes, cudaMemcpyHostToDevice, st));
    if (timed) span_end(s);
    return DGC_OK;
}

int ensure_alignment_arrays(Slot& s, const dgc_dbatch& db) {
    const size_t A = db.A;
    CU(s.o_off.ensure(8 * (A + 1))); CU(s.o_len.ensure(4 * (A + 1))); CU(s.o_start.ensure(4 * (A + 1)));
    CU(s.o_flags.ensure(4 * (A + 1))); CU(s.o_nI.ensure(4 * (A + 1))); CU(s.o_nE.ensure(4 * (A + 1)));
    CU(s.o_tkept.ensure(4 * (A + 1)));
    CU(s.norm.ensure(2 * (db.norm_cols + 8)));
    return DGC_OK;
}

// K1 for alignments [k0, k1)
int launch_normalize(Slot& s, const dgc_dbatch& db, uint32_t k0, uint32_t k1) {
    const dgc_opts& o = s.h->opts;
    NormArgs na;
    na.arena = db.arena.as<u8>(); na.aln_start = db.aln_start.as<u32>(); na.aln_len = db.aln_len.as<u32>();
    na.aln_q_off = db.aln_q_off.as<u64>(); na.aln_t_off = db.aln_t_off.as<u64>(); na.aln_norm_off = db.aln_norm_off.as<u64>();
    na.n_alns = k1; na.aln_first = k0; na.min_len = o.min_len; na.trim = o.trim; na.flags = (o.flags & 0xFFFFu) | (db.packed4 ? F_PACKED4 : 0u);
    na.norm = s.norm.as<u16>(); na.o_off = s.o_off.as<u64>(); na.o_len = s.o_len.as<u32>(); na.o_start = s.o_start.as<u32>();
    na.o_flags = s.o_flags.as<u32>(); na.o_nI = s.o_nI.as<u32>(); na.o_nE = s.o_nE.as<u32>(); na.o_tkept = s.o_tkept.as<u32>();
    const uint32_t n = k1 - k0;
    if (n) {
        if (o.flags & DGC_F_NO_NORMALIZE) {
            na.retry_only = 0;
            k_normalize_trim<<<(n + 127) / 128, 128, 0, s.stream>>>(na);
            s.tm.launches++;
        } else {
            na.retry_only = 1;
            k_normalize_warp<<<(n + NW_WARPS - 1) / NW_WARPS, NW_WARPS * 32, 0, s.stream>>>(na);
            k_normalize_trim<<<(n + 127) / 128, 128, 0, s.stream>>>(na);  // the few the warp kernel handed back
            s.tm.launches += 2;
        }
    }
    CU(cudaGetLastError());
    return DGC_OK;
}
int launch_normalize(Slot& s, const dgc_dbatch& db) { return launch_normalize(s, db, 0, db.A); }

// min-coverage filter and graph sizes for targets [t0, t1)
int launch_plan(Slot& s, const dgc_dbatch& db, uint32_t t0, uint32_t t1) {
    const dgc_opts& o = s.h->opts;
    PlanArgs pa;
    pa.n_targets = t1; pa.t_first = t0; pa.min_cov = o.min_cov; pa.tgt_tlen = db.tgt_tlen.as<u32>(); pa.tgt_aln_begin = db.tgt_aln_begin.as<u32>();
    pa.a_start = s.o_start.as<u32>(); pa.a_flags = s.o_flags.as<u32>(); pa.a_nI = s.o_nI.as<u32>(); pa.a_nE = s.o_nE.as<u32>();
    pa.a_tkept = s.o_tkept.as<u32>(); pa.a_len = s.o_len.as<u32>();
    pa.tgt_status = s.t_status.as<u32>(); pa.tgt_V = s.t_V.as<u32>(); pa.tgt_E = s.t_E.as<u32>(); pa.tgt_cols = s.t_cols.as<u64>();
    if (t1 > t0) k_target_plan<<<(t1 - t0 + 127) / 128, 128, 0, s.stream>>>(pa);
    s.tm.launches++;
    CU(cudaGetLastError());
    return DGC_OK;
}

struct Chunk { uint32_t first, count; uint64_t nUnits, nPos, nCons, nFrag, nGs; uint32_t maxL2; };
// per-position accumulators of the build (gacc): sized for the wide (u16 list length) variant, 11 bytes per position;
// every target's share is rounded up to 16 positions so that the u16 arrays inside stay aligned whatever L is
constexpr uint64_t ACC_BYTES = 11;
inline uint64_t pos16(uint64_t L2) { return (L2 + 15) & ~15ull; }
uint64_t chunk_bytes(const Chunk& c) { return 128 * c.nUnits + ACC_BYTES * c.nPos + c.nCons + 8 * c.nFrag + 4 * c.nGs; }

// Pool of list blocks behind a target's records, in 8-byte words (a multiple of 16 keeps every arena 128-byte
// aligned): enter's out-list and exit's in-list always, the rest only when a list outgrows its record.  The normal
// size covers the coverages the record layout was shaped for; `worst` is the bound for the rebuild of a target whose
// pool ran dry (T_POOL): lists double when they grow, so a backbone in-list costs <= 2 words and a backbone
// out-list <= 4 words per edge that can reach a backbone vertex (Bin = sum of nE - nI over the alignments), and the
// merges re-home at most a few slots per edge.
bool g_tiny_pool = false;  // DGC_TEST_TINY_POOL: first-pass pools hold the enter/exit lists only, so that the tests reach the rebuild
uint64_t pool_words(uint64_t V, uint64_t E, uint64_t L, uint32_t A, bool worst) {
    if (!worst && g_tiny_pool) return (4ull * (A + 8) + 15) & ~15ull;
    if (!worst) return (4ull * (A + 8) + E / 2 + 256 + 15) & ~15ull;
    const uint64_t sE = E > L + 1 + A + 16 ? E - (L + 1) - A - 16 : 0, sI = V > L + 2 ? V - (L + 2) : 0;
    const uint64_t bin = sE > sI ? sE - sI : 0;
    return (4ull * (A + 8) + 256 + 2 * bin + 4 * (bin + A) + 8 * E + 15) & ~15ull;
}

// K2 for targets [first, first + count) of the current chunk (their arena offsets are on the device already)
int launch_build(Slot& s, dgc_dbatch& db, uint32_t first, uint32_t count, uint32_t maxL2) {
    dgc_handle* h = s.h;
    cudaStream_t st = s.stream;
    Build2Args ba;
    ba.first = first; ba.tgt_tlen = db.tgt_tlen.as<u32>(); ba.tgt_aln_begin = db.tgt_aln_begin.as<u32>();
    ba.tgt_bb_off = db.has_bb ? db.tgt_bb_off.as<u64>() : nullptr; ba.arena = db.arena.as<u8>();
    ba.tgt_status = s.t_status.as<u32>(); ba.rec_off = s.rec_off.as<u64>(); ba.tgt_V = s.t_V.as<u32>();
    ba.pool_cap = s.pool_cap.as<u32>(); ba.acc_off = s.pos_off.as<u64>();
    ba.norm = s.norm.as<u16>(); ba.a_off = s.o_off.as<u64>(); ba.a_len = s.o_len.as<u32>();
    ba.a_start = s.o_start.as<u32>(); ba.a_flags = s.o_flags.as<u32>(); ba.a_tkept = s.o_tkept.as<u32>();
    ba.rec = s.rec.as<u8>();
    ba.tgt_status_out = s.t_status.as<u32>(); ba.tgt_nv = s.t_nv.as<u32>(); ba.tgt_pool_top = s.t_pool_top.as<u32>();
    // List lengths per position are bytes unless a target of the launch has so many alignments that a backbone vertex
    // could collect 254 edges (A + 1 >= 254): then the u16 variant runs (11 instead of 9 bytes per position).
    uint32_t maxA = 0;
    for (uint32_t t = first; t < first + count; t++) maxA = std::max(maxA, db.h_aln_begin[t + 1] - db.h_aln_begin[t]);
    const bool wide = maxA + 2 >= acc_max_deg<u8>();
    const uint64_t bpp = wide ? acc_bytes_per_pos<u16>() : acc_bytes_per_pos<u8>();
    // per-position running sums in shared memory when the longest backbone of the chunk fits
    // (a launch whose longest backbone does not fit asks for no dynamic shared memory at all: the accumulators of all
    // its targets go to `gacc`, which stays L2-resident, and the blocks per SM are no longer limited by shared memory)
    ba.smem_positions = bpp * maxL2 <= h->build_smem_max ? maxL2 : 0;
    ba.gacc = s.gacc.as<u8>();
    const size_t build_smem = (bpp * ba.smem_positions + 15) & ~15ull;
    span_begin(s, PH_BUILD);
    if (wide) k_build_graph2<u16><<<count, BT, build_smem, st>>>(ba);
    else k_build_graph2<u8><<<count, BT, build_smem, st>>>(ba);
    s.tm.launches++;
    CU(cudaGetLastError());
    span_end(s);
    return DGC_OK;
}

// K3, K4 and the output compaction for one chunk.  stop_after: 1 merge only, 2 everything.
int launch_serial(Slot& s, dgc_dbatch& db, const Chunk& c, int stop_after) {
    const dgc_opts& o = s.h->opts;
    cudaStream_t st = s.stream;

    Serial2Args sa;
    sa.first = c.first; sa.count = c.count;
    sa.flags = o.flags; sa.min_weight = o.min_cov; sa.min_len = o.min_len;
    sa.tgt_tlen = db.tgt_tlen.as<u32>(); sa.tgt_aln_begin = db.tgt_aln_begin.as<u32>(); sa.tgt_status = s.t_status.as<u32>();
    sa.rec_off = s.rec_off.as<u64>(); sa.tgt_V = s.t_V.as<u32>();
    sa.pool_cap = s.pool_cap.as<u32>(); sa.tgt_pool_top = s.t_pool_top.as<u32>(); sa.tgt_nv = s.t_nv.as<u32>();
    sa.rec = s.rec.as<u8>();
    sa.gscratch = s.gscratch.as<u32>(); sa.gscratch_off = s.gs_off.as<u64>(); sa.gscratch_cap = s.gs_cap.as<u32>();
    sa.cons = s.cons.as<u8>(); sa.cons_off = s.cons_off.as<u64>(); sa.frags = s.frags.as<int>(); sa.frag_off = s.frag_off.as<u64>();
    sa.frag_cap = s.frag_cap.as<u32>(); sa.tgt_cons_len = s.t_cons_len.as<u32>(); sa.tgt_nfrag = s.t_nfrag.as<u32>();
    // Longest graphs first: a launch of more targets than resident warps runs in waves, and a long chain that starts
    // late is the tail of the whole launch.  The four warps of a block also get graphs of similar size this way.
    {
        const size_t T = db.T;
        CU(s.hp_order.ensure(4 * (T + 1))); CU(s.order.ensure(4 * (T + 1)));
        const uint32_t* hp_V = s.hp_plan_in.as<uint32_t>() + (T + 1);
        uint32_t* ord = s.hp_order.as<uint32_t>() + c.first;
        for (uint32_t i = 0; i < c.count; i++) ord[i] = c.first + i;
        std::stable_sort(ord, ord + c.count, [&](uint32_t x, uint32_t y) { return hp_V[x] > hp_V[y]; });
        CU(cudaMemcpyAsync(s.order.as<u32>() + c.first, ord, 4ull * c.count, cudaMemcpyHostToDevice, st));
        sa.order = s.order.as<u32>();
    }
    const uint32_t warps_per_block = WARP_KERNEL_THREADS / 32;   // one warp per target
    const uint32_t grid = (c.count + warps_per_block - 1) / warps_per_block;
    if (stop_after >= 2 && !(o.flags & (DGC_F_NO_MERGE | DGC_F_SPLIT_SWEEPS))) {
        span_begin(s, PH_MERGE);  // both sweeps in one launch; reported as ms_merge
        k_merge_best2<<<grid, WARP_KERNEL_THREADS, 0, st>>>(sa);
        s.tm.launches++;
        CU(cudaGetLastError());
        span_end(s);
    } else {
    if (!(o.flags & DGC_F_NO_MERGE)) {
        span_begin(s, PH_MERGE);
        k_merge_nodes2<<<grid, WARP_KERNEL_THREADS, 0, st>>>(sa);
        s.tm.launches++;
        CU(cudaGetLastError());
        span_end(s);
    }
    if (stop_after < 2) return DGC_OK;
    span_begin(s, PH_PATH);
    k_best_path2<<<grid, WARP_KERNEL_THREADS, 0, st>>>(sa);
    s.tm.launches++;
    CU(cudaGetLastError());
    span_end(s);
    }
    span_begin(s, PH_COMPACT);
    // scan over the chunk's targets, chained through `totals` so chunks append to one output
    k_scan_outputs<<<1, 1024, 0, st>>>(c.count, s.t_status.as<u32>() + c.first, s.t_cons_len.as<u32>() + c.first,
                                       s.t_nfrag.as<u32>() + c.first, s.t_cons_coff.as<u64>() + c.first,
                                       s.t_frag_coff.as<u32>() + c.first, s.totals.as<u64>());
    k_compact<<<c.count, 128, 0, st>>>(c.count, s.t_status.as<u32>() + c.first, s.t_cons_len.as<u32>() + c.first,
                                       s.t_nfrag.as<u32>() + c.first, s.cons.as<u8>(), s.cons_off.as<u64>() + c.first,
                                       s.frags.as<int>(), s.frag_off.as<u64>() + c.first, s.t_cons_coff.as<u64>() + c.first,
                                       s.t_frag_coff.as<u32>() + c.first, s.out_cons.as<u8>(), s.out_frags.as<int>());
    s.tm.launches += 2;
    CU(cudaGetLastError());
    span_end(s);
    return DGC_OK;
}

// D2H of the compacted results and the CnsResult records in target order
int fetch_results(Slot& s, dgc_dbatch& db) {
    const size_t T = db.T;
    cudaStream_t st = s.stream;
    // ---- D2H of the compacted results
    span_begin(s, PH_D2H);
    CU(s.r_status.ensure(4 * (T + 1))); CU(s.r_cons_coff.ensure(8 * (T + 1))); CU(s.r_frag_coff.ensure(4 * (T + 1)));
    CU(s.r_totals.ensure(64)); CU(s.r_nfrag.ensure(4 * (T + 1)));
    CU(cudaMemcpyAsync(s.r_nfrag.p, s.t_nfrag.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(s.r_status.p, s.t_status.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(s.r_cons_coff.p, s.t_cons_coff.p, 8 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(s.r_frag_coff.p, s.t_frag_coff.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(s.r_totals.p, s.totals.p, 32, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint64_t tot_cons = s.r_totals.as<uint64_t>()[0], tot_frag = s.r_totals.as<uint64_t>()[1];
    CU(s.r_cons.ensure(tot_cons + 16)); CU(s.r_frags.ensure(8 * tot_frag + 16));
    if (tot_cons) CU(cudaMemcpyAsync(s.r_cons.p, s.out_cons.p, tot_cons, cudaMemcpyDeviceToHost, st));
    if (tot_frag) CU(cudaMemcpyAsync(s.r_frags.p, s.out_frags.p, 8 * tot_frag, cudaMemcpyDeviceToHost, st));
    span_end(s);
    CU(cudaStreamSynchronize(st));

    // ---- CnsResult records in target order
    uint32_t* rs = s.r_status.as<uint32_t>();
    for (size_t t = 0; t < T; t++)
        if (rs[t] == T_POOL) rs[t] = T_INTERNAL;  // not even the worst-case pool was enough (or addressable)
    const uint64_t* cc = s.r_cons_coff.as<uint64_t>();
    const uint32_t* fc = s.r_frag_coff.as<uint32_t>();
    const int32_t* fr = s.r_frags.as<int32_t>();
    const uint32_t* rn = s.r_nfrag.as<uint32_t>();
    s.res_r0.reserve(tot_frag); s.res_r1.reserve(tot_frag); s.res_seq_off.reserve(tot_frag); s.res_seq_len.reserve(tot_frag);
    for (size_t t = 0; t < T; t++) {
        s.res_frag_begin[t] = (uint32_t)s.res_r0.size();
        if (rs[t] != T_OK) continue;
        const uint32_t f0 = fc[t], nf = rn[t];
        for (uint32_t i = 0; i < nf; i++) {
            const int32_t r0 = fr[2 * (size_t)(f0 + i)], r1 = fr[2 * (size_t)(f0 + i) + 1];
            s.res_r0.push_back(r0); s.res_r1.push_back(r1);
            s.res_seq_off.push_back(cc[t] + (uint64_t)r0);
            s.res_seq_len.push_back((uint32_t)(r1 - r0));
        }
    }
    s.res_frag_begin[T] = (uint32_t)s.res_r0.size();
    s.results.n_frags = (uint32_t)s.res_r0.size();
    s.results.tgt_status = rs;
    s.results.tgt_frag_begin = s.res_frag_begin.data();
    s.results.frag_range0 = s.res_r0.data(); s.results.frag_range1 = s.res_r1.data();
    s.results.frag_seq_off = s.res_seq_off.data(); s.results.frag_seq_len = s.res_seq_len.data();
    s.results.seq_arena = s.r_cons.as<uint8_t>(); s.results.seq_bytes = tot_cons;
    return DGC_OK;
}


// Second pass for the targets whose list pool ran dry (status T_POOL after the sweeps of their chunk): each run of
// such targets is planned again with the worst-case pool, rebuilt and swept.  By now the arenas hold nothing that is
// still needed (every chunk's results have been compacted), so they may be re-allocated.  hp_V / hp_E / the p_* offset
// arrays are the pinned planner arrays of this run.
int retry_pool_failures(Slot& s, dgc_dbatch& db, int stop_after) {
    dgc_handle* h = s.h;
    const dgc_opts& o = h->opts;
    const size_t T = db.T;
    cudaStream_t st = s.stream;
    uint32_t* hp_status = s.hp_plan_in.as<uint32_t>();
    const uint32_t* hp_V = hp_status + (T + 1);
    const uint32_t* hp_E = hp_V + (T + 1);
    uint64_t* p_rec = s.hp_plan_out.as<uint64_t>();
    uint64_t* p_pos = p_rec + (T + 1);
    uint64_t* p_cons = p_pos + (T + 1);
    uint64_t* p_frag = p_cons + (T + 1);
    uint64_t* p_gs = p_frag + (T + 1);
    uint32_t* p_fcap = (uint32_t*)(p_gs + (T + 1));
    uint32_t* p_gcap = p_fcap + (T + 1);
    uint32_t* p_pcap = p_gcap + (T + 1);
    CU(cudaMemcpyAsync(hp_status, s.t_status.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    const uint64_t budget = o.scratch_bytes ? o.scratch_bytes : (112ull << 30);
    size_t t = 0;
    while (t < T) {
        if (hp_status[t] != T_POOL) { t++; continue; }
        // one chunk: the longest run of failed targets from t that fits the budget
        Chunk c{(uint32_t)t, 0, 0, 0, 0, 0, 0, 0};
        while (t < T && hp_status[t] == T_POOL) {
            const uint64_t V = hp_V[t], E = hp_E[t], L2 = (uint64_t)db.h_tlen[t] + 2;
            const uint32_t A = db.h_aln_begin[t + 1] - db.h_aln_begin[t];
            const uint64_t pcap = pool_words(V, E, db.h_tlen[t], A, true);
            const uint64_t units = L2 + V + pcap / 16;
            if (16 * units > 0xFFFFFFF0ull || pcap > 0xFFFFFFF0ull) break;  // stays T_POOL -> reported as DGC_T_INTERNAL
            Chunk nxt = c;
            nxt.nUnits += units; nxt.nPos += pos16(L2); nxt.nCons += (V + 15) & ~15ull; nxt.nFrag += p_fcap[t]; nxt.nGs += p_gcap[t]; nxt.count++;
            if (chunk_bytes(nxt) > budget) break;
            p_rec[t] = c.nUnits; p_pos[t] = c.nPos; p_cons[t] = c.nCons; p_frag[t] = c.nFrag; p_gs[t] = c.nGs;
            p_pcap[t] = (uint32_t)pcap;
            nxt.maxL2 = std::max(c.maxL2, (uint32_t)L2);
            c = nxt;
            t++;
        }
        if (c.count == 0) { t++; continue; }  // a single target beyond the budget or the 32 GiB a list pointer can address
        const size_t a = c.first, n = c.count;
        CU(s.rec.ensure(128 * c.nUnits + 256));
        CU(s.cons.ensure(c.nCons + 64)); CU(s.frags.ensure(8 * c.nFrag + 64)); CU(s.gscratch.ensure(4 * c.nGs + 64));
        if (ACC_BYTES * c.maxL2 > h->build_smem_max) CU(s.gacc.ensure(ACC_BYTES * c.nPos + 64));
        CU(cudaMemcpyAsync(s.rec_off.as<u64>() + a, p_rec + a, 8 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.pos_off.as<u64>() + a, p_pos + a, 8 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.cons_off.as<u64>() + a, p_cons + a, 8 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.frag_off.as<u64>() + a, p_frag + a, 8 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.gs_off.as<u64>() + a, p_gs + a, 8 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.pool_cap.as<u32>() + a, p_pcap + a, 4 * n, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(s.t_status.as<u32>() + a, 0, 4 * n, st));  // T_OK
        int rc = launch_build(s, db, c.first, c.count, c.maxL2);
        if (rc == DGC_OK && stop_after >= 1) rc = launch_serial(s, db, c, stop_after);
        if (rc) return rc;
        s.tm.arena_bytes += chunk_bytes(c);
        s.tm.chunks++;
    }
    return DGC_OK;
}

// The whole hot path for a resident batch.  stop_after: 0 build, 1 merge, 2 everything.
int run_pipeline(Slot& s, dgc_dbatch& db, int stop_after, uint32_t only_target) {
    dgc_handle* h = s.h;
    const dgc_opts& o = h->opts;
    const size_t T = db.T;
    cudaStream_t st = s.stream;
    s.results = dgc_results{};
    s.results.n_targets = db.T;
    s.res_frag_begin.assign(T + 1, 0);
    s.res_r0.clear(); s.res_r1.clear(); s.res_seq_off.clear(); s.res_seq_len.clear();
    if (T == 0) {
        s.results.tgt_frag_begin = s.res_frag_begin.data();
        return DGC_OK;
    }
    trace(s, "run_pipeline: begin");
    int rc = ensure_alignment_arrays(s, db);
    if (rc) return rc;
    for (DevBuf* b : {&s.t_status, &s.t_V, &s.t_E, &s.t_ne, &s.t_nv, &s.t_cons_len, &s.t_nfrag, &s.frag_cap, &s.gs_cap, &s.t_frag_coff,
                      &s.pool_cap, &s.t_pool_top})
        CU(b->ensure(4 * (T + 1)));
    for (DevBuf* b : {&s.t_cols, &s.t_cons_coff, &s.rec_off, &s.pos_off, &s.cons_off, &s.frag_off, &s.gs_off})
        CU(b->ensure(8 * (T + 1)));
    CU(s.totals.ensure(64));
    CU(s.hp_plan_in.ensure(20 * (T + 1) + 16));
    CU(s.hp_plan_out.ensure(64 * (T + 1)));
    trace(s, "run_pipeline: small buffers + norm allocated");

    span_begin(s, PH_NORM);
    rc = launch_normalize(s, db);
    if (rc) return rc;
    span_end(s);

    span_begin(s, PH_COUNT);
    rc = launch_plan(s, db, 0, db.T);
    if (rc) return rc;
    span_end(s);

    uint32_t* hp_status = s.hp_plan_in.as<uint32_t>();
    uint32_t* hp_V = hp_status + (T + 1);
    uint32_t* hp_E = hp_V + (T + 1);
    uint64_t* hp_cols = (uint64_t*)(hp_E + (T + 1) + ((T + 1) & 1));
    CU(cudaMemcpyAsync(hp_status, s.t_status.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hp_V, s.t_V.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hp_E, s.t_E.p, 4 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaMemcpyAsync(hp_cols, s.t_cols.p, 8 * T, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    trace(s, "run_pipeline: normalise + plan done");

    // ---- host planning: arena offsets per target (128-byte units), chunks that fit the scratch budget
    uint64_t* p_rec = s.hp_plan_out.as<uint64_t>();
    uint64_t* p_pos = p_rec + (T + 1);
    uint64_t* p_cons = p_pos + (T + 1);
    uint64_t* p_frag = p_cons + (T + 1);
    uint64_t* p_gs = p_frag + (T + 1);
    uint32_t* p_fcap = (uint32_t*)(p_gs + (T + 1));
    uint32_t* p_gcap = p_fcap + (T + 1);
    uint32_t* p_pcap = p_gcap + (T + 1);
    uint64_t budget = o.scratch_bytes ? o.scratch_bytes : (112ull << 30);
    std::vector<Chunk> chunks;
    Chunk cur{0, 0, 0, 0, 0, 0, 0, 0};
    uint64_t out_cons_cap = 0, out_frag_cap = 0;
    for (size_t t = 0; t < T; t++) {
        const bool ok = hp_status[t] == T_OK && (only_target == 0xFFFFFFFFu || only_target == t);
        if (!ok && hp_status[t] == T_OK) hp_status[t] = T_SKIPPED;  // debug single-target run
        const uint64_t V = ok ? hp_V[t] : 0, E = ok ? hp_E[t] : 0, L2 = ok ? (uint64_t)db.h_tlen[t] + 2 : 0;
        const uint32_t A = db.h_aln_begin[t + 1] - db.h_aln_begin[t];
        const uint64_t pcap = ok ? pool_words(V, E, db.h_tlen[t], A, false) : 0;
        const uint64_t units = ok ? L2 + V + pcap / 16 : 0;  // backbone records take two units
        const uint64_t fcap = ok ? V / ((uint64_t)o.min_len + 1) + 2 : 0, gcap = ok ? 8ull * (A + 2) + 64 : 0;
        const uint64_t consb = (V + 15) & ~15ull;
        Chunk one{0, 1, units, L2, consb, fcap, gcap, (uint32_t)L2};
        if (16 * units > 0xFFFFFFF0ull) { s.err = "one target's graph exceeds the 32 GiB a list pointer can address"; return DGC_E_NOMEM; }
        if (chunk_bytes(one) > budget) { s.err = "one target needs more than scratch_bytes of graph arena"; return DGC_E_NOMEM; }
        Chunk nxt = cur;
        nxt.nUnits += units; nxt.nPos += pos16(L2); nxt.nCons += consb; nxt.nFrag += fcap; nxt.nGs += gcap; nxt.count++;
        if (cur.count && (chunk_bytes(nxt) > budget || cur.count >= (1u << 20))) {
            chunks.push_back(cur);
            cur = Chunk{(uint32_t)t, 0, 0, 0, 0, 0, 0, 0};
        }
        p_rec[t] = cur.nUnits; p_pos[t] = cur.nPos; p_cons[t] = cur.nCons; p_frag[t] = cur.nFrag; p_gs[t] = cur.nGs;
        p_fcap[t] = (uint32_t)fcap; p_gcap[t] = (uint32_t)gcap; p_pcap[t] = (uint32_t)pcap;
        cur.nUnits += units; cur.nPos += pos16(L2); cur.nCons += consb; cur.nFrag += fcap; cur.nGs += gcap; cur.count++;
        cur.maxL2 = std::max(cur.maxL2, (uint32_t)L2);
        out_cons_cap += V; out_frag_cap += fcap;
        s.tm.nodes += V; s.tm.norm_cols += ok ? hp_cols[t] : 0;
    }
    if (cur.count) chunks.push_back(cur);
    if (only_target != 0xFFFFFFFFu) CU(cudaMemcpyAsync(s.t_status.p, hp_status, 4 * T, cudaMemcpyHostToDevice, st));
    uint64_t mUnits = 0, mPos = 0, mCons = 0, mFrag = 0, mGs = 0;
    bool need_gacc = false;
    for (auto& c : chunks) {
        mUnits = std::max(mUnits, c.nUnits); mPos = std::max(mPos, c.nPos);
        mCons = std::max(mCons, c.nCons); mFrag = std::max(mFrag, c.nFrag); mGs = std::max(mGs, c.nGs);
        s.tm.arena_bytes += chunk_bytes(c);
        if (ACC_BYTES * c.maxL2 > h->build_smem_max) need_gacc = true;
    }
    s.tm.chunks = (uint32_t)chunks.size();
    CU(s.rec.ensure(128 * mUnits + 256));
    CU(s.cons.ensure(mCons + 64)); CU(s.frags.ensure(8 * mFrag + 64)); CU(s.gscratch.ensure(4 * mGs + 64));
    if (need_gacc) CU(s.gacc.ensure(ACC_BYTES * mPos + 64));
    CU(s.out_cons.ensure(out_cons_cap + 64)); CU(s.out_frags.ensure(8 * out_frag_cap + 64));
    trace(s, "run_pipeline: graph arenas allocated");
    CU(cudaMemcpyAsync(s.rec_off.p, p_rec, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.pos_off.p, p_pos, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.cons_off.p, p_cons, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.frag_off.p, p_frag, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.gs_off.p, p_gs, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.frag_cap.p, p_fcap, 4 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.gs_cap.p, p_gcap, 4 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.pool_cap.p, p_pcap, 4 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemsetAsync(s.totals.p, 0, 64, st));
    CU(cudaMemsetAsync(s.t_cons_len.p, 0, 4 * (T + 1), st));
    CU(cudaMemsetAsync(s.t_nfrag.p, 0, 4 * (T + 1), st));

    for (auto& c : chunks) {
        rc = launch_build(s, db, c.first, c.count, c.maxL2);
        if (rc) return rc;
        if (stop_after == 0) continue;
        rc = launch_serial(s, db, c, stop_after);
        if (rc) return rc;
    }
    rc = retry_pool_failures(s, db, stop_after);
    if (rc) return rc;
    if (stop_after < 2) { CU(cudaStreamSynchronize(st)); return DGC_OK; }
    trace(s, "run_pipeline: kernels enqueued");
    rc = fetch_results(s, db);
    trace(s, "run_pipeline: results fetched");
    return rc;
}

// ------------------------------------------------------------------------------------------
// Staged form of the same path for a batch that arrives from the host (dgc_submit).  The alignment strings are
// the bulk of a batch (0.54 MB per 8 kb x 30x target); copying all of them before the first kernel leaves the SMs
// idle for the whole transfer (100 ms for 10 000 targets).  When the targets' strings lie in the arena in target
// order, a range of targets is a range of bytes: the arena travels in pieces on the slot's copy stream while the
// compute stream normalises, sizes and builds the targets of each piece as soon as its bytes have landed.  The
// latency-bound sweeps (mergeNodes, bestPath) then run once over every target that fits the arena, which is what
// keeps ~10 000 chains in flight.
struct Piece { uint32_t t0, t1, k0, k1; uint64_t b0, b1; cudaEvent_t copied; };

struct TurnMsg { dgc_handle* h; uint64_t next; };
void release_h2d_turn_now(dgc_handle* h, uint64_t next) {
    {
        std::lock_guard<std::mutex> lk(h->h2d_mu);
        h->h2d_turn = next;
    }
    h->h2d_cv.notify_all();
}
void CUDART_CB release_h2d_turn_cb(void* p) {  // host function on the copy stream: the batch's bytes have landed
    TurnMsg* m = (TurnMsg*)p;
    release_h2d_turn_now(m->h, m->next);
    delete m;
}

// Splits a batch into about `want` target ranges of equal byte size; false when the arena is not laid out in
// target order (then the batch is copied in one piece).
bool split_pieces(const dgc_batch* b, uint32_t want, std::vector<Piece>& out) {
    const uint32_t T = b->n_targets;
    if (want < 2 || T < 2 * want) return false;
    std::vector<uint64_t> hi(T);
    uint64_t prev_hi = 0;
    for (uint32_t t = 0; t < T; t++) {
        uint64_t lo = ~0ull, h = 0;
        if (b->tgt_bb_off) { lo = b->tgt_bb_off[t]; h = lo + b->tgt_tlen[t]; }
        for (uint32_t k = b->tgt_aln_begin[t]; k < b->tgt_aln_begin[t + 1]; k++) {
            const uint64_t n = (b->flags & DGC_BATCH_PACKED4) ? ((uint64_t)b->aln_len[k] + 1) / 2 : b->aln_len[k], q = b->aln_q_off[k], tt = b->aln_t_off[k];
            lo = std::min(lo, std::min(q, tt));
            h = std::max(h, std::max(q, tt) + n);
        }
        if (lo == ~0ull) { lo = prev_hi; h = prev_hi; }
        if (lo < prev_hi) return false;
        prev_hi = std::max(prev_hi, h);
        hi[t] = prev_hi;
    }
    const uint64_t per = b->arena_bytes / want + 1;
    Piece p{0, 0, 0, 0, 0, 0, nullptr};
    for (uint32_t t = 0; t < T; t++) {
        if (hi[t] - p.b0 >= per || t + 1 == T) {
            p.t1 = t + 1; p.k1 = b->tgt_aln_begin[t + 1];
            p.b1 = t + 1 == T ? b->arena_bytes : hi[t];
            out.push_back(p);
            p = Piece{t + 1, 0, p.k1, 0, p.b1, 0, nullptr};
        }
    }
    return out.size() >= 2;
}


int run_staged(Slot& s, const dgc_batch* b, std::vector<Piece>& pieces) {
    dgc_handle* h = s.h;
    const dgc_opts& o = h->opts;
    dgc_dbatch& db = s.db;
    cudaStream_t st = s.stream, cs = s.copy_stream;
    bool turn_released = false;
    auto fail = [&](int rc) {
        if (!turn_released) { release_h2d_turn_now(h, s.seq + 1); turn_released = true; }
        return rc;
    };
    int rc = upload_batch(s, b, db, true, false);  // the small arrays, on the compute stream
    if (rc) return fail(rc);
    {
        cudaEvent_t e0 = s.events.get(), e1 = s.events.get();
        cudaEventRecord(e0, cs);
        for (auto& p : pieces) {
            if (p.b1 > p.b0 &&
                cudaMemcpyAsync((u8*)db.arena.p + p.b0, b->arena + p.b0, p.b1 - p.b0, cudaMemcpyHostToDevice, cs) != cudaSuccess) {
                s.err = "staged arena copy failed";
                return fail(DGC_E_CUDA);
            }
            p.copied = s.events.get();
            cudaEventRecord(p.copied, cs);
        }
        cudaEventRecord(e1, cs);
        s.spans.push_back({PH_H2D, {e0, e1}});
        TurnMsg* m = new TurnMsg{h, s.seq + 1};
        if (cudaLaunchHostFunc(cs, release_h2d_turn_cb, m) == cudaSuccess) turn_released = true;
        else { delete m; cudaGetLastError(); cudaStreamSynchronize(cs); fail(0); }
    }

    const size_t T = db.T;
    s.results = dgc_results{};
    s.results.n_targets = db.T;
    s.res_frag_begin.assign(T + 1, 0);
    s.res_r0.clear(); s.res_r1.clear(); s.res_seq_off.clear(); s.res_seq_len.clear();
    rc = ensure_alignment_arrays(s, db);
    if (rc) return rc;
    for (DevBuf* d : {&s.t_status, &s.t_V, &s.t_E, &s.t_ne, &s.t_nv, &s.t_cons_len, &s.t_nfrag, &s.frag_cap, &s.gs_cap, &s.t_frag_coff,
                      &s.pool_cap, &s.t_pool_top})
        CU(d->ensure(4 * (T + 1)));
    for (DevBuf* d : {&s.t_cols, &s.t_cons_coff, &s.rec_off, &s.pos_off, &s.cons_off, &s.frag_off, &s.gs_off})
        CU(d->ensure(8 * (T + 1)));
    CU(s.totals.ensure(64));
    CU(s.hp_plan_in.ensure(20 * (T + 1) + 16));
    CU(s.hp_plan_out.ensure(64 * (T + 1)));
    CU(cudaMemsetAsync(s.totals.p, 0, 64, st));
    CU(cudaMemsetAsync(s.t_cons_len.p, 0, 4 * (T + 1), st));
    CU(cudaMemsetAsync(s.t_nfrag.p, 0, 4 * (T + 1), st));

    uint32_t* hp_status = s.hp_plan_in.as<uint32_t>();
    uint32_t* hp_V = hp_status + (T + 1);
    uint32_t* hp_E = hp_V + (T + 1);
    uint64_t* hp_cols = (uint64_t*)(hp_E + (T + 1) + ((T + 1) & 1));
    uint64_t* p_rec = s.hp_plan_out.as<uint64_t>();
    uint64_t* p_pos = p_rec + (T + 1);
    uint64_t* p_cons = p_pos + (T + 1);
    uint64_t* p_frag = p_cons + (T + 1);
    uint64_t* p_gs = p_frag + (T + 1);
    uint32_t* p_fcap = (uint32_t*)(p_gs + (T + 1));
    uint32_t* p_gcap = p_fcap + (T + 1);
    uint32_t* p_pcap = p_gcap + (T + 1);
    const uint64_t budget = o.scratch_bytes ? o.scratch_bytes : (112ull << 30);

    Chunk cur{0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t built_to = 0, planned_to = 0, pend_maxL2 = 0, nchunks = 0;
    uint64_t out_cons_need = 0, out_frag_need = 0, live_cons = 0, live_frag = 0, raw_seen = 0;

    // does chunk c fit the arenas as they are allocated now?
    auto cap_ok = [&](const Chunk& c) {
        return 128 * c.nUnits + 256 <= s.rec.cap && c.nCons + 64 <= s.cons.cap && 8 * c.nFrag + 64 <= s.frags.cap &&
               4 * c.nGs + 64 <= s.gscratch.cap && (ACC_BYTES * c.maxL2 <= h->build_smem_max || ACC_BYTES * c.nPos + 64 <= s.gacc.cap);
    };
    // (re)allocate the chunk arenas for c; only while no chunk is open (cudaFree waits for the kernels still using them)
    auto cap_ensure = [&](const Chunk& c) -> int {
        CU(s.rec.ensure(128 * c.nUnits + 256));
        CU(s.cons.ensure(c.nCons + 64)); CU(s.frags.ensure(8 * c.nFrag + 64)); CU(s.gscratch.ensure(4 * c.nGs + 64));
        if (ACC_BYTES * c.maxL2 > h->build_smem_max) CU(s.gacc.ensure(ACC_BYTES * c.nPos + 64));
        return DGC_OK;
    };
    auto flush_build = [&]() -> int {
        if (built_to == planned_to) return DGC_OK;
        const size_t a = built_to, n = planned_to - built_to;
        PullArgs pa;
        const uint32_t nn = (uint32_t)n;
        pa.dst[0] = (u32*)(s.rec_off.as<u64>() + a); pa.src[0] = (const u32*)(p_rec + a); pa.words[0] = 2 * nn;
        pa.dst[1] = (u32*)(s.pos_off.as<u64>() + a); pa.src[1] = (const u32*)(p_pos + a); pa.words[1] = 2 * nn;
        pa.dst[2] = (u32*)(s.cons_off.as<u64>() + a); pa.src[2] = (const u32*)(p_cons + a); pa.words[2] = 2 * nn;
        pa.dst[3] = (u32*)(s.frag_off.as<u64>() + a); pa.src[3] = (const u32*)(p_frag + a); pa.words[3] = 2 * nn;
        pa.dst[4] = (u32*)(s.gs_off.as<u64>() + a); pa.src[4] = (const u32*)(p_gs + a); pa.words[4] = 2 * nn;
        pa.dst[5] = s.frag_cap.as<u32>() + a; pa.src[5] = p_fcap + a; pa.words[5] = nn;
        pa.dst[6] = s.gs_cap.as<u32>() + a; pa.src[6] = p_gcap + a; pa.words[6] = nn;
        pa.dst[7] = s.pool_cap.as<u32>() + a; pa.src[7] = p_pcap + a; pa.words[7] = nn;
        k_pull_words<<<16, 256, 0, st>>>(pa);  // not a memcpy: see k_pull_words
        s.tm.launches++;
        CU(cudaGetLastError());
        const int r = launch_build(s, db, built_to, planned_to - built_to, pend_maxL2);
        built_to = planned_to; pend_maxL2 = 0;
        return r;
    };
    auto close_chunk = [&]() -> int {
        int r = flush_build();
        if (r || !cur.count) return r;
        CU(s.out_cons.ensure_keep(out_cons_need + 64, live_cons, st));
        CU(s.out_frags.ensure_keep(8 * out_frag_need + 64, 8 * live_frag, st));
        live_cons = out_cons_need; live_frag = out_frag_need;
        r = launch_serial(s, db, cur, 2);
        s.tm.arena_bytes += chunk_bytes(cur);
        nchunks++;
        cur = Chunk{cur.first + cur.count, 0, 0, 0, 0, 0, 0, 0};
        return r;
    };

    std::vector<Chunk> need;  // per target of the current piece: what it adds to a chunk
    Chunk total{0, 0, 0, 0, 0, 0, 0, 0};
    for (size_t pi = 0; pi < pieces.size(); pi++) {
        const Piece& p = pieces[pi];
        const size_t t0 = p.t0, nT = p.t1 - p.t0;
        CU(cudaStreamWaitEvent(st, p.copied, 0));
        span_begin(s, PH_NORM);
        rc = launch_normalize(s, db, p.k0, p.k1);
        if (rc) return rc;
        span_end(s);
        span_begin(s, PH_COUNT);
        rc = launch_plan(s, db, p.t0, p.t1);
        if (rc) return rc;
        span_end(s);
        CU(cudaMemcpyAsync(hp_status + t0, s.t_status.as<u32>() + t0, 4 * nT, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(hp_V + t0, s.t_V.as<u32>() + t0, 4 * nT, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(hp_E + t0, s.t_E.as<u32>() + t0, 4 * nT, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(hp_cols + t0, s.t_cols.as<u64>() + t0, 8 * nT, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));

        // ---- what every target of the piece needs (same sizing rules as run_pipeline)
        need.assign(nT, Chunk{0, 1, 0, 0, 0, 0, 0, 0});
        Chunk sum{0, 0, 0, 0, 0, 0, 0, 0};
        uint64_t piece_cons = 0, piece_frag = 0;
        for (size_t i = 0; i < nT; i++) {
            const size_t t = t0 + i;
            const bool ok = hp_status[t] == T_OK;
            const uint64_t V = ok ? hp_V[t] : 0, E = ok ? hp_E[t] : 0, L2 = ok ? (uint64_t)db.h_tlen[t] + 2 : 0;
            const uint32_t A = db.h_aln_begin[t + 1] - db.h_aln_begin[t];
            const uint64_t pcap = ok ? pool_words(V, E, db.h_tlen[t], A, false) : 0;
            const uint64_t units = ok ? L2 + V + pcap / 16 : 0;
            const uint64_t fcap = ok ? V / ((uint64_t)o.min_len + 1) + 2 : 0, gcap = ok ? 8ull * (A + 2) + 64 : 0;
            Chunk& n = need[i];
            n.nUnits = units; n.nPos = pos16(L2); n.nCons = (V + 15) & ~15ull; n.nFrag = fcap; n.nGs = gcap; n.maxL2 = (uint32_t)L2;
            if (16 * units > 0xFFFFFFF0ull) { s.err = "one target's graph exceeds the 32 GiB a list pointer can address"; return DGC_E_NOMEM; }
            if (chunk_bytes(n) > budget) { s.err = "one target needs more than scratch_bytes of graph arena"; return DGC_E_NOMEM; }
            p_fcap[t] = (uint32_t)fcap; p_gcap[t] = (uint32_t)gcap; p_pcap[t] = (uint32_t)pcap;
            sum.nUnits += units; sum.nPos += pos16(L2); sum.nCons += n.nCons; sum.nFrag += fcap; sum.nGs += gcap;
            sum.maxL2 = std::max(sum.maxL2, n.maxL2);
            piece_cons += V; piece_frag += fcap;
            s.tm.nodes += V; s.tm.norm_cols += ok ? hp_cols[t] : 0;
        }
        total.nUnits += sum.nUnits; total.nPos += sum.nPos; total.nCons += sum.nCons; total.nFrag += sum.nFrag; total.nGs += sum.nGs;
        uint64_t piece_raw = 0;
        for (uint32_t k = p.k0; k < p.k1; k++) piece_raw += b->aln_len[k];
        raw_seen += piece_raw;
        // ---- with no chunk open the arenas may be (re)allocated: size them for the rest of the batch, scaled from
        // this piece by raw columns, within the budget
        if (cur.count == 0) {
            const double scale = piece_raw ? 1.06 * (double)(db.raw_cols - raw_seen + piece_raw) / (double)piece_raw : 1.0;
            Chunk est = sum;
            est.nUnits = (uint64_t)(sum.nUnits * scale); est.nPos = (uint64_t)(sum.nPos * scale); est.nCons = (uint64_t)(sum.nCons * scale);
            est.nFrag = (uint64_t)(sum.nFrag * scale); est.nGs = (uint64_t)(sum.nGs * scale);
            if (chunk_bytes(est) > budget) {
                const double f = (double)budget / (double)chunk_bytes(est);
                est.nUnits = (uint64_t)(est.nUnits * f); est.nPos = (uint64_t)(est.nPos * f); est.nCons = (uint64_t)(est.nCons * f);
                est.nFrag = (uint64_t)(est.nFrag * f); est.nGs = (uint64_t)(est.nGs * f);
            }
            if (pi == 0) {  // an estimate that fell short last time costs a second sweep: remember what a whole batch took
                Chunk hx = est;
                hx.nUnits = std::max(est.nUnits, s.hint[0]); hx.nPos = std::max(est.nPos, s.hint[1]); hx.nCons = std::max(est.nCons, s.hint[2]);
                hx.nFrag = std::max(est.nFrag, s.hint[3]); hx.nGs = std::max(est.nGs, s.hint[4]);
                if (chunk_bytes(hx) <= budget) est = hx;
            }
            if (!cap_ok(est) && (rc = cap_ensure(est)) != DGC_OK) {  // cannot have it all: fall back to what this piece needs
                cudaGetLastError();
                s.err.clear();
            }
            const uint64_t oc = (uint64_t)((out_cons_need + piece_cons) + piece_cons * (scale - 1.0));
            const uint64_t of = (uint64_t)((out_frag_need + piece_frag) + piece_frag * (scale - 1.0));
            CU(s.out_cons.ensure_keep(oc + 64, live_cons, st));
            CU(s.out_frags.ensure_keep(8 * of + 64, 8 * live_frag, st));
        }
        // ---- arena offsets target by target; a chunk closes when the next target does not fit
        for (size_t i = 0; i < nT; i++) {
            const size_t t = t0 + i;
            const Chunk& n = need[i];
            Chunk nxt = cur;
            nxt.nUnits += n.nUnits; nxt.nPos += n.nPos; nxt.nCons += n.nCons; nxt.nFrag += n.nFrag; nxt.nGs += n.nGs; nxt.count++;
            nxt.maxL2 = std::max(cur.maxL2, n.maxL2);
            if (cur.count && (chunk_bytes(nxt) > budget || !cap_ok(nxt) || cur.count >= (1u << 20))) {
                if ((rc = close_chunk()) != DGC_OK) return rc;
                nxt = n; nxt.first = (uint32_t)t; nxt.count = 1;
            }
            if (cur.count == 0) {
                cur.first = (uint32_t)t;
                if (!cap_ok(nxt)) {  // an empty chunk that still does not fit: grow to the rest of this piece (or this target)
                    Chunk rest{0, 0, 0, 0, 0, 0, 0, 0};
                    for (size_t j = i; j < nT; j++) {
                        rest.nUnits += need[j].nUnits; rest.nPos += need[j].nPos; rest.nCons += need[j].nCons;
                        rest.nFrag += need[j].nFrag; rest.nGs += need[j].nGs; rest.maxL2 = std::max(rest.maxL2, need[j].maxL2);
                    }
                    if (chunk_bytes(rest) > budget || cap_ensure(rest) != DGC_OK) {
                        cudaGetLastError();
                        s.err.clear();
                        if ((rc = cap_ensure(nxt)) != DGC_OK) return rc;
                    }
                }
            }
            p_rec[t] = cur.nUnits; p_pos[t] = cur.nPos; p_cons[t] = cur.nCons; p_frag[t] = cur.nFrag; p_gs[t] = cur.nGs;
            cur.nUnits += n.nUnits; cur.nPos += n.nPos; cur.nCons += n.nCons; cur.nFrag += n.nFrag; cur.nGs += n.nGs; cur.count++;
            cur.maxL2 = std::max(cur.maxL2, n.maxL2);
            pend_maxL2 = std::max(pend_maxL2, n.maxL2);
            out_cons_need += hp_status[t] == T_OK ? hp_V[t] : 0; out_frag_need += n.nFrag;
            planned_to = (uint32_t)t + 1;
        }
        if ((rc = flush_build()) != DGC_OK) return rc;
    }
    if ((rc = close_chunk()) != DGC_OK) return rc;
    s.tm.chunks = nchunks;
    if (chunk_bytes(total) <= budget) {
        const uint64_t add = 1024;
        s.hint[0] = total.nUnits + add; s.hint[1] = total.nPos + add; s.hint[2] = total.nCons + add; s.hint[3] = total.nFrag + add;
        s.hint[4] = total.nGs + add;
    }
    if ((rc = retry_pool_failures(s, db, 2)) != DGC_OK) return rc;
    return fetch_results(s, db);
}

void finish_timings(Slot& s) {
    float acc[PH_N] = {0};
    for (auto& sp : s.spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.second.first, sp.second.second) == cudaSuccess) acc[sp.first] += ms;
    }
    s.tm.ms_h2d = acc[PH_H2D]; s.tm.ms_normalize = acc[PH_NORM]; s.tm.ms_count = acc[PH_COUNT]; s.tm.ms_build = acc[PH_BUILD];
    s.tm.ms_merge = acc[PH_MERGE]; s.tm.ms_bestpath = acc[PH_PATH]; s.tm.ms_compact = acc[PH_COMPACT]; s.tm.ms_d2h = acc[PH_D2H];
    s.tm.ms_realign = acc[PH_DAZ];
    s.tm.ms_total = 0;
    for (int i = 0; i < PH_N; i++) s.tm.ms_total += acc[i];
    s.tm.ms_span = 0;
    if (!s.spans.empty()) cudaEventElapsedTime(&s.tm.ms_span, s.spans.front().second.first, s.spans.back().second.second);
}

void begin_run(Slot& s) {
    s.events.reset();
    s.spans.clear();
    s.tm = dgc_timings{};
    s.err.clear();
}

int slot_process_daz(Slot& s, const dgc_daz_batch* b);

int slot_process_pre(Slot& s, const dgc_pre_batch* b);

int slot_process(Slot& s, const dgc_batch* b) {
    if (s.is_daz) return slot_process_daz(s, &s.pending_daz);
    if (s.is_pre) return slot_process_pre(s, &s.pending_pre);
    begin_run(s);
    dgc_handle* h = s.h;
    {
        std::unique_lock<std::mutex> lk(h->h2d_mu);
        h->h2d_cv.wait(lk, [&] { return h->h2d_turn == s.seq; });
    }
    int rc;
    std::vector<Piece> pieces;
    const uint32_t want = b->arena_bytes >= h->stage_min_bytes ? (uint32_t)std::min<uint64_t>(16, b->arena_bytes / h->stage_piece_bytes) : 0;
    if (split_pieces(b, want, pieces)) {
        rc = run_staged(s, b, pieces);  // hands the link to the next batch as soon as its own bytes have landed
    } else {
        trace(s, "slot: got the H2D turn");
        rc = upload_batch(s, b, s.db, true);
        cudaStreamSynchronize(s.stream);
        trace(s, "slot: batch uploaded");
        release_h2d_turn_now(h, s.seq + 1);
        if (rc == DGC_OK) rc = run_pipeline(s, s.db, 2, 0xFFFFFFFFu);
    }
    cudaStreamSynchronize(s.copy_stream);
    cudaStreamSynchronize(s.stream);
    finish_timings(s);
    return rc;
}


// ------------------------------------------------------------------------------------------
// dgc_submit_daz: the overlaps' gapped strings are built on the device (daz_realign.cuh) inside the slot's batch arena,
// then the batch runs through the same kernels as any other.  All sizes the planner needs are bounds that follow
// from the trace points: an overlap has at most (aepos - abpos + 1) + sum(diffs) columns.
int validate_daz(const dgc_handle* h, const dgc_daz_batch* b, std::string& err) {
    if (!b) { err = "null batch"; return DGC_E_INVALID; }
    if (!h->daz_n_reads) { err = "dgc_daz_set_db has not been called"; return DGC_E_STATE; }
    if (b->tspace == 0 || b->tspace > 125) { err = "trace spacing above 125 (two-byte trace values): realign on the host"; return DGC_E_DAZ; }
    if (b->n_targets && (!b->tgt_aread || !b->tgt_ovl_begin)) { err = "null target arrays"; return DGC_E_INVALID; }
    if (b->n_ovls && (!b->ovl_bread || !b->ovl_comp || !b->ovl_abpos || !b->ovl_aepos || !b->ovl_bbpos || !b->ovl_bepos ||
                      !b->ovl_trace_off || !b->ovl_tlen || !b->trace)) { err = "null overlap arrays"; return DGC_E_INVALID; }
    if (b->n_targets && (b->tgt_ovl_begin[0] != 0 || b->tgt_ovl_begin[b->n_targets] != b->n_ovls)) { err = "tgt_ovl_begin is not a CSR over n_ovls"; return DGC_E_INVALID; }
    for (uint32_t t = 0; t < b->n_targets; t++) {
        if (b->tgt_ovl_begin[t + 1] < b->tgt_ovl_begin[t]) { err = "tgt_ovl_begin not monotone"; return DGC_E_INVALID; }
        if (b->tgt_aread[t] >= h->daz_n_reads) { err = "A read outside the DB"; return DGC_E_INVALID; }
        const int alen = (int)h->daz_h_rlen[b->tgt_aread[t]];
        for (uint32_t k = b->tgt_ovl_begin[t]; k < b->tgt_ovl_begin[t + 1]; k++) {
            if (b->ovl_bread[k] >= h->daz_n_reads) { err = "B read outside the DB"; return DGC_E_INVALID; }
            const int blen = (int)h->daz_h_rlen[b->ovl_bread[k]];
            const int ab = b->ovl_abpos[k], ae = b->ovl_aepos[k], bb = b->ovl_bbpos[k], be = b->ovl_bepos[k];
            if (ab < 0 || ae < ab || ae > alen || bb < 0 || be < bb || be > blen) { err = "overlap coordinates outside its reads"; return DGC_E_INVALID; }
            const uint32_t tl = b->ovl_tlen[k];
            if (tl < 2 || (tl & 1)) { err = "overlap without trace points"; return DGC_E_DAZ; }
            if (b->ovl_trace_off[k] > b->trace_bytes || tl > b->trace_bytes - b->ovl_trace_off[k]) { err = "trace outside its array"; return DGC_E_INVALID; }
            // the segments the trace points imply must tile [abpos, aepos) x [bbpos, bepos)
            const uint32_t nseg = tl / 2, want = (uint32_t)(ae / (int)b->tspace - ab / (int)b->tspace) + ((ae % (int)b->tspace) ? 1u : 0u);
            if (nseg != (want ? want : 1u)) { err = "trace length does not match the overlap's A interval"; return DGC_E_DAZ; }
            const uint8_t* p = b->trace + b->ovl_trace_off[k];
            int64_t sb = bb;
            for (uint32_t i = 0; i + 1 < nseg; i++) sb += p[2 * i + 1];
            if (sb > be || be - sb > DZ_MAX_N) { err = "trace points do not fit the overlap's B interval"; return DGC_E_DAZ; }
        }
    }
    return DGC_OK;
}

int build_daz_batch(Slot& s, const dgc_daz_batch* b, bool strings_only) {
    dgc_handle* h = s.h;
    dgc_dbatch& db = s.db;
    cudaStream_t st = s.stream;
    const size_t T = b->n_targets, O = b->n_ovls;
    // ---- host planning: offsets of everything from the trace-point bounds
    // plan layout: u32 tlen[T+1] | u32 start[O+1] | u32 segb[O+1] | u64 bb_off[T+1] | u64 q_off[O+1] | u64 t_off[O+1] | u64 norm_off[O+1] | u64 entry_off[O+1]
    CU(s.zh_plan.ensure(4 * (T + 1) + 8 * (O + 1) + 8 * (T + 1) + 32 * (O + 1) + 64));
    uint32_t* p_tlen = s.zh_plan.as<uint32_t>();
    uint32_t* p_start = p_tlen + (T + 1);
    uint32_t* p_segb = p_start + (O + 1);
    uint64_t* p_bb = (uint64_t*)(p_segb + (O + 1) + ((T + 1 + 2 * (O + 1)) & 1));
    uint64_t* p_q = p_bb + (T + 1);
    uint64_t* p_t = p_q + (O + 1);
    uint64_t* p_norm = p_t + (O + 1);
    uint64_t* p_ent = p_norm + (O + 1);
    uint64_t arena = 0, norm = 0, ent = 0, raw = 0;
    uint32_t nseg_tot = 0, dmax = 0, nmax = 0;
    for (size_t t = 0; t < T; t++) {
        const uint32_t L = h->daz_h_rlen[b->tgt_aread[t]];
        p_tlen[t] = L;
        p_bb[t] = arena;
        arena += ((uint64_t)L + 15) & ~15ull;
        for (uint32_t k = b->tgt_ovl_begin[t]; k < b->tgt_ovl_begin[t + 1]; k++) {
            const uint8_t* p = b->trace + b->ovl_trace_off[k];
            const uint32_t nseg = b->ovl_tlen[k] / 2;
            uint64_t sumd = 0, sb = (uint64_t)b->ovl_bbpos[k];
            for (uint32_t i = 0; i < nseg; i++) {
                sumd += p[2 * i];
                dmax = std::max<uint32_t>(dmax, p[2 * i]);
                if (i + 1 < nseg) { nmax = std::max<uint32_t>(nmax, p[2 * i + 1]); sb += p[2 * i + 1]; }
            }
            nmax = std::max<uint32_t>(nmax, (uint32_t)((uint64_t)b->ovl_bepos[k] - sb));
            const uint64_t cols = (uint64_t)(b->ovl_aepos[k] - b->ovl_abpos[k] + 1) + sumd;
            p_start[k] = (uint32_t)(b->ovl_abpos[k] + 1);            // dagcon::Alignment::start (:356)
            p_segb[k] = nseg_tot;
            p_q[k] = arena; arena += (cols + 15) & ~15ull;
            p_t[k] = arena; arena += (cols + 15) & ~15ull;
            p_norm[k] = norm; norm += (2 * cols + 7) & ~7ull;
            p_ent[k] = ent; ent += sumd;
            raw += cols;
            nseg_tot += nseg;
        }
    }
    p_segb[O] = nseg_tot;
    if (arena > (1ull << 40) || nseg_tot > 0x7FFFFFF0u) { s.err = "dazcon batch too large"; return DGC_E_NOMEM; }
    db.T = (uint32_t)T; db.A = (uint32_t)O; db.has_bb = true; db.packed4 = false; db.arena_bytes = arena; db.norm_cols = norm; db.raw_cols = raw;
    db.h_tlen.assign(p_tlen, p_tlen + T);
    db.h_aln_begin.assign(b->tgt_ovl_begin, b->tgt_ovl_begin + T + 1);
    if (T == 0) return DGC_OK;
    CU(db.tgt_tlen.ensure(4 * (T + 1))); CU(db.tgt_aln_begin.ensure(4 * (T + 1))); CU(db.tgt_bb_off.ensure(8 * (T + 1)));
    CU(db.aln_start.ensure(4 * (O + 1))); CU(db.aln_len.ensure(4 * (O + 1)));
    CU(db.aln_q_off.ensure(8 * (O + 1))); CU(db.aln_t_off.ensure(8 * (O + 1))); CU(db.aln_norm_off.ensure(8 * (O + 1)));
    CU(db.arena.ensure(arena + 64));
    for (DevBuf* d : {&s.z_bread, &s.z_comp, &s.z_abpos, &s.z_aepos, &s.z_bbpos, &s.z_bepos, &s.z_tlen, &s.z_segb, &s.z_aread}) CU(d->ensure(4 * (O + 1)));
    CU(s.z_troff.ensure(8 * (O + 1))); CU(s.z_eoff.ensure(8 * (O + 1))); CU(s.z_tgt_aread.ensure(4 * (T + 1)));
    CU(s.z_trace.ensure(b->trace_bytes + 16));
    CU(s.z_seg_bb.ensure(4ull * nseg_tot + 16)); CU(s.z_seg_eoff.ensure(4ull * nseg_tot + 16));
    for (DevBuf* d : {&s.z_seg_cnt, &s.z_seg_neg, &s.z_seg_ppos, &s.z_seg_pneg}) CU(d->ensure(2ull * nseg_tot + 16));
    CU(s.z_entries.ensure(4 * ent + 16));
    CU(s.z_err.ensure(16));
    // wave slabs: rows x width cells per thread, as many threads as the budget allows
    const uint32_t rows = dmax + 3;
    const uint32_t koff = nmax + dmax / 2 + 4;
    const uint32_t width = koff + std::max<uint32_t>(b->tspace, DZ_MAX_M) + dmax / 2 + 6;
    const uint64_t slab_cells = (uint64_t)rows * width;
    uint64_t n_slabs = std::min<uint64_t>(nseg_tot, (uint64_t)h->sm_count * 512);
    n_slabs = std::max<uint64_t>(1, std::min<uint64_t>(n_slabs, (3ull << 30) / (3 * slab_cells)));
    CU(s.z_vf.ensure(2 * slab_cells * n_slabs + 64)); CU(s.z_hf.ensure(slab_cells * n_slabs + 64));
    // per-overlap A read (the kernels index overlaps, not targets)
    std::vector<uint32_t> aread(O);
    for (size_t t = 0; t < T; t++)
        for (uint32_t k = b->tgt_ovl_begin[t]; k < b->tgt_ovl_begin[t + 1]; k++) aread[k] = b->tgt_aread[t];
    span_begin(s, PH_H2D);
    CU(cudaMemcpyAsync(db.tgt_tlen.p, p_tlen, 4 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(db.tgt_aln_begin.p, b->tgt_ovl_begin, 4 * (T + 1), cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(db.tgt_bb_off.p, p_bb, 8 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(s.z_tgt_aread.p, b->tgt_aread, 4 * T, cudaMemcpyHostToDevice, st));
    if (O) {
        CU(cudaMemcpyAsync(db.aln_start.p, p_start, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_q_off.p, p_q, 8 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_t_off.p, p_t, 8 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_norm_off.p, p_norm, 8 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_aread.p, aread.data(), 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_bread.p, b->ovl_bread, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_comp.p, b->ovl_comp, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_abpos.p, b->ovl_abpos, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_aepos.p, b->ovl_aepos, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_bbpos.p, b->ovl_bbpos, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_bepos.p, b->ovl_bepos, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_tlen.p, b->ovl_tlen, 4 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_troff.p, b->ovl_trace_off, 8 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_segb.p, p_segb, 4 * (O + 1), cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_eoff.p, p_ent, 8 * O, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.z_trace.p, b->trace, b->trace_bytes, cudaMemcpyHostToDevice, st));
    }
    CU(cudaMemsetAsync(s.z_err.p, 0, 16, st));
    span_end(s);
    CU(cudaStreamSynchronize(st));   // `aread` and the caller's arrays may go; the link is free for the next batch
    release_h2d_turn_now(h, s.seq + 1);
    if (O == 0 && !strings_only) return DGC_OK;
    DazArgs a;
    a.db.bases = h->daz_bases.as<u8>(); a.db.boff = h->daz_boff.as<u64>(); a.db.rlen = h->daz_rlen.as<u32>();
    a.n_ovl = (u32)O; a.n_seg = nseg_tot; a.tspace = b->tspace;
    a.ovl_aread = s.z_aread.as<u32>(); a.ovl_bread = s.z_bread.as<u32>(); a.ovl_comp = s.z_comp.as<u32>();
    a.ovl_abpos = s.z_abpos.as<int>(); a.ovl_aepos = s.z_aepos.as<int>(); a.ovl_bbpos = s.z_bbpos.as<int>(); a.ovl_bepos = s.z_bepos.as<int>();
    a.ovl_trace_off = s.z_troff.as<u64>(); a.ovl_tlen = s.z_tlen.as<u32>(); a.ovl_seg_begin = s.z_segb.as<u32>();
    a.ovl_entry_off = s.z_eoff.as<u64>(); a.trace = s.z_trace.as<u8>();
    a.seg_bb = s.z_seg_bb.as<int>(); a.seg_eoff = s.z_seg_eoff.as<u32>(); a.seg_cnt = s.z_seg_cnt.as<u16>(); a.seg_neg = s.z_seg_neg.as<u16>();
    a.seg_ppos = s.z_seg_ppos.as<u16>(); a.seg_pneg = s.z_seg_pneg.as<u16>(); a.entries = s.z_entries.as<int>();
    a.vf = s.z_vf.as<short>(); a.hf = s.z_hf.as<signed char>(); a.rows = rows; a.width = width; a.koff = koff; a.n_slabs = (u32)n_slabs;
    a.arena = db.arena.as<u8>(); a.aln_q_off = db.aln_q_off.as<u64>(); a.aln_t_off = db.aln_t_off.as<u64>(); a.aln_len = db.aln_len.as<u32>();
    a.error = s.z_err.as<u32>();
    span_begin(s, PH_DAZ);
    if (O) {
        k_daz_prep<<<(unsigned)((O + 127) / 128), 128, 0, st>>>(a);
        k_daz_iter<<<(unsigned)((n_slabs + 127) / 128), 128, 0, st>>>(a);
        k_daz_scan<<<(unsigned)((O + 127) / 128), 128, 0, st>>>(a);
        k_daz_emit<<<(nseg_tot + 127) / 128, 128, 0, st>>>(a);
        s.tm.launches += 4;
    }
    k_daz_backbone<<<(unsigned)T, 256, 0, st>>>(a.db, (u32)T, s.z_tgt_aread.as<u32>(), db.tgt_bb_off.as<u64>(), db.arena.as<u8>());
    s.tm.launches++;
    CU(cudaGetLastError());
    span_end(s);
    uint32_t zerr = 0;
    CU(cudaMemcpyAsync(&zerr, s.z_err.p, 4, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (zerr) { s.err = "device realignment failed on a segment (code " + std::to_string(zerr) + "): the trace points do not bound its edits"; return DGC_E_DAZ; }
    return DGC_OK;
}

int slot_process_daz(Slot& s, const dgc_daz_batch* b) {
    begin_run(s);
    dgc_handle* h = s.h;
    {
        std::unique_lock<std::mutex> lk(h->h2d_mu);
        h->h2d_cv.wait(lk, [&] { return h->h2d_turn == s.seq; });
    }
    bool released = false;
    int rc = build_daz_batch(s, b, false);
    {   // build_daz_batch hands the link on once its arrays have landed; on an early error do it here
        std::lock_guard<std::mutex> lk(h->h2d_mu);
        released = h->h2d_turn > s.seq;
    }
    if (!released) release_h2d_turn_now(h, s.seq + 1);
    if (rc == DGC_OK) rc = run_pipeline(s, s.db, 2, 0xFFFFFFFFu);
    cudaStreamSynchronize(s.stream);
    finish_timings(s);
    return rc;
}

// ------------------------------------------------------------------------------------------
// dgc_submit_pre: the `-a` step on the device (pre_align.cuh), then the batch runs through the same kernels as any
// other.  A record's gapped strings have at most qlen + tseglen columns: every offset is planned from that bound.
int validate_pre(const dgc_pre_batch* b, std::string& err) {
    if (!b) { err = "null batch"; return DGC_E_INVALID; }
    if (b->n_targets && (!b->tgt_tlen || !b->tgt_rec_begin)) { err = "null target arrays"; return DGC_E_INVALID; }
    if (b->n_recs && (!b->rec_start || !b->rec_strand || !b->rec_qlen || !b->rec_tseglen || !b->rec_q_off || !b->rec_t_off || !b->arena)) {
        err = "null record arrays"; return DGC_E_INVALID;
    }
    if (b->n_targets && (b->tgt_rec_begin[0] != 0 || b->tgt_rec_begin[b->n_targets] != b->n_recs)) { err = "tgt_rec_begin is not a CSR over n_recs"; return DGC_E_INVALID; }
    for (uint32_t t = 0; t < b->n_targets; t++) {
        if (b->tgt_rec_begin[t + 1] < b->tgt_rec_begin[t]) { err = "tgt_rec_begin not monotone"; return DGC_E_INVALID; }
        if (b->tgt_tlen[t] > 0x3FFFFFFFu) { err = "tlen too large"; return DGC_E_INVALID; }
    }
    for (uint32_t k = 0; k < b->n_recs; k++) {
        const uint64_t nq = b->rec_qlen[k], nt = b->rec_tseglen[k];
        if (b->rec_q_off[k] > b->arena_bytes || nq > b->arena_bytes - b->rec_q_off[k] || b->rec_t_off[k] > b->arena_bytes ||
            nt > b->arena_bytes - b->rec_t_off[k]) { err = "record segment outside arena"; return DGC_E_INVALID; }
        if (nq > 0x0FFFFFFFu || nt > 0x0FFFFFFFu) { err = "record too long"; return DGC_E_INVALID; }
    }
    return DGC_OK;
}

int build_pre_batch(Slot& s, const dgc_pre_batch* b) {
    dgc_handle* h = s.h;
    dgc_dbatch& db = s.db;
    cudaStream_t st = s.stream;
    const size_t T = b->n_targets, R = b->n_recs;
    // plan: u64 q_out[R+1] | u64 t_out[R+1] | u64 norm_off[R+1] | u64 tb_off[R+1] | u32 tglen[R+1] | u32 ids[R+1] | u32 state[R+1]
    CU(s.ph_plan.ensure(32 * (R + 1) + 12 * (R + 1) + 64));
    uint64_t* p_q = s.ph_plan.as<uint64_t>();
    uint64_t* p_t = p_q + (R + 1);
    uint64_t* p_norm = p_t + (R + 1);
    uint64_t* p_tb = p_norm + (R + 1);
    uint32_t* p_tg = (uint32_t*)(p_tb + (R + 1));
    uint32_t* p_ids = p_tg + (R + 1);
    uint32_t* p_state = p_ids + (R + 1);
    uint64_t arena = 0, norm = 0, raw = 0;
    for (size_t t = 0; t < T; t++)
        for (uint32_t k = b->tgt_rec_begin[t]; k < b->tgt_rec_begin[t + 1]; k++) {
            const uint64_t cap = (uint64_t)b->rec_qlen[k] + b->rec_tseglen[k];
            p_q[k] = arena; arena += (cap + 15) & ~15ull;
            p_t[k] = arena; arena += (cap + 15) & ~15ull;
            p_norm[k] = norm; norm += (2 * cap + 7) & ~7ull;
            p_tg[k] = b->tgt_tlen[t];
            raw += cap;
        }
    db.T = (uint32_t)T; db.A = (uint32_t)R; db.has_bb = false; db.packed4 = false; db.arena_bytes = arena; db.norm_cols = norm; db.raw_cols = raw;
    db.h_tlen.assign(b->tgt_tlen, b->tgt_tlen + T);
    db.h_aln_begin.assign(b->tgt_rec_begin, b->tgt_rec_begin + T + 1);
    if (T == 0) return DGC_OK;
    CU(db.tgt_tlen.ensure(4 * (T + 1))); CU(db.tgt_aln_begin.ensure(4 * (T + 1)));
    CU(db.aln_start.ensure(4 * (R + 1))); CU(db.aln_len.ensure(4 * (R + 1)));
    CU(db.aln_q_off.ensure(8 * (R + 1))); CU(db.aln_t_off.ensure(8 * (R + 1))); CU(db.aln_norm_off.ensure(8 * (R + 1)));
    CU(db.arena.ensure(arena + 64));
    CU(s.p_in.ensure(b->arena_bytes + 64));
    for (DevBuf* d : {&s.p_qlen, &s.p_tlen, &s.p_strand, &s.p_start, &s.p_tglen, &s.p_ids, &s.p_tbeg, &s.p_tend, &s.p_state}) CU(d->ensure(4 * (R + 1)));
    CU(s.p_qoff.ensure(8 * (R + 1))); CU(s.p_toff.ensure(8 * (R + 1))); CU(s.p_tboff.ensure(8 * (R + 1)));
    span_begin(s, PH_H2D);
    CU(cudaMemcpyAsync(db.tgt_tlen.p, b->tgt_tlen, 4 * T, cudaMemcpyHostToDevice, st));
    CU(cudaMemcpyAsync(db.tgt_aln_begin.p, b->tgt_rec_begin, 4 * (T + 1), cudaMemcpyHostToDevice, st));
    if (R) {
        CU(cudaMemcpyAsync(db.aln_q_off.p, p_q, 8 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_t_off.p, p_t, 8 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_norm_off.p, p_norm, 8 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_qoff.p, b->rec_q_off, 8 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_toff.p, b->rec_t_off, 8 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_qlen.p, b->rec_qlen, 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_tlen.p, b->rec_tseglen, 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_strand.p, b->rec_strand, 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_start.p, b->rec_start, 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_tglen.p, p_tg, 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(s.p_in.p, b->arena, b->arena_bytes, cudaMemcpyHostToDevice, st));
        CU(cudaMemsetAsync(s.p_state.p, 0, 4 * R, st));
    }
    span_end(s);
    CU(cudaStreamSynchronize(st));
    release_h2d_turn_now(h, s.seq + 1);
    if (R == 0) return DGC_OK;
    PreArgs a;
    a.rec_start = s.p_start.as<u32>(); a.rec_tlen = s.p_tglen.as<u32>(); a.aln_start = db.aln_start.as<u32>();
    a.arena_in = s.p_in.as<u8>(); a.q_off = s.p_qoff.as<u64>(); a.t_off = s.p_toff.as<u64>(); a.q_len = s.p_qlen.as<u32>(); a.t_len = s.p_tlen.as<u32>();
    a.strand = s.p_strand.as<u32>(); a.tb_off = s.p_tboff.as<u64>();
    a.arena_out = db.arena.as<u8>(); a.out_q_off = db.aln_q_off.as<u64>(); a.out_t_off = db.aln_t_off.as<u64>(); a.out_len = db.aln_len.as<u32>();
    a.t_begin = s.p_tbeg.as<u32>(); a.t_end = s.p_tend.as<u32>(); a.state = s.p_state.as<u32>();
    // ---- bands 64, 128, ...: the records whose path came too close to the band's edge go round again
    std::vector<uint32_t> pending(R);
    for (size_t k = 0; k < R; k++) pending[k] = (uint32_t)k;
    const uint64_t tb_budget = 8ull << 30;
    span_begin(s, PH_DAZ);   // reported as ms_realign
    for (uint32_t band = 64; !pending.empty(); band *= 2) {
        if (band > PA_MAX_BAND) {   // hopeless records (unrelated sequences longer than the widest band): nothing aligns
            break;
        }
        const uint32_t warps = band <= 512 ? 4u : 1u;
        const size_t smem = (size_t)warps * 2 * (2 * band + 1 + 2 * PA_G) * sizeof(int);
        if (smem > 48 * 1024) CU(cudaFuncSetAttribute(k_pre_align, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        size_t done = 0;
        while (done < pending.size()) {   // waves that fit the move-matrix budget
            uint64_t tb = 0;
            size_t n = 0;
            while (done + n < pending.size()) {
                const uint32_t k = pending[done + n];
                const uint64_t full = (uint64_t)std::max(b->rec_qlen[k], b->rec_tseglen[k]) + 1;
                const uint64_t W = 2 * std::min<uint64_t>(band, full) + 1;
                const uint64_t need = (((uint64_t)b->rec_qlen[k] + 1) * W + 15) & ~15ull;
                if (n && tb + need > tb_budget) break;
                p_tb[n] = tb; p_ids[n] = k;
                tb += need;
                n++;
            }
            CU(s.p_tb.ensure(tb + 64));
            CU(s.p_ids.ensure(4 * (n + 1)));
            CU(cudaMemcpyAsync(s.p_tboff.p, p_tb, 8 * n, cudaMemcpyHostToDevice, st));
            CU(cudaMemcpyAsync(s.p_ids.p, p_ids, 4 * n, cudaMemcpyHostToDevice, st));
            a.rec_ids = s.p_ids.as<u32>(); a.n_launch = (u32)n; a.band = band; a.warps = warps; a.tb = s.p_tb.as<u8>();
            k_pre_align<<<(unsigned)((n + warps - 1) / warps), warps * 32, smem, st>>>(a);
            s.tm.launches++;
            CU(cudaGetLastError());
            CU(cudaStreamSynchronize(st));   // p_tb / p_ids are reused by the next wave
            done += n;
        }
        CU(cudaMemcpyAsync(p_state, s.p_state.p, 4 * R, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        std::vector<uint32_t> next;
        for (uint32_t k : pending)
            if (p_state[k] == 2) next.push_back(k);
        pending.swap(next);
    }
    if (!pending.empty()) {   // beyond the widest band: zero-column alignments (they still count towards the target's coverage)
        for (uint32_t k : pending) p_state[k] = 3;
        std::vector<uint32_t> zero(R, 0), st1(R, 0);
        CU(cudaMemcpyAsync(zero.data(), db.aln_len.p, 4 * R, cudaMemcpyDeviceToHost, st));
        CU(cudaMemcpyAsync(st1.data(), db.aln_start.p, 4 * R, cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        for (uint32_t k : pending) { zero[k] = 0; st1[k] = b->rec_start[k] + 1; }
        CU(cudaMemcpyAsync(db.aln_len.p, zero.data(), 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(db.aln_start.p, st1.data(), 4 * R, cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }
    span_end(s);
    return DGC_OK;
}

int slot_process_pre(Slot& s, const dgc_pre_batch* b) {
    begin_run(s);
    dgc_handle* h = s.h;
    {
        std::unique_lock<std::mutex> lk(h->h2d_mu);
        h->h2d_cv.wait(lk, [&] { return h->h2d_turn == s.seq; });
    }
    int rc = build_pre_batch(s, b);
    bool released;
    {
        std::lock_guard<std::mutex> lk(h->h2d_mu);
        released = h->h2d_turn > s.seq;
    }
    if (!released) release_h2d_turn_now(h, s.seq + 1);
    if (rc == DGC_OK) rc = run_pipeline(s, s.db, 2, 0xFFFFFFFFu);
    cudaStreamSynchronize(s.stream);
    finish_timings(s);
    return rc;
}

void worker_main(Slot* sp) {
    Slot& s = *sp;
    cudaSetDevice(s.h->device);
    std::unique_lock<std::mutex> lk(s.mu);
    for (;;) {
        s.cv.wait(lk, [&] { return s.state == Slot::QUEUED || s.state == Slot::STOP; });
        if (s.state == Slot::STOP) return;
        s.state = Slot::RUNNING;
        const dgc_batch b = s.pending;
        lk.unlock();
        int rc = slot_process(s, &b);
        lk.lock();
        s.rc = rc;
        s.state = Slot::DONE;
        s.cv.notify_all();
    }
}

}  // namespace

// ============================================================================ dgc_pre_align_noalloc: static scratch
// The one scratch array of the allocation-free entry point (include/dagcon_b200.h): 64 KB of record lists for
// k_pre_align (launch ids, slab offsets) followed by the move matrices.  Part of the module image: every GPU that
// loads the library holds it until exit.  In HOST code the name below is only a stand-in for the symbol -- never a
// device address: the address comes from cudaGetSymbolAddress, once per device (na_scratch()).
namespace {

constexpr size_t NA_LIST_BYTES = 64 * 1024;      // u32 ids[MAX_RECS] at 0, u64 tb_off[MAX_RECS] at 16 KB
constexpr size_t NA_TBOFF_AT = 4 * (size_t)DGC_NOALLOC_MAX_RECS;
static_assert(NA_TBOFF_AT + 8 * (size_t)DGC_NOALLOC_MAX_RECS <= NA_LIST_BYTES, "record lists fit their 64 KB");
static __device__ __align__(256) unsigned char g_na_scratch[NA_LIST_BYTES + (size_t)DGC_NOALLOC_SCRATCH_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mu;
unsigned char* g_na_ptr[NA_MAX_DEVICES] = {};

// the device address of g_na_scratch on `device`, which must be the current device
cudaError_t na_scratch(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_na_scratch);
        if (e != cudaSuccess) return e;
        g_na_ptr[device] = (unsigned char*)p;
    }
    *out = g_na_ptr[device];
    return cudaSuccess;
}

// records no band can take (the existing path's tail of build_pre_batch): zero columns, start = tstart + 1
__global__ void k_na_hopeless(const u32* ids, u32 n, const u32* rec_start, u32* out_len, u32* out_start, u32* tbeg, u32* tend, u32* state) {
    const u32 i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const u32 r = ids[i];
    out_len[r] = 0; out_start[r] = rec_start[r] + 1; tbeg[r] = 0; tend[r] = 0; state[r] = 3;
}

int na_fail(char* err, uint64_t cap, int rc, const char* fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, (size_t)cap, fmt, ap);
        va_end(ap);
    }
    return rc;
}

}  // namespace

// ============================================================================ C ABI
extern "C" {

int dgc_abi_version(void) { return DGC_ABI_VERSION; }

int dgc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

const char* dgc_last_error(dgc_handle* h) { return h ? h->err.c_str() : g_create_error.c_str(); }

int dgc_create(const dgc_opts* opts, dgc_handle** out) {
    if (!opts || !out) { g_create_error = "null argument"; return DGC_E_INVALID; }
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        cudaGetLastError();
        g_create_error = std::string("no CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0") +
                         " (libdagcon_b200 has no CPU path)";
        return DGC_E_NODEVICE;
    }
    if (opts->device < 0 || opts->device >= n) { g_create_error = "device ordinal out of range"; return DGC_E_INVALID; }
    if ((e = cudaSetDevice(opts->device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return DGC_E_CUDA; }
    cudaDeviceProp prop;
    if ((e = cudaGetDeviceProperties(&prop, opts->device)) != cudaSuccess) { g_create_error = cudaGetErrorString(e); return DGC_E_CUDA; }
    if (prop.major < 10) {
        g_create_error = "device is sm_" + std::to_string(prop.major * 10 + prop.minor) + "; this library is built for sm_100a only";
        return DGC_E_NODEVICE;
    }
    dgc_handle* h = new dgc_handle();
    h->opts = *opts;
    h->device = opts->device;
    h->sm_count = prop.multiProcessorCount;
    if (const char* e1 = getenv("DGC_STAGE_PIECE_BYTES")) h->stage_piece_bytes = std::max<uint64_t>(1, strtoull(e1, nullptr, 10));
    if (const char* e3 = getenv("DGC_TEST_TINY_POOL")) g_tiny_pool = e3[0] == '1';
    if (const char* e2 = getenv("DGC_STAGE_MIN_BYTES")) h->stage_min_bytes = strtoull(e2, nullptr, 10);
    {   // opt in to large dynamic shared memory for the build kernel (static use is ~4.2 KB)
        size_t optin = prop.sharedMemPerBlockOptin > 8192 ? prop.sharedMemPerBlockOptin - 8192 : 0;
        // 74 KB = three blocks per SM.  Measured on B200: 8 kb backbones 89 ms in shared memory vs 103 ms in gacc;
        // 20 kb backbones 184 ms in shared memory (180 KB, one block per SM) vs 114 ms in gacc
        optin = std::min<size_t>(optin, 74 * 1024);
        if (cudaFuncSetAttribute(k_build_graph2<u8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)optin) == cudaSuccess &&
            cudaFuncSetAttribute(k_build_graph2<u16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)optin) == cudaSuccess)
            h->build_smem_max = optin;
        else { cudaGetLastError(); h->build_smem_max = 40 * 1024; }
    }
    if (const char* e4 = getenv("DGC_BUILD_SMEM_MAX")) h->build_smem_max = std::min<size_t>(h->build_smem_max, strtoull(e4, nullptr, 10));
    for (int i = 0; i < NSLOT; i++) {
        Slot& s = h->slots[i];
        s.h = h; s.index = i;
        if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess ||
            (e = cudaStreamCreateWithFlags(&s.copy_stream, cudaStreamNonBlocking)) != cudaSuccess) {
            g_create_error = cudaGetErrorString(e);
            delete h;
            return DGC_E_CUDA;
        }
    }
    for (int i = 0; i < NSLOT; i++) h->slots[i].th = std::thread(worker_main, &h->slots[i]);
    *out = h;
    return DGC_OK;
}

int dgc_destroy(dgc_handle* h) {
    if (!h) return DGC_OK;
    cudaSetDevice(h->device);
    for (int i = 0; i < NSLOT; i++) {
        Slot& s = h->slots[i];
        {
            std::unique_lock<std::mutex> lk(s.mu);
            s.cv.wait(lk, [&] { return s.state != Slot::QUEUED && s.state != Slot::RUNNING; });
            s.state = Slot::STOP;
            s.cv.notify_all();
        }
        if (s.th.joinable()) s.th.join();
        cudaStreamSynchronize(s.stream);
        s.db.release();
        for (DevBuf* b : {&s.o_off, &s.o_len, &s.o_start, &s.o_flags, &s.o_nI, &s.o_nE, &s.o_tkept, &s.norm, &s.t_status, &s.t_V,
                          &s.t_E, &s.t_cols, &s.t_ne, &s.t_nv, &s.t_cons_len, &s.t_nfrag, &s.t_cons_coff, &s.t_frag_coff, &s.totals,
                          &s.rec_off, &s.pos_off, &s.pool_cap, &s.t_pool_top, &s.cons_off, &s.frag_off, &s.frag_cap, &s.gs_off, &s.gs_cap,
                          &s.rec, &s.cons, &s.frags, &s.gscratch, &s.gacc, &s.out_cons, &s.out_frags, &s.z_aread, &s.z_bread, &s.z_comp,
                          &s.z_abpos, &s.z_aepos, &s.z_bbpos, &s.z_bepos, &s.z_troff, &s.z_tlen, &s.z_segb, &s.z_eoff, &s.z_trace, &s.z_tgt_aread,
                          &s.z_seg_bb, &s.z_seg_eoff, &s.z_seg_cnt, &s.z_seg_neg, &s.z_seg_ppos, &s.z_seg_pneg, &s.z_entries, &s.z_vf, &s.z_hf, &s.z_err,
                          &s.p_in, &s.p_qoff, &s.p_toff, &s.p_qlen, &s.p_tlen, &s.p_strand, &s.p_start, &s.p_tglen, &s.p_ids, &s.p_tboff, &s.p_tb,
                          &s.p_tbeg, &s.p_tend, &s.p_state})
            b->release();
        s.zh_plan.release();
        s.ph_plan.release();
        for (HostBuf* b : {&s.h_norm_off, &s.hp_plan_in, &s.hp_plan_out, &s.r_status, &s.r_cons_coff, &s.r_frag_coff, &s.r_totals,
                           &s.r_cons, &s.r_frags, &s.r_nfrag})
            b->release();
        s.events.destroy();
        cudaStreamDestroy(s.stream);
        if (s.copy_stream) cudaStreamDestroy(s.copy_stream);
    }
    h->daz_bases.release(); h->daz_boff.release(); h->daz_rlen.release();
    delete h;
    return DGC_OK;
}

int dgc_host_alloc(void** p, uint64_t bytes) {
    if (!p) return DGC_E_INVALID;
    cudaError_t e = cudaMallocHost(p, bytes ? bytes : 1);
    if (e != cudaSuccess) { cudaGetLastError(); *p = nullptr; return DGC_E_CUDA; }
    return DGC_OK;
}
int dgc_host_free(void* p) {
    if (p && cudaFreeHost(p) != cudaSuccess) { cudaGetLastError(); return DGC_E_CUDA; }
    return DGC_OK;
}

int dgc_submit(dgc_handle* h, const dgc_batch* batch) {
    if (!h) return DGC_E_INVALID;
    int rc = validate_batch(batch, h->err);
    if (rc) return rc;
    int si = -1;
    {
        std::lock_guard<std::mutex> lk(h->q_mu);
        if (h->inflight.size() >= (size_t)DGC_MAX_INFLIGHT) { h->err = "too many batches in flight: collect first"; return DGC_E_STATE; }
        for (int i = 0; i < NSLOT && si < 0; i++)
            if (!h->busy[i]) si = i;
        h->busy[si] = true;
        h->inflight.push_back(si);
    }
    Slot& s = h->slots[si];
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::IDLE || s.state == Slot::DONE; });
        s.pending = *batch;
        s.is_daz = false;
        s.is_pre = false;
        s.seq = h->submitted;
        s.state = Slot::QUEUED;
        s.cv.notify_all();
    }
    h->submitted++;
    return DGC_OK;
}

int dgc_collect(dgc_handle* h, dgc_results* out) {
    if (!h || !out) return DGC_E_INVALID;
    int si;
    {
        std::lock_guard<std::mutex> lk(h->q_mu);
        if (h->inflight.empty()) { h->err = "nothing submitted"; return DGC_E_STATE; }
        si = h->inflight.front();
    }
    Slot& s = h->slots[si];
    int rc;
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::DONE; });
        rc = s.rc;
        s.state = Slot::IDLE;
    }
    h->collected++;
    h->last_tm = s.tm;
    {   // the slot is free again: its result buffers stay untouched until a later submit takes it
        std::lock_guard<std::mutex> lk(h->q_mu);
        h->inflight.pop_front();
        h->busy[si] = false;
    }
    if (rc) { h->err = s.err; return rc; }
    *out = s.results;
    return DGC_OK;
}


int dgc_daz_set_db(dgc_handle* h, const dgc_daz_db* d) {
    if (!h || !d || !d->n_reads || !d->read_boff || !d->read_len || !d->bases) { if (h) h->err = "bad dgc_daz_db"; return DGC_E_INVALID; }
    if (h->submitted != h->collected) { h->err = "dgc_daz_set_db while batches are in flight"; return DGC_E_STATE; }
    for (uint32_t i = 0; i < d->n_reads; i++)
        if (d->read_boff[i] > d->bases_bytes || ((uint64_t)d->read_len[i] + 3) / 4 > d->bases_bytes - d->read_boff[i]) { h->err = "read outside the base array"; return DGC_E_INVALID; }
    cudaSetDevice(h->device);
    cudaError_t e;
    if ((e = h->daz_bases.ensure(d->bases_bytes + 16)) != cudaSuccess || (e = h->daz_boff.ensure(8ull * d->n_reads)) != cudaSuccess ||
        (e = h->daz_rlen.ensure(4ull * d->n_reads)) != cudaSuccess ||
        (e = cudaMemcpy(h->daz_bases.p, d->bases, d->bases_bytes, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(h->daz_boff.p, d->read_boff, 8ull * d->n_reads, cudaMemcpyHostToDevice)) != cudaSuccess ||
        (e = cudaMemcpy(h->daz_rlen.p, d->read_len, 4ull * d->n_reads, cudaMemcpyHostToDevice)) != cudaSuccess) {
        h->err = cudaGetErrorString(e);
        return DGC_E_CUDA;
    }
    h->daz_h_rlen.assign(d->read_len, d->read_len + d->n_reads);
    h->daz_n_reads = d->n_reads;
    return DGC_OK;
}

int dgc_submit_daz(dgc_handle* h, const dgc_daz_batch* batch) {
    if (!h) return DGC_E_INVALID;
    int rc = validate_daz(h, batch, h->err);
    if (rc) return rc;
    int si = -1;
    {
        std::lock_guard<std::mutex> lk(h->q_mu);
        if (h->inflight.size() >= (size_t)DGC_MAX_INFLIGHT) { h->err = "too many batches in flight: collect first"; return DGC_E_STATE; }
        for (int i = 0; i < NSLOT && si < 0; i++)
            if (!h->busy[i]) si = i;
        h->busy[si] = true;
        h->inflight.push_back(si);
    }
    Slot& s = h->slots[si];
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::IDLE || s.state == Slot::DONE; });
        s.pending_daz = *batch;
        s.is_daz = true;
        s.is_pre = false;
        s.seq = h->submitted;
        s.state = Slot::QUEUED;
        s.cv.notify_all();
    }
    h->submitted++;
    return DGC_OK;
}

int dgc_debug_daz_strings(dgc_handle* h, const dgc_daz_batch* batch, uint8_t* out_q, uint8_t* out_t, uint64_t out_cap,
                          uint64_t* out_off, uint32_t* out_len) {
    if (!h || !batch || !out_off || !out_len) return DGC_E_INVALID;
    if (h->submitted != h->collected) { h->err = "debug call while batches are in flight"; return DGC_E_STATE; }
    int rc = validate_daz(h, batch, h->err);
    if (rc) return rc;
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    begin_run(s);
    s.seq = h->h2d_turn;                 // build_daz_batch hands the H2D turn on: keep the counter where it was afterwards
    const uint64_t turn = h->h2d_turn;
    rc = build_daz_batch(s, batch, true);
    release_h2d_turn_now(h, turn);
    if (rc) { h->err = s.err; return rc; }
    const size_t O = batch->n_ovls;
    std::vector<uint64_t> qo(O), to(O);
    std::vector<uint32_t> ln(O);
    cudaError_t e = cudaSuccess;
    if (O) {
        e = cudaMemcpy(qo.data(), s.db.aln_q_off.p, 8 * O, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(to.data(), s.db.aln_t_off.p, 8 * O, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(ln.data(), s.db.aln_len.p, 4 * O, cudaMemcpyDeviceToHost);
    }
    uint64_t pos = 0;
    for (size_t k = 0; k < O && e == cudaSuccess; k++) {
        if (pos + ln[k] > out_cap) { h->err = "output buffers too small"; return DGC_E_INVALID; }
        out_off[k] = pos; out_len[k] = ln[k];
        e = cudaMemcpy(out_q + pos, (const u8*)s.db.arena.p + qo[k], ln[k], cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(out_t + pos, (const u8*)s.db.arena.p + to[k], ln[k], cudaMemcpyDeviceToHost);
        pos += ln[k];
    }
    if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return DGC_E_CUDA; }
    return DGC_OK;
}

int dgc_submit_pre(dgc_handle* h, const dgc_pre_batch* batch) {
    if (!h) return DGC_E_INVALID;
    int rc = validate_pre(batch, h->err);
    if (rc) return rc;
    int si = -1;
    {
        std::lock_guard<std::mutex> lk(h->q_mu);
        if (h->inflight.size() >= (size_t)DGC_MAX_INFLIGHT) { h->err = "too many batches in flight: collect first"; return DGC_E_STATE; }
        for (int i = 0; i < NSLOT && si < 0; i++)
            if (!h->busy[i]) si = i;
        h->busy[si] = true;
        h->inflight.push_back(si);
    }
    Slot& s = h->slots[si];
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::IDLE || s.state == Slot::DONE; });
        s.pending_pre = *batch;
        s.is_pre = true;
        s.is_daz = false;
        s.seq = h->submitted;
        s.state = Slot::QUEUED;
        s.cv.notify_all();
    }
    h->submitted++;
    return DGC_OK;
}

int dgc_debug_pre_strings(dgc_handle* h, const dgc_pre_batch* batch, uint8_t* out_q, uint8_t* out_t, uint64_t out_cap,
                          uint64_t* out_off, uint32_t* out_len, uint32_t* out_start) {
    if (!h || !batch || !out_off || !out_len || !out_start) return DGC_E_INVALID;
    if (h->submitted != h->collected) { h->err = "debug call while batches are in flight"; return DGC_E_STATE; }
    int rc = validate_pre(batch, h->err);
    if (rc) return rc;
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    begin_run(s);
    const uint64_t turn = h->h2d_turn;
    s.seq = turn;
    rc = build_pre_batch(s, batch);
    release_h2d_turn_now(h, turn);
    if (rc) { h->err = s.err; return rc; }
    const size_t R = batch->n_recs;
    std::vector<uint64_t> qo(R), to(R);
    std::vector<uint32_t> ln(R);
    cudaError_t e = cudaSuccess;
    if (R) {
        e = cudaMemcpy(qo.data(), s.db.aln_q_off.p, 8 * R, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(to.data(), s.db.aln_t_off.p, 8 * R, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(ln.data(), s.db.aln_len.p, 4 * R, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(out_start, s.db.aln_start.p, 4 * R, cudaMemcpyDeviceToHost);
    }
    uint64_t pos = 0;
    for (size_t k = 0; k < R && e == cudaSuccess; k++) {
        if (pos + ln[k] > out_cap) { h->err = "output buffers too small"; return DGC_E_INVALID; }
        out_off[k] = pos; out_len[k] = ln[k];
        if (ln[k]) {
            e = cudaMemcpy(out_q + pos, (const u8*)s.db.arena.p + qo[k], ln[k], cudaMemcpyDeviceToHost);
            if (e == cudaSuccess) e = cudaMemcpy(out_t + pos, (const u8*)s.db.arena.p + to[k], ln[k], cudaMemcpyDeviceToHost);
        }
        pos += ln[k];
    }
    if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return DGC_E_CUDA; }
    return DGC_OK;
}

int dgc_consensus(dgc_handle* h, const dgc_batch* batch, dgc_results* out) {
    int rc = dgc_submit(h, batch);
    if (rc) return rc;
    return dgc_collect(h, out);
}

int dgc_upload(dgc_handle* h, const dgc_batch* batch, dgc_dbatch** out) {
    if (!h || !out) return DGC_E_INVALID;
    int rc = validate_batch(batch, h->err);
    if (rc) return rc;
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    dgc_dbatch* db = new dgc_dbatch();
    rc = upload_batch(s, batch, *db, false);
    if (rc == DGC_OK && cudaStreamSynchronize(s.stream) != cudaSuccess) { s.err = "upload sync failed"; rc = DGC_E_CUDA; }
    if (rc) { h->err = s.err; db->release(); delete db; return rc; }
    *out = db;
    return DGC_OK;
}

int dgc_run(dgc_handle* h, dgc_dbatch* db, dgc_results* out) {
    if (!h || !db || !out) return DGC_E_INVALID;
    if (h->submitted != h->collected) { h->err = "dgc_run while batches are in flight"; return DGC_E_STATE; }
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    begin_run(s);
    int rc = run_pipeline(s, *db, 2, 0xFFFFFFFFu);
    cudaStreamSynchronize(s.stream);
    finish_timings(s);
    h->last_tm = s.tm;
    if (rc) { h->err = s.err; return rc; }
    *out = s.results;
    return DGC_OK;
}

int dgc_dbatch_free(dgc_handle* h, dgc_dbatch* db) {
    if (!db) return DGC_OK;
    if (h) cudaSetDevice(h->device);
    db->release();
    delete db;
    return DGC_OK;
}

int dgc_get_timings(dgc_handle* h, dgc_timings* out) {
    if (!h || !out) return DGC_E_INVALID;
    *out = h->last_tm;
    return DGC_OK;
}

int dgc_debug_normalize(dgc_handle* h, dgc_dbatch* db, uint16_t* out_cols, uint64_t out_cap, uint64_t* out_off,
                        uint32_t* out_len, uint32_t* out_start, uint32_t* out_flags) {
    if (!h || !db) return DGC_E_INVALID;
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    begin_run(s);
    int rc = ensure_alignment_arrays(s, *db);
    if (rc == DGC_OK) rc = launch_normalize(s, *db);
    if (rc) { h->err = s.err; return rc; }
    if (out_cap < db->norm_cols) { h->err = "out_cols too small"; return DGC_E_INVALID; }
    cudaError_t e = cudaStreamSynchronize(s.stream);
    if (e == cudaSuccess && db->norm_cols) e = cudaMemcpy(out_cols, s.norm.p, 2 * db->norm_cols, cudaMemcpyDeviceToHost);
    if (e == cudaSuccess && db->A) {
        e = cudaMemcpy(out_off, s.o_off.p, 8ull * db->A, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(out_len, s.o_len.p, 4ull * db->A, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(out_start, s.o_start.p, 4ull * db->A, cudaMemcpyDeviceToHost);
        if (e == cudaSuccess) e = cudaMemcpy(out_flags, s.o_flags.p, 4ull * db->A, cudaMemcpyDeviceToHost);
    }
    if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return DGC_E_CUDA; }
    return DGC_OK;
}

int dgc_debug_graph(dgc_handle* h, dgc_dbatch* db, uint32_t target, int stage, char** text, uint64_t* n) {
    if (!h || !db || !text || !n || target >= db->T) return DGC_E_INVALID;
    cudaSetDevice(h->device);
    Slot& s = h->slots[0];
    begin_run(s);
    int rc = run_pipeline(s, *db, stage ? 1 : 0, target);
    if (rc) { h->err = s.err; return rc; }
    uint32_t status = 0, nv = 0;
    cudaMemcpy(&status, s.t_status.as<u32>() + target, 4, cudaMemcpyDeviceToHost);
    std::string out;
    if (status != T_OK) {
        out = "status " + std::to_string(status) + "\n";
    } else {
        cudaMemcpy(&nv, s.t_nv.as<u32>() + target, 4, cudaMemcpyDeviceToHost);
        uint32_t ptop = 0;
        cudaMemcpy(&ptop, s.t_pool_top.as<u32>() + target, 4, cudaMemcpyDeviceToHost);
        const uint32_t L = db->h_tlen[target], nbb = L + 2;
        // single-target run: the arena offset is 0; records, then the pool, ptop = first unused 8-byte word
        std::vector<uint64_t> AR((size_t)ptop + 16);
        cudaError_t e = cudaMemcpy(AR.data(), s.rec.p, 8 * (size_t)ptop, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) { h->err = cudaGetErrorString(e); return DGC_E_CUDA; }
        const uint8_t* base = (const uint8_t*)AR.data();
        auto rec = [&](uint32_t v) -> const uint8_t* { return base + (size_t)REC_UNIT * gv_handle(v, nbb); };
        auto outs = [&](uint32_t v) -> const OSlot* { return (const OSlot*)(base + 8 * (size_t)((const Hdr*)rec(v))->out_ptr); };
        auto ins = [&](uint32_t v) -> const uint32_t* { return (const uint32_t*)(base + 8 * (size_t)((const Hdr*)rec(v))->in_ptr); };
        char buf[160];
        for (uint32_t v = 0; v < nv; v++) {
            const Hdr* nd = (const Hdr*)rec(v);
            const uint32_t fl = (nd->bfv >> 8) & 0xFF;
            snprintf(buf, sizeof buf, "%u %c %d %d %d %d |", v, (char)(nd->bfv & 0xFF), nd->weight, v < nbb ? nd->cov : 0,
                     (int)((fl & NF_BACKBONE) != 0), (int)((fl & NF_DELETED) != 0));
            out += buf;
            const OSlot* o = outs(v);
            for (uint32_t i = 0; i < (nd->deg >> 16); i++) {
                snprintf(buf, sizeof buf, " %u:%d:%d", gv_vertex(o[i].dst, nbb), (int)(o[i].cv & ~VISBIT), (int)((o[i].cv & VISBIT) != 0));
                out += buf;
            }
            out += " |";
            const uint32_t* in = ins(v);
            for (uint32_t i = 0; i < (nd->deg & 0xFFFFu); i++) {
                const uint32_t sv = gv_vertex(in[i], nbb);
                int cnt = -1, vis = -1;
                if (sv < nv) {
                    const Hdr* hs = (const Hdr*)rec(sv);
                    const OSlot* os = outs(sv);
                    for (uint32_t k = 0; k < (hs->deg >> 16); k++)
                        if (os[k].dst == gv_handle(v, nbb)) { cnt = (int)(os[k].cv & ~VISBIT); vis = (int)((os[k].cv & VISBIT) != 0); break; }
                }
                snprintf(buf, sizeof buf, " %u:%d:%d", sv, cnt, vis);
                out += buf;
            }
            out += "\n";
        }
    }
    char* p = (char*)malloc(out.size() + 1);
    if (!p) return DGC_E_NOMEM;
    memcpy(p, out.data(), out.size());
    p[out.size()] = 0;
    *text = p;
    *n = out.size();
    return DGC_OK;
}

void dgc_debug_free(char* text) { free(text); }

// The `-a` aligner without a single allocation (header: no counterpart in the reference).
// The band schedule and the wave planning are build_pre_batch's, on stack buffers; the kernel is
// k_pre_align itself, its scratch the static array above.  No cudaMalloc / cudaFree / std::vector here.
int dgc_pre_align_noalloc(int device, const dgc_pre_dev* a, void* stream, char* err, uint64_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (!a) return na_fail(err, err_cap, DGC_E_INVALID, "null arguments");
    const uint32_t R = a->n_recs;
    if (R > DGC_NOALLOC_MAX_RECS)
        return na_fail(err, err_cap, DGC_E_INVALID, "n_recs %u is beyond DGC_NOALLOC_MAX_RECS (%u): split the batch", R, DGC_NOALLOC_MAX_RECS);
    if (R == 0) return DGC_OK;
    if (!a->rec_start || !a->rec_strand || !a->rec_qlen || !a->rec_tseglen || !a->rec_tlen || !a->rec_q_off || !a->rec_t_off || !a->arena ||
        !a->out_arena || !a->out_q_off || !a->out_t_off || !a->out_len || !a->out_start || !a->out_tbegin || !a->out_tend || !a->out_state)
        return na_fail(err, err_cap, DGC_E_INVALID, "null device pointer in dgc_pre_dev");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0) return na_fail(err, err_cap, DGC_E_NODEVICE, "no CUDA device (there is no CPU path)");
    if (device < 0 || device >= ndev || device >= NA_MAX_DEVICES) return na_fail(err, err_cap, DGC_E_INVALID, "device %d out of range", device);
    int prev_dev = device;
    cudaGetDevice(&prev_dev);
    cudaError_t e = cudaSetDevice(device);
    unsigned char* scratch = nullptr;
    if (e == cudaSuccess) e = na_scratch(device, &scratch);   // cudaGetSymbolAddress, once per device, with it current
    if (e != cudaSuccess) {
        if (prev_dev != device) cudaSetDevice(prev_dev);
        return na_fail(err, err_cap, DGC_E_CUDA, "static scratch: %s", cudaGetErrorString(e));
    }
    cudaStream_t st = (cudaStream_t)stream;
    u32* d_ids = (u32*)scratch;
    u64* d_tboff = (u64*)(scratch + NA_TBOFF_AT);
    u8* d_tb = scratch + NA_LIST_BYTES;
    const uint64_t tb_budget = DGC_NOALLOC_SCRATCH_BYTES;

    uint32_t h_qlen[DGC_NOALLOC_MAX_RECS], h_tlen[DGC_NOALLOC_MAX_RECS], h_state[DGC_NOALLOC_MAX_RECS];
    uint32_t pending[DGC_NOALLOC_MAX_RECS], h_ids[DGC_NOALLOC_MAX_RECS];
    uint64_t h_tboff[DGC_NOALLOC_MAX_RECS];
    int rc = DGC_OK;
#define NA_CU(call)                                                                                              \
    do {                                                                                                         \
        if (rc == DGC_OK && (e = (call)) != cudaSuccess) rc = na_fail(err, err_cap, DGC_E_CUDA, "%s: %s", #call, cudaGetErrorString(e)); \
    } while (0)
    NA_CU(cudaMemcpyAsync(h_qlen, a->rec_qlen, 4ull * R, cudaMemcpyDeviceToHost, st));
    NA_CU(cudaMemcpyAsync(h_tlen, a->rec_tseglen, 4ull * R, cudaMemcpyDeviceToHost, st));
    NA_CU(cudaMemsetAsync(a->out_state, 0, 4ull * R, st));
    NA_CU(cudaStreamSynchronize(st));

    PreArgs pa;
    pa.rec_start = a->rec_start; pa.rec_tlen = a->rec_tlen; pa.aln_start = a->out_start;
    pa.arena_in = a->arena; pa.q_off = (const u64*)a->rec_q_off; pa.t_off = (const u64*)a->rec_t_off; pa.q_len = a->rec_qlen; pa.t_len = a->rec_tseglen;
    pa.strand = a->rec_strand; pa.tb = d_tb; pa.tb_off = d_tboff; pa.rec_ids = d_ids;
    pa.arena_out = a->out_arena; pa.out_q_off = (const u64*)a->out_q_off; pa.out_t_off = (const u64*)a->out_t_off; pa.out_len = a->out_len;
    pa.t_begin = a->out_tbegin; pa.t_end = a->out_tend; pa.state = a->out_state;

    uint32_t n_pending = R;
    for (uint32_t k = 0; k < R; k++) pending[k] = k;
    for (uint32_t band = 64; rc == DGC_OK && n_pending && band <= PA_MAX_BAND; band *= 2) {
        const uint32_t warps = band <= 512 ? 4u : 1u;
        const size_t smem = (size_t)warps * 2 * (2 * band + 1 + 2 * PA_G) * sizeof(int);
        if (smem > 48 * 1024) NA_CU(cudaFuncSetAttribute(k_pre_align, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        uint32_t done = 0;
        while (rc == DGC_OK && done < n_pending) {   // waves that fit the static move-matrix area
            uint64_t tb = 0;
            uint32_t n = 0;
            while (done + n < n_pending) {
                const uint32_t k = pending[done + n];
                const uint64_t full = (uint64_t)std::max(h_qlen[k], h_tlen[k]) + 1;
                const uint64_t W = 2 * std::min<uint64_t>(band, full) + 1;
                const uint64_t need = (((uint64_t)h_qlen[k] + 1) * W + 15) & ~15ull;
                if (need > tb_budget) {
                    rc = na_fail(err, err_cap, DGC_E_INVALID,
                                 "record %u (%u x %u bases) needs a move matrix of %llu bytes at band %u: beyond the static scratch "
                                 "(DGC_NOALLOC_SCRATCH_BYTES = %u); use dgc_submit_pre for it",
                                 k, h_qlen[k], h_tlen[k], (unsigned long long)need, band, (unsigned)DGC_NOALLOC_SCRATCH_BYTES);
                    break;
                }
                if (tb + need > tb_budget) break;
                h_tboff[n] = tb; h_ids[n] = k;
                tb += need;
                n++;
            }
            if (rc != DGC_OK) break;
            NA_CU(cudaMemcpyAsync(d_tboff, h_tboff, 8ull * n, cudaMemcpyHostToDevice, st));
            NA_CU(cudaMemcpyAsync(d_ids, h_ids, 4ull * n, cudaMemcpyHostToDevice, st));
            if (rc == DGC_OK) {
                pa.n_launch = n; pa.band = band; pa.warps = warps;
                k_pre_align<<<(n + warps - 1) / warps, warps * 32, smem, st>>>(pa);
                NA_CU(cudaGetLastError());
            }
            NA_CU(cudaStreamSynchronize(st));   // the lists and the move matrices are reused by the next wave
            done += n;
        }
        NA_CU(cudaMemcpyAsync(h_state, a->out_state, 4ull * R, cudaMemcpyDeviceToHost, st));
        NA_CU(cudaStreamSynchronize(st));
        if (rc != DGC_OK) break;
        uint32_t m = 0;
        for (uint32_t i = 0; i < n_pending; i++)
            if (h_state[pending[i]] == 2) pending[m++] = pending[i];
        n_pending = m;
    }
    if (rc == DGC_OK && n_pending) {   // beyond the widest band: zero-column alignments
        NA_CU(cudaMemcpyAsync(d_ids, pending, 4ull * n_pending, cudaMemcpyHostToDevice, st));
        if (rc == DGC_OK) {
            k_na_hopeless<<<(n_pending + 127) / 128, 128, 0, st>>>(d_ids, n_pending, a->rec_start, a->out_len, a->out_start, a->out_tbegin, a->out_tend,
                                                                  a->out_state);
            NA_CU(cudaGetLastError());
        }
        NA_CU(cudaStreamSynchronize(st));
    }
#undef NA_CU
    if (prev_dev != device) cudaSetDevice(prev_dev);
    return rc;
}

}  // extern "C"
