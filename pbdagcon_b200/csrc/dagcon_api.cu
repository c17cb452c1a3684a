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

#include <algorithm>
#include <condition_variable>
#include <cstdio>
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
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMalloc(&p, want);
        if (e != cudaSuccess) { e = cudaMalloc(&p, bytes + 256); want = bytes + 256; }
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return (T*)p; }
};
struct HostBuf {  // pinned
    void* p = nullptr;
    size_t cap = 0;
    cudaError_t ensure(size_t bytes) {
        if (bytes <= cap) return cudaSuccess;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        size_t want = bytes + bytes / 8 + 256;
        cudaError_t e = cudaMallocHost(&p, want);
        if (e == cudaSuccess) cap = want;
        return e;
    }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
    template <class T> T* as() const { return (T*)p; }
};

}  // namespace

// A batch resident in HBM: the dgc_batch arrays plus the derived normalisation slots.
struct dgc_dbatch {
    uint32_t T = 0, A = 0;
    bool has_bb = false;
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

enum Phase { PH_H2D, PH_NORM, PH_COUNT, PH_BUILD, PH_MERGE, PH_PATH, PH_COMPACT, PH_D2H, PH_N };

struct Slot {
    dgc_handle* h = nullptr;
    int index = 0;
    cudaStream_t stream = nullptr;
    dgc_dbatch db;  // device copy of a submitted batch
    HostBuf h_norm_off;
    // per-alignment work arrays
    DevBuf o_off, o_len, o_start, o_flags, o_nI, o_nE, o_tkept, norm;
    // per-target arrays
    DevBuf t_status, t_V, t_E, t_cols, t_ne, t_nv, t_cons_len, t_nfrag, t_cons_coff, t_frag_coff, totals;
    DevBuf rec_off, pos_off, pool_cap, t_pool_top, cons_off, frag_off, frag_cap, gs_off, gs_cap;
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
    std::string err;
    Slot slots[NSLOT];
    uint64_t submitted = 0, collected = 0;
    dgc_timings last_tm{};
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
            if (b->tgt_bb_off && b->tgt_bb_off[t] + b->tgt_tlen[t] > b->arena_bytes) { err = "backbone outside arena"; return DGC_E_INVALID; }
        }
    } else if (b->n_alns) { err = "alignments without targets"; return DGC_E_INVALID; }
    for (uint32_t k = 0; k < b->n_alns; k++) {
        uint64_t n = b->aln_len[k];
        if (b->aln_q_off[k] + n > b->arena_bytes || b->aln_t_off[k] + n > b->arena_bytes) {
            err = "alignment string outside arena"; return DGC_E_INVALID;
        }
    }
    return DGC_OK;
}

// H2D of a batch into db (async on s.stream; host arrays must stay valid until the stream syncs)
int upload_batch(Slot& s, const dgc_batch* b, dgc_dbatch& db, bool timed) {
    db.T = b->n_targets; db.A = b->n_alns; db.has_bb = b->tgt_bb_off != nullptr; db.arena_bytes = b->arena_bytes;
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
    if (b->arena_bytes) CU(cudaMemcpyAsync(db.arena.p, b->arena, b->arena_bytes, cudaMemcpyHostToDevice, st));
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

int launch_normalize(Slot& s, const dgc_dbatch& db) {
    const dgc_opts& o = s.h->opts;
    NormArgs na;
    na.arena = db.arena.as<u8>(); na.aln_start = db.aln_start.as<u32>(); na.aln_len = db.aln_len.as<u32>();
    na.aln_q_off = db.aln_q_off.as<u64>(); na.aln_t_off = db.aln_t_off.as<u64>(); na.aln_norm_off = db.aln_norm_off.as<u64>();
    na.n_alns = db.A; na.min_len = o.min_len; na.trim = o.trim; na.flags = o.flags;
    na.norm = s.norm.as<u16>(); na.o_off = s.o_off.as<u64>(); na.o_len = s.o_len.as<u32>(); na.o_start = s.o_start.as<u32>();
    na.o_flags = s.o_flags.as<u32>(); na.o_nI = s.o_nI.as<u32>(); na.o_nE = s.o_nE.as<u32>(); na.o_tkept = s.o_tkept.as<u32>();
    if (db.A) {
        if (o.flags & DGC_F_NO_NORMALIZE) {
            na.retry_only = 0;
            k_normalize_trim<<<(db.A + 127) / 128, 128, 0, s.stream>>>(na);
            s.tm.launches++;
        } else {
            na.retry_only = 1;
            k_normalize_warp<<<(db.A + NW_WARPS - 1) / NW_WARPS, NW_WARPS * 32, 0, s.stream>>>(na);
            k_normalize_trim<<<(db.A + 127) / 128, 128, 0, s.stream>>>(na);  // the few the warp kernel handed back
            s.tm.launches += 2;
        }
    }
    CU(cudaGetLastError());
    return DGC_OK;
}

struct Chunk { uint32_t first, count; uint64_t nUnits, nPos, nCons, nFrag, nGs; uint32_t maxL2; };

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

    span_begin(s, PH_NORM);
    rc = launch_normalize(s, db);
    if (rc) return rc;
    span_end(s);

    span_begin(s, PH_COUNT);
    PlanArgs pa;
    pa.n_targets = db.T; pa.min_cov = o.min_cov; pa.tgt_tlen = db.tgt_tlen.as<u32>(); pa.tgt_aln_begin = db.tgt_aln_begin.as<u32>();
    pa.a_start = s.o_start.as<u32>(); pa.a_flags = s.o_flags.as<u32>(); pa.a_nI = s.o_nI.as<u32>(); pa.a_nE = s.o_nE.as<u32>();
    pa.a_tkept = s.o_tkept.as<u32>(); pa.a_len = s.o_len.as<u32>();
    pa.tgt_status = s.t_status.as<u32>(); pa.tgt_V = s.t_V.as<u32>(); pa.tgt_E = s.t_E.as<u32>(); pa.tgt_cols = s.t_cols.as<u64>();
    k_target_plan<<<(db.T + 127) / 128, 128, 0, st>>>(pa);
    s.tm.launches++;
    CU(cudaGetLastError());
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

    // ---- host planning: arena offsets per target (128-byte units), chunks that fit the scratch budget
    uint64_t* p_rec = s.hp_plan_out.as<uint64_t>();
    uint64_t* p_pos = p_rec + (T + 1);
    uint64_t* p_cons = p_pos + (T + 1);
    uint64_t* p_frag = p_cons + (T + 1);
    uint64_t* p_gs = p_frag + (T + 1);
    uint32_t* p_fcap = (uint32_t*)(p_gs + (T + 1));
    uint32_t* p_gcap = p_fcap + (T + 1);
    uint32_t* p_pcap = p_gcap + (T + 1);
    uint64_t budget = o.scratch_bytes ? o.scratch_bytes : (32ull << 30);
    std::vector<Chunk> chunks;
    Chunk cur{0, 0, 0, 0, 0, 0, 0, 0};
    auto chunk_bytes = [](const Chunk& c) { return 128 * c.nUnits + 9 * c.nPos + c.nCons + 8 * c.nFrag + 4 * c.nGs; };
    uint64_t out_cons_cap = 0, out_frag_cap = 0;
    for (size_t t = 0; t < T; t++) {
        const bool ok = hp_status[t] == T_OK && (only_target == 0xFFFFFFFFu || only_target == t);
        if (!ok && hp_status[t] == T_OK) hp_status[t] = T_SKIPPED;  // debug single-target run
        const uint64_t V = ok ? hp_V[t] : 0, E = ok ? hp_E[t] : 0, L2 = ok ? (uint64_t)db.h_tlen[t] + 2 : 0;
        const uint32_t A = db.h_aln_begin[t + 1] - db.h_aln_begin[t];
        // pool of list blocks behind the records: enter's out-list and exit's in-list always, the rest only when a
        // list outgrows its record; a multiple of 16 words keeps every target's arena 128-byte aligned
        const uint64_t pcap = ok ? (4ull * (A + 8) + E / 2 + 256 + 15) & ~15ull : 0;
        const uint64_t units = ok ? L2 + V + pcap / 16 : 0;  // backbone records take two units
        const uint64_t fcap = ok ? V / ((uint64_t)o.min_len + 1) + 2 : 0, gcap = ok ? 8ull * (A + 2) + 64 : 0;
        const uint64_t consb = (V + 15) & ~15ull;
        Chunk one{0, 1, units, L2, consb, fcap, gcap, (uint32_t)L2};
        if (16 * units > 0xFFFFFFF0ull) { s.err = "one target's graph exceeds the 32 GiB a list pointer can address"; return DGC_E_NOMEM; }
        if (chunk_bytes(one) > budget) { s.err = "one target needs more than scratch_bytes of graph arena"; return DGC_E_NOMEM; }
        Chunk nxt = cur;
        nxt.nUnits += units; nxt.nPos += L2; nxt.nCons += consb; nxt.nFrag += fcap; nxt.nGs += gcap; nxt.count++;
        if (cur.count && (chunk_bytes(nxt) > budget || cur.count >= (1u << 20))) {
            chunks.push_back(cur);
            cur = Chunk{(uint32_t)t, 0, 0, 0, 0, 0, 0, 0};
        }
        p_rec[t] = cur.nUnits; p_pos[t] = cur.nPos; p_cons[t] = cur.nCons; p_frag[t] = cur.nFrag; p_gs[t] = cur.nGs;
        p_fcap[t] = (uint32_t)fcap; p_gcap[t] = (uint32_t)gcap; p_pcap[t] = (uint32_t)pcap;
        cur.nUnits += units; cur.nPos += L2; cur.nCons += consb; cur.nFrag += fcap; cur.nGs += gcap; cur.count++;
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
        if (9ull * c.maxL2 > h->build_smem_max) need_gacc = true;
    }
    s.tm.chunks = (uint32_t)chunks.size();
    CU(s.rec.ensure(128 * mUnits + 256));
    CU(s.cons.ensure(mCons + 64)); CU(s.frags.ensure(8 * mFrag + 64)); CU(s.gscratch.ensure(4 * mGs + 64));
    if (need_gacc) CU(s.gacc.ensure(9 * mPos + 64));
    CU(s.out_cons.ensure(out_cons_cap + 64)); CU(s.out_frags.ensure(8 * out_frag_cap + 64));
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
        Build2Args ba;
        ba.first = c.first; ba.tgt_tlen = db.tgt_tlen.as<u32>(); ba.tgt_aln_begin = db.tgt_aln_begin.as<u32>();
        ba.tgt_bb_off = db.has_bb ? db.tgt_bb_off.as<u64>() : nullptr; ba.arena = db.arena.as<u8>();
        ba.tgt_status = s.t_status.as<u32>(); ba.rec_off = s.rec_off.as<u64>(); ba.tgt_V = s.t_V.as<u32>();
        ba.pool_cap = s.pool_cap.as<u32>(); ba.acc_off = s.pos_off.as<u64>();
        ba.norm = s.norm.as<u16>(); ba.a_off = s.o_off.as<u64>(); ba.a_len = s.o_len.as<u32>();
        ba.a_start = s.o_start.as<u32>(); ba.a_flags = s.o_flags.as<u32>();
        ba.rec = s.rec.as<u8>();
        ba.tgt_status_out = s.t_status.as<u32>(); ba.tgt_nv = s.t_nv.as<u32>(); ba.tgt_pool_top = s.t_pool_top.as<u32>();
        // per-position running sums: 9 B per position in shared memory when the longest backbone of the chunk fits
        ba.smem_positions = (uint32_t)std::min<uint64_t>(c.maxL2, h->build_smem_max / 9);
        ba.gacc = s.gacc.as<u8>();
        const size_t build_smem = (9ull * ba.smem_positions + 15) & ~15ull;
        span_begin(s, PH_BUILD);
        k_build_graph2<<<c.count, BT, build_smem, st>>>(ba);
        s.tm.launches++;
        CU(cudaGetLastError());
        span_end(s);
        if (stop_after == 0) continue;

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
        const uint32_t warps_per_block = WARP_KERNEL_THREADS / 32;   // one warp per target
        const uint32_t grid = (c.count + warps_per_block - 1) / warps_per_block;
        if (!(o.flags & DGC_F_NO_MERGE)) {
            span_begin(s, PH_MERGE);
            k_merge_nodes2<<<grid, WARP_KERNEL_THREADS, 0, st>>>(sa);
            s.tm.launches++;
            CU(cudaGetLastError());
            span_end(s);
        }
        if (stop_after < 2) continue;
        span_begin(s, PH_PATH);
        k_best_path2<<<grid, WARP_KERNEL_THREADS, 0, st>>>(sa);
        s.tm.launches++;
        CU(cudaGetLastError());
        span_end(s);
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
    }
    if (stop_after < 2) { CU(cudaStreamSynchronize(st)); return DGC_OK; }

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
    const uint32_t* rs = s.r_status.as<uint32_t>();
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

void finish_timings(Slot& s) {
    float acc[PH_N] = {0};
    for (auto& sp : s.spans) {
        float ms = 0;
        if (cudaEventElapsedTime(&ms, sp.second.first, sp.second.second) == cudaSuccess) acc[sp.first] += ms;
    }
    s.tm.ms_h2d = acc[PH_H2D]; s.tm.ms_normalize = acc[PH_NORM]; s.tm.ms_count = acc[PH_COUNT]; s.tm.ms_build = acc[PH_BUILD];
    s.tm.ms_merge = acc[PH_MERGE]; s.tm.ms_bestpath = acc[PH_PATH]; s.tm.ms_compact = acc[PH_COMPACT]; s.tm.ms_d2h = acc[PH_D2H];
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

int slot_process(Slot& s, const dgc_batch* b) {
    begin_run(s);
    int rc = upload_batch(s, b, s.db, true);
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
    {   // opt in to large dynamic shared memory for the build kernel (static use is ~4.2 KB)
        size_t optin = prop.sharedMemPerBlockOptin > 8192 ? prop.sharedMemPerBlockOptin - 8192 : 0;
        optin = std::min<size_t>(optin, 200 * 1024);
        if (cudaFuncSetAttribute(k_build_graph2, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)optin) == cudaSuccess)
            h->build_smem_max = optin;
        else { cudaGetLastError(); h->build_smem_max = 40 * 1024; }
    }
    for (int i = 0; i < NSLOT; i++) {
        Slot& s = h->slots[i];
        s.h = h; s.index = i;
        if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess) {
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
                          &s.rec, &s.cons, &s.frags, &s.gscratch, &s.gacc, &s.out_cons, &s.out_frags})
            b->release();
        for (HostBuf* b : {&s.h_norm_off, &s.hp_plan_in, &s.hp_plan_out, &s.r_status, &s.r_cons_coff, &s.r_frag_coff, &s.r_totals,
                           &s.r_cons, &s.r_frags, &s.r_nfrag})
            b->release();
        s.events.destroy();
        cudaStreamDestroy(s.stream);
    }
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
    if (h->submitted - h->collected >= NSLOT - 1) { h->err = "too many batches in flight: collect first"; return DGC_E_STATE; }
    Slot& s = h->slots[h->submitted % NSLOT];
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::IDLE || s.state == Slot::DONE; });
        s.pending = *batch;
        s.state = Slot::QUEUED;
        s.cv.notify_all();
    }
    h->submitted++;
    return DGC_OK;
}

int dgc_collect(dgc_handle* h, dgc_results* out) {
    if (!h || !out) return DGC_E_INVALID;
    if (h->collected == h->submitted) { h->err = "nothing submitted"; return DGC_E_STATE; }
    Slot& s = h->slots[h->collected % NSLOT];
    int rc;
    {
        std::unique_lock<std::mutex> lk(s.mu);
        s.cv.wait(lk, [&] { return s.state == Slot::DONE; });
        rc = s.rc;
        s.state = Slot::IDLE;
    }
    h->collected++;
    h->last_tm = s.tm;
    if (rc) { h->err = s.err; return rc; }
    *out = s.results;
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

}  // extern "C"
