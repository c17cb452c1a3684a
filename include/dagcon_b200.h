/* dagcon_b200.h -- C ABI of libdagcon_b200.so: the DAGCon consensus hot path on B200 (sm_100a).
 *
 * The reference (verdurin/pbdagcon @ 3c382f2) has no FFI; its boundary for this path is the C++
 * call pattern of the Consensus functor.  Each entry point below names the reference interface
 * it replaces (file:line relative to the reference tree).  All functions return 0 on success
 * or a negative DGC_E_* code; none throws, calls exit(), or falls back to a CPU path: when no
 * CUDA device is usable dgc_create fails.
 *
 * Batch = what AlnProvider::nextTarget hands to the consumer (src/cpp/AlnProvider.hpp:39,43),
 * for many targets at once: per target {tlen, [optional backbone string], its alignments in
 * input order}; per alignment {1-based start, raw column count, qstr, tstr} exactly as the
 * provider produced them (dagcon::Alignment, src/cpp/Alignment.hpp:12-42).  Normalisation,
 * trimming, the min-length / min-coverage filters, graph construction, node merging, the
 * best-path DP and fragment extraction all run on the device.
 */
#ifndef DAGCON_B200_H
#define DAGCON_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DGC_ABI_VERSION 2
#define DGC_MAX_INFLIGHT 4 /* batches that may be submitted and not yet collected */

/* error codes (function return values) */
#define DGC_OK 0
#define DGC_E_INVALID (-1)   /* bad argument / malformed batch */
#define DGC_E_CUDA (-2)      /* CUDA runtime error; see dgc_last_error */
#define DGC_E_NOMEM (-3)     /* a single target does not fit the scratch budget */
#define DGC_E_STATE (-4)     /* collect without submit, etc. */
#define DGC_E_NODEVICE (-5)  /* no usable CUDA device (there is no CPU fallback) */
#define DGC_E_DAZ (-6)       /* dgc_submit_daz: an overlap the device realignment cannot take (trace values out of range) */

/* per-target status (dgc_results.tgt_status) */
#define DGC_T_OK 0
#define DGC_T_SKIPPED 1       /* fewer than min_cov alignments: main.cpp:66-72,118; dazcon.cpp:69 */
#define DGC_T_BAD_ALIGNMENT 2 /* an alignment runs off the backbone (reference: out-of-bounds
                                 vertex access, AlnGraphBoost.cpp:73-79) or start < 1 */
#define DGC_T_TOO_MANY_ALNS 3 /* more than 65000 alignments in one target (16-bit degrees) */
#define DGC_T_INTERNAL 4      /* scratch overflow / inconsistent graph (cycle); never on valid input */

/* dgc_opts.flags */
#define DGC_F_NO_NORMALIZE 1u /* skip normalizeGaps (tests: AlnGraphBoostTest.cpp:11-47 feeds raw alignments) */
#define DGC_F_NO_TRIM 2u      /* skip trimAln */
#define DGC_F_NO_MERGE 4u     /* skip mergeNodes */
#define DGC_F_SINGLE 8u       /* consensus(minWeight) -> one longest run (AlnGraphBoost.cpp:285-325)
                                 instead of consensus(seqs,minWeight,minLen) (:327-373) */
#define DGC_F_SPLIT_SWEEPS 512u /* tuning/diagnostic: mergeNodes and bestPath as two launches (per-kernel timings) */
#define DGC_F_NO_PREFETCH 256u /* tuning/diagnostic: the graph sweeps do not prefetch vertex records (results identical) */

/* Replaces ProgramOpts (src/cpp/ProgramOpts.hpp:8-36) for this path. */
typedef struct dgc_opts {
    uint32_t min_cov;       /* -c: min alignments per target AND consensus minWeight (main.cpp:261) */
    uint32_t min_len;       /* -m / dazcon -l: min raw alignment columns AND min fragment length */
    uint32_t trim;          /* -t: target bases trimmed from both alignment ends */
    uint32_t flags;         /* DGC_F_* */
    int32_t device;         /* CUDA device ordinal */
    uint32_t reserved0;     /* must be 0 (was targets_per_warp; the sweeps are always one warp per target) */
    uint64_t scratch_bytes; /* HBM budget for the per-chunk graph arena; 0 = default (112 GiB: 10 000 targets of 8 kb x 30x or 4 800 of 20 kb x 50x in one chunk) */
} dgc_opts;

/* dgc_batch.flags */
#define DGC_BATCH_PACKED4 1u /* qstr/tstr travel as 4-bit codes, two columns per byte (column c of a string in byte c/2 of
                                it, even columns in the low nibble): half the bytes on the PCIe link.  Codes index
                                DGC_PACK4_ALPHABET; a string holding any other character cannot be packed (send that
                                batch as ASCII).  Offsets and aln_len keep their meaning (byte offset of the string in the
                                arena, number of columns); backbone strings (tgt_bb_off) stay ASCII. */
#define DGC_PACK4_ALPHABET "-ACGTN."  /* codes 0..6 */

/* Packed targets (SoA).  Host pointers; pinned memory (dgc_host_alloc) makes the H2D copy async.
 * Offsets are byte offsets into `arena`.  Strings need not be aligned or NUL-terminated. */
typedef struct dgc_batch {
    uint32_t n_targets;
    uint32_t n_alns;
    const uint32_t* tgt_tlen;      /* [n_targets]   Alignment::tlen of alns[0] */
    const uint32_t* tgt_aln_begin; /* [n_targets+1] CSR into the alignment arrays */
    const uint64_t* tgt_bb_off;    /* [n_targets] backbone string (tlen bytes) for the
                                      AlnGraphBoost(const std::string&) ctor (dazcon.cpp:76),
                                      or NULL: AlnGraphBoost(size_t) (main.cpp:130) */
    const uint32_t* aln_start;     /* [n_alns] Alignment::start, 1-based */
    const uint32_t* aln_len;       /* [n_alns] qstr.length() == tstr.length() */
    const uint64_t* aln_q_off;     /* [n_alns] */
    const uint64_t* aln_t_off;     /* [n_alns] */
    const uint8_t* arena;
    uint64_t arena_bytes;
    uint32_t flags;                /* DGC_BATCH_* */
    uint32_t reserved;
} dgc_batch;

/* Replaces std::vector<CnsResult> per target (src/cpp/AlnGraphBoost.hpp CnsResult; filled at
 * AlnGraphBoost.cpp:327-373).  Memory is owned by the handle and stays valid until the next
 * dgc_submit / dgc_collect / dgc_consensus / dgc_run on it, or dgc_destroy. */
typedef struct dgc_results {
    uint32_t n_targets;
    uint32_t n_frags;
    const uint32_t* tgt_status;     /* [n_targets] DGC_T_* */
    const uint32_t* tgt_frag_begin; /* [n_targets+1] */
    const int32_t* frag_range0;     /* [n_frags] CnsResult::range[0] */
    const int32_t* frag_range1;     /* [n_frags] CnsResult::range[1] */
    const uint64_t* frag_seq_off;   /* [n_frags] into seq_arena */
    const uint32_t* frag_seq_len;   /* [n_frags] */
    const uint8_t* seq_arena;
    uint64_t seq_bytes;
} dgc_results;

/* device time of the last run, per kernel (CUDA events on the launching stream) */
typedef struct dgc_timings {
    float ms_h2d, ms_normalize, ms_count, ms_build, ms_merge, ms_bestpath, ms_compact, ms_d2h, ms_total;
    float ms_span;           /* first event to last event on the stream: includes host planning gaps */
    uint32_t launches;       /* kernels launched by the last run */
    uint32_t chunks;
    uint64_t arena_bytes;    /* graph arena bytes used (sum over chunks) */
    uint64_t norm_cols;      /* normalised+trimmed columns fed to the graph */
    uint64_t nodes, edges;   /* pre-merge graph size, summed over targets */
    float ms_realign;        /* dgc_submit_daz: trace-point realignment + string emission */
    uint32_t reserved1;
} dgc_timings;

typedef struct dgc_handle dgc_handle;
typedef struct dgc_dbatch dgc_dbatch; /* a batch resident in HBM */

int dgc_abi_version(void);
int dgc_device_count(void);

/* One handle per GPU and per submitting thread; replaces the N Consensus threads + the two
 * BoundedBuffers (main.cpp:248-274, dazcon.cpp:211-230, BoundedBuffer.hpp:18-50). */
int dgc_create(const dgc_opts* opts, dgc_handle** out);
int dgc_destroy(dgc_handle* h);
const char* dgc_last_error(dgc_handle* h); /* h may be NULL: error of the last failed dgc_create */

/* Pinned host staging memory for dgc_batch arrays. */
int dgc_host_alloc(void** p, uint64_t bytes);
int dgc_host_free(void* p);

/* Asynchronous pipeline: submit copies the batch H2D and enqueues the kernels on one of the
 * handle's streams (the dgc_batch struct is copied; the arrays it points to must stay untouched
 * until the matching collect);
 * collect returns results in submit order.  At most DGC_MAX_INFLIGHT batches may be in flight
 * (submit returns DGC_E_STATE beyond that); their kernels and copies overlap on separate streams.
 * Replaces AlnBuf::push / CnsBuf::pop (main.cpp:87,120,144,168). */
int dgc_submit(dgc_handle* h, const dgc_batch* batch);
int dgc_collect(dgc_handle* h, dgc_results* out);

/* ---- dazcon: overlaps in, consensus out (the trace-point realignment runs on the device) -------------------------
 * Replaces the per-overlap work of the reference's Reader thread, Target::getAlignments
 * (src/cpp/DazAlnProvider.cpp:304-369: Load_Subread, Complement_Seq, Compute_Trace_PTS(GREEDIEST) =
 * DALIGNER/align.c:4799-4896 + iter_np :4132-4459, decodeAlignment :383-417) and the target-sequence load of
 * DazAlnProvider::nextTarget(targSeq, dest) (:119-131).  Hit selection (TargetHit scoring, sortHits, the max-hit cap
 * and the provider's min-coverage filter) stays with the caller: a batch lists, per target, the overlaps that
 * getAlignments would turn into alignments, in that order. */
typedef struct dgc_daz_db {      /* the trimmed DAZZ_DB (Open_DB + Trim_DB): stays resident on the device */
    uint32_t n_reads;
    const uint64_t* read_boff;   /* [n_reads] HITS_READ.boff: byte offset of the read in `bases` */
    const uint32_t* read_len;    /* [n_reads] HITS_READ.rlen */
    const uint8_t* bases;        /* the .bps file: 2 bits per base, 4 bases per byte, first base in the top bits */
    uint64_t bases_bytes;
} dgc_daz_db;

typedef struct dgc_daz_batch {
    uint32_t n_targets;
    uint32_t n_ovls;
    uint32_t tspace;                /* trace spacing of the LAS file; must be <= 125 (byte-wide trace values) */
    const uint32_t* tgt_aread;      /* [n_targets] the A read (0-based DB index); FASTA id = aread + 1 */
    const uint32_t* tgt_ovl_begin;  /* [n_targets+1] CSR into the overlap arrays */
    const uint32_t* ovl_bread;      /* [n_ovls] */
    const uint32_t* ovl_comp;       /* [n_ovls] COMP(flags): B is reverse-complemented */
    const int32_t* ovl_abpos;       /* [n_ovls] Path.abpos / aepos / bbpos / bepos */
    const int32_t* ovl_aepos;
    const int32_t* ovl_bbpos;
    const int32_t* ovl_bepos;
    const uint64_t* ovl_trace_off;  /* [n_ovls] first trace value of the overlap in `trace` */
    const uint32_t* ovl_tlen;       /* [n_ovls] Path.tlen: number of trace values (pairs: diffs, B advance); even, >= 2 */
    const uint8_t* trace;
    uint64_t trace_bytes;
} dgc_daz_batch;

int dgc_daz_set_db(dgc_handle* h, const dgc_daz_db* db);
/* Same pipeline discipline as dgc_submit (results through dgc_collect, in submit order, dgc_results unchanged). */
int dgc_submit_daz(dgc_handle* h, const dgc_daz_batch* batch);
/* Test hook: the gapped strings the device writes for every overlap of a batch (what decodeAlignment produces):
 * q/t of overlap k at out_off[k] .. +out_len[k] in out_q / out_t (capacity out_cap bytes each). */
int dgc_debug_daz_strings(dgc_handle* h, const dgc_daz_batch* batch, uint8_t* out_q, uint8_t* out_t, uint64_t out_cap,
                          uint64_t* out_off, uint32_t* out_len);

/* ---- `pbdagcon -a`: ungapped 'pre' records in, consensus out (the alignment step runs on the device) ---------------
 * Replaces `std::for_each(alns.begin(), alns.end(), SimpleAligner())` (src/cpp/main.cpp:127-129) and
 * SimpleAligner::align (src/cpp/SimpleAligner.cpp:25-63) in front of the consensus path.  The reference's arithmetic
 * there is blasr_libcpp's (source absent): this aligner is the repository's own banded DP; the
 * coordinate and strand bookkeeping of SimpleAligner.cpp:49-62 is the reference's. */
typedef struct dgc_pre_batch {
    uint32_t n_targets;
    uint32_t n_recs;
    const uint32_t* tgt_tlen;       /* [n_targets] */
    const uint32_t* tgt_rec_begin;  /* [n_targets+1] CSR into the record arrays */
    const uint32_t* rec_start;      /* [n_recs] tstart of the 'pre' line (0-based, on the record's strand; Alignment.cpp:95-111) */
    const uint32_t* rec_strand;     /* [n_recs] 0 = '+', 1 = '-' */
    const uint32_t* rec_qlen;       /* [n_recs] ungapped query segment length */
    const uint32_t* rec_tseglen;    /* [n_recs] ungapped target segment length */
    const uint64_t* rec_q_off;      /* [n_recs] into arena (ASCII) */
    const uint64_t* rec_t_off;
    const uint8_t* arena;
    uint64_t arena_bytes;
} dgc_pre_batch;
int dgc_submit_pre(dgc_handle* h, const dgc_pre_batch* batch);
/* Test hook: the gapped strings and Alignment::start the device produced for every record (0 columns = nothing aligns). */
int dgc_debug_pre_strings(dgc_handle* h, const dgc_pre_batch* batch, uint8_t* out_q, uint8_t* out_t, uint64_t out_cap,
                          uint64_t* out_off, uint32_t* out_len, uint32_t* out_start);

/* ---- the `-a` aligner for callers that must not allocate --------------------------------------------------------
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same kernel and same
 * band schedule (64, 128, ... 8192) as dgc_submit_pre / dgc_debug_pre_strings, so the strings are the same bytes, but
 * nothing is allocated, on the device or through any library: every pointer of dgc_pre_dev is a DEVICE pointer the
 * caller owns, and the kernel's scratch (move matrices, record lists, slab offsets) is one file-scope `__device__`
 * array of DGC_NOALLOC_SCRATCH_BYTES (+ 64 KB of lists) inside the library, which every GPU that loads the library holds until exit.
 * Limits: n_recs <= DGC_NOALLOC_MAX_RECS, and every record's move matrix, (qlen + 1) * (2 * min(band, max(qlen,
 * tseglen) + 1) + 1) bytes at the band it needs, must fit the scratch on its own; beyond either the call returns
 * DGC_E_INVALID with the reason in `err` and launches nothing more (a batch larger than the scratch runs in waves).
 * The call works on `stream` (a cudaStream_t, NULL = the default stream) and returns after synchronising it: the
 * band schedule reads the records' states back between passes (into a buffer on its stack).
 * ALL CALLS SHARE THE ONE ARRAY (per GPU): calls that overlap on the same GPU -- two host threads, two streams --
 * race on it and corrupt each other's move matrices; serialise them.  The array keeps the last call's bytes; nothing
 * in it has to start at zero (every cell a traceback reads is written by the same launch), and out_state, which does,
 * is cleared by the call. */
#define DGC_NOALLOC_MAX_RECS 4096u
#define DGC_NOALLOC_SCRATCH_BYTES (64u << 20)
typedef struct dgc_pre_dev {
    uint32_t n_recs;
    uint32_t reserved;
    /* inputs, device memory */
    const uint32_t* rec_start;      /* [n_recs] as dgc_pre_batch.rec_start */
    const uint32_t* rec_strand;     /* [n_recs] 0 = '+', 1 = '-' */
    const uint32_t* rec_qlen;       /* [n_recs] */
    const uint32_t* rec_tseglen;    /* [n_recs] */
    const uint32_t* rec_tlen;       /* [n_recs] length of the record's whole target (dgc_pre_batch.tgt_tlen of its target) */
    const uint64_t* rec_q_off;      /* [n_recs] into arena (ASCII) */
    const uint64_t* rec_t_off;
    const uint8_t* arena;
    /* outputs, device memory */
    uint8_t* out_arena;             /* gapped strings; record k owns qlen + tseglen bytes at out_q_off[k] and at out_t_off[k] */
    const uint64_t* out_q_off;      /* [n_recs] chosen by the caller */
    const uint64_t* out_t_off;
    uint32_t* out_len;              /* [n_recs] gapped columns (0 = nothing aligns) */
    uint32_t* out_start;            /* [n_recs] Alignment::start (1-based, target orientation; SimpleAligner.cpp:49-62) */
    uint32_t* out_tbegin;           /* [n_recs] target bases cut in front of / kept up to, within the segment */
    uint32_t* out_tend;
    uint32_t* out_state;            /* [n_recs] 1 = aligned, 3 = nothing aligns */
} dgc_pre_dev;
int dgc_pre_align_noalloc(int device, const dgc_pre_dev* args, void* stream, char* err, uint64_t err_cap);

/* Synchronous convenience: submit + collect. */
int dgc_consensus(dgc_handle* h, const dgc_batch* batch, dgc_results* out);

/* HBM-resident batches (benchmarking the kernels without the PCIe copy). */
int dgc_upload(dgc_handle* h, const dgc_batch* batch, dgc_dbatch** out);
int dgc_run(dgc_handle* h, dgc_dbatch* db, dgc_results* out);
int dgc_dbatch_free(dgc_handle* h, dgc_dbatch* db);
int dgc_get_timings(dgc_handle* h, dgc_timings* out);

/* Test hooks (parity against the oracle at intermediate stages).
 * dgc_debug_normalize: normalizeGaps+trimAln (Alignment.cpp:131-242) of every alignment of a
 *   resident batch; out_cols receives (q | t<<8) per column at out_off[a] .. +out_len[a].
 * dgc_debug_graph: text dump of one target's graph after build (stage 0) or after mergeNodes
 *   (stage 1), one line per vertex: "id base weight coverage backbone deleted | dst:count:visited
 *   ... | src:count:visited ..." (vertex ids = the reference's creation order). */
int dgc_debug_normalize(dgc_handle* h, dgc_dbatch* db, uint16_t* out_cols, uint64_t out_cap,
                        uint64_t* out_off, uint32_t* out_len, uint32_t* out_start, uint32_t* out_flags);
int dgc_debug_graph(dgc_handle* h, dgc_dbatch* db, uint32_t target, int stage, char** text, uint64_t* n);
void dgc_debug_free(char* text);

#ifdef __cplusplus
}
#endif
#endif
