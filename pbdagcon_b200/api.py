"""ctypes mirror of include/dagcon_b200.h -- the host-side view tests and bench.py use.

`Consensus` plays the role of the reference's Consensus functor (reference src/cpp/main.cpp:97-152,
src/cpp/dazcon.cpp:61-107): it takes what an AlnProvider yields (targets with their alignments), and
returns the FASTA records `>{id}/{range0}_{range1}\\n{seq}\\n` in input-target order.  All compute
happens in libdagcon_b200.so on the GPU; there is no Python or CPU implementation behind it and
construction fails when the library or a device is missing."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "csrc", "libdagcon_b200.so")

T_OK, T_SKIPPED, T_BAD_ALIGNMENT, T_TOO_MANY_ALNS, T_INTERNAL = 0, 1, 2, 3, 4
F_NO_NORMALIZE, F_NO_TRIM, F_NO_MERGE, F_SINGLE = 1, 2, 4, 8
MAX_INFLIGHT = 4
BATCH_PACKED4 = 1

EXPORTS = ["dgc_abi_version", "dgc_device_count", "dgc_create", "dgc_destroy", "dgc_last_error", "dgc_host_alloc",
           "dgc_host_free", "dgc_submit", "dgc_collect", "dgc_consensus", "dgc_upload", "dgc_run", "dgc_dbatch_free",
           "dgc_get_timings", "dgc_debug_normalize", "dgc_debug_graph", "dgc_debug_free",
           "dgc_daz_set_db", "dgc_submit_daz", "dgc_debug_daz_strings", "dgc_submit_pre", "dgc_debug_pre_strings",
           "dgc_pre_align_noalloc"]


class DagconError(RuntimeError):
    pass


class Opts(C.Structure):
    _fields_ = [("min_cov", C.c_uint32), ("min_len", C.c_uint32), ("trim", C.c_uint32), ("flags", C.c_uint32),
                ("device", C.c_int32), ("reserved0", C.c_uint32), ("scratch_bytes", C.c_uint64)]


class CBatch(C.Structure):
    _fields_ = [("n_targets", C.c_uint32), ("n_alns", C.c_uint32),
                ("tgt_tlen", C.c_void_p), ("tgt_aln_begin", C.c_void_p), ("tgt_bb_off", C.c_void_p),
                ("aln_start", C.c_void_p), ("aln_len", C.c_void_p), ("aln_q_off", C.c_void_p), ("aln_t_off", C.c_void_p),
                ("arena", C.c_void_p), ("arena_bytes", C.c_uint64), ("flags", C.c_uint32), ("reserved", C.c_uint32)]


class CPreBatch(C.Structure):
    _fields_ = [("n_targets", C.c_uint32), ("n_recs", C.c_uint32),
                ("tgt_tlen", C.c_void_p), ("tgt_rec_begin", C.c_void_p), ("rec_start", C.c_void_p), ("rec_strand", C.c_void_p),
                ("rec_qlen", C.c_void_p), ("rec_tseglen", C.c_void_p), ("rec_q_off", C.c_void_p), ("rec_t_off", C.c_void_p),
                ("arena", C.c_void_p), ("arena_bytes", C.c_uint64)]


class CPreDev(C.Structure):
    """dgc_pre_dev: every pointer is a DEVICE pointer the caller owns (dgc_pre_align_noalloc allocates nothing)."""
    _fields_ = [("n_recs", C.c_uint32), ("reserved", C.c_uint32)] + \
               [(k, C.c_void_p) for k in ("rec_start", "rec_strand", "rec_qlen", "rec_tseglen", "rec_tlen", "rec_q_off", "rec_t_off", "arena",
                                          "out_arena", "out_q_off", "out_t_off", "out_len", "out_start", "out_tbegin", "out_tend", "out_state")]


NOALLOC_MAX_RECS = 4096
NOALLOC_SCRATCH_BYTES = 64 << 20


class PreBatch:
    """'pre' records (reference Alignment.cpp:82-112): targets = [(tlen, [(qseq, tseq, strand '+'|'-', tstart)])]."""

    def __init__(self, targets):
        tl, rb, st, sd, ql, tl2, qo, to, chunks, pos = [], [0], [], [], [], [], [], [], [], 0
        for tlen, recs in targets:
            tl.append(tlen)
            for q, t, strand, start in recs:
                st.append(start); sd.append(1 if strand in ("-", b"-") else 0)
                ql.append(len(q)); tl2.append(len(t))
                qo.append(pos); chunks.append(q); pos += len(q)
                to.append(pos); chunks.append(t); pos += len(t)
            rb.append(len(st))
        self.a = [np.array(x, dt) for x, dt in ((tl, np.uint32), (rb, np.uint32), (st, np.uint32), (sd, np.uint32), (ql, np.uint32),
                                                (tl2, np.uint32), (qo, np.uint64), (to, np.uint64))]
        self.arena = np.frombuffer(b"".join(chunks) + b"\0" * 16, np.uint8).copy()
        self.n_targets, self.n_recs = len(tl), len(st)
        self.cap = int(sum(ql) + sum(tl2))
        c = CPreBatch()
        c.n_targets, c.n_recs = self.n_targets, self.n_recs
        (c.tgt_tlen, c.tgt_rec_begin, c.rec_start, c.rec_strand, c.rec_qlen, c.rec_tseglen, c.rec_q_off, c.rec_t_off) = [_ptr(x) for x in self.a]
        c.arena, c.arena_bytes = _ptr(self.arena), int(self.arena.size) - 16
        self.c = c


class CResults(C.Structure):
    _fields_ = [("n_targets", C.c_uint32), ("n_frags", C.c_uint32),
                ("tgt_status", C.POINTER(C.c_uint32)), ("tgt_frag_begin", C.POINTER(C.c_uint32)),
                ("frag_range0", C.POINTER(C.c_int32)), ("frag_range1", C.POINTER(C.c_int32)),
                ("frag_seq_off", C.POINTER(C.c_uint64)), ("frag_seq_len", C.POINTER(C.c_uint32)),
                ("seq_arena", C.POINTER(C.c_uint8)), ("seq_bytes", C.c_uint64)]


class Timings(C.Structure):
    _fields_ = [(k, C.c_float) for k in ("ms_h2d", "ms_normalize", "ms_count", "ms_build", "ms_merge", "ms_bestpath",
                                         "ms_compact", "ms_d2h", "ms_total", "ms_span")] + \
               [("launches", C.c_uint32), ("chunks", C.c_uint32), ("arena_bytes", C.c_uint64), ("norm_cols", C.c_uint64),
                ("nodes", C.c_uint64), ("edges", C.c_uint64), ("ms_realign", C.c_float), ("reserved1", C.c_uint32)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


_lib = None


def load_library():
    """Load libdagcon_b200.so (built by __graft_entry__.build()).  Raises if it is missing: the
    product has no other implementation to fall back to."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise DagconError(f"{LIB_PATH} is missing: run `python __graft_entry__.py` to build the CUDA library")
    lib = C.CDLL(LIB_PATH)
    for name in EXPORTS:
        if not hasattr(lib, name):
            raise DagconError(f"{LIB_PATH} does not export {name}")
    H = C.c_void_p
    lib.dgc_abi_version.restype = C.c_int
    lib.dgc_device_count.restype = C.c_int
    lib.dgc_create.argtypes = [C.POINTER(Opts), C.POINTER(H)]
    lib.dgc_destroy.argtypes = [H]
    lib.dgc_last_error.restype = C.c_char_p
    lib.dgc_last_error.argtypes = [H]
    lib.dgc_host_alloc.argtypes = [C.POINTER(C.c_void_p), C.c_uint64]
    lib.dgc_host_free.argtypes = [C.c_void_p]
    lib.dgc_submit.argtypes = [H, C.POINTER(CBatch)]
    lib.dgc_collect.argtypes = [H, C.POINTER(CResults)]
    lib.dgc_consensus.argtypes = [H, C.POINTER(CBatch), C.POINTER(CResults)]
    lib.dgc_upload.argtypes = [H, C.POINTER(CBatch), C.POINTER(H)]
    lib.dgc_run.argtypes = [H, H, C.POINTER(CResults)]
    lib.dgc_dbatch_free.argtypes = [H, H]
    lib.dgc_get_timings.argtypes = [H, C.POINTER(Timings)]
    lib.dgc_debug_normalize.argtypes = [H, H, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dgc_debug_graph.argtypes = [H, H, C.c_uint32, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]
    lib.dgc_debug_free.argtypes = [C.c_void_p]
    lib.dgc_submit_pre.argtypes = [H, C.POINTER(CPreBatch)]
    lib.dgc_debug_pre_strings.argtypes = [H, C.POINTER(CPreBatch), C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    lib.dgc_pre_align_noalloc.argtypes = [C.c_int, C.POINTER(CPreDev), C.c_void_p, C.c_char_p, C.c_uint64]
    _lib = lib
    return lib


def version():
    return load_library().dgc_abi_version()


class PinnedArray:
    """numpy view over cudaMallocHost memory (dgc_host_alloc): the pinned staging buffers a
    dispatcher packs targets into."""

    def __init__(self, shape, dtype):
        lib = load_library()
        self.nbytes = int(np.prod(shape)) * np.dtype(dtype).itemsize
        p = C.c_void_p()
        if lib.dgc_host_alloc(C.byref(p), max(self.nbytes, 1)) != 0:
            raise DagconError("dgc_host_alloc failed")
        self.ptr = p.value
        buf = (C.c_uint8 * max(self.nbytes, 1)).from_address(self.ptr)
        self.array = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)

    def free(self):
        if self.ptr:
            self.array = None
            load_library().dgc_host_free(self.ptr)
            self.ptr = None


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class PackedBatch:
    """Holds the numpy arrays of a dgc_batch alive together with the C struct."""

    def __init__(self, tgt_tlen, tgt_aln_begin, aln_start, aln_len, aln_q_off, aln_t_off, arena, tgt_bb_off=None, ids=None, packed4=False):
        self.tgt_tlen = np.ascontiguousarray(tgt_tlen, np.uint32)
        self.tgt_aln_begin = np.ascontiguousarray(tgt_aln_begin, np.uint32)
        self.aln_start = np.ascontiguousarray(aln_start, np.uint32)
        self.aln_len = np.ascontiguousarray(aln_len, np.uint32)
        self.aln_q_off = np.ascontiguousarray(aln_q_off, np.uint64)
        self.aln_t_off = np.ascontiguousarray(aln_t_off, np.uint64)
        self.arena = np.ascontiguousarray(arena, np.uint8)
        self.tgt_bb_off = None if tgt_bb_off is None else np.ascontiguousarray(tgt_bb_off, np.uint64)
        self.ids = ids
        self.n_targets = int(self.tgt_tlen.shape[0])
        self.n_alns = int(self.aln_len.shape[0])
        assert self.tgt_aln_begin.shape[0] == self.n_targets + 1
        c = CBatch()
        c.n_targets, c.n_alns = self.n_targets, self.n_alns
        c.tgt_tlen, c.tgt_aln_begin = _ptr(self.tgt_tlen), _ptr(self.tgt_aln_begin)
        c.tgt_bb_off = _ptr(self.tgt_bb_off) if self.tgt_bb_off is not None else None
        c.aln_start, c.aln_len = _ptr(self.aln_start), _ptr(self.aln_len)
        c.aln_q_off, c.aln_t_off = _ptr(self.aln_q_off), _ptr(self.aln_t_off)
        c.arena, c.arena_bytes = _ptr(self.arena), int(self.arena.size)
        c.flags = BATCH_PACKED4 if packed4 else 0
        self.packed4 = packed4
        self.c = c

    @classmethod
    def from_synth(cls, b, ids=None):
        return cls(b.tgt_tlen, b.tgt_aln_begin, b.aln_start, b.aln_len, b.aln_q_off, b.aln_t_off, b.arena, b.tgt_bb_off, ids)

    @classmethod
    def from_synth_packed4(cls, b, arena_out=None):
        """The batch with its strings as 4-bit codes (DGC_BATCH_PACKED4): what a host packer sends to halve the bytes on
        the link.  None when a string cannot be packed.  arena_out: optional preallocated (pinned) uint8 array factory
        taking a size and returning an array to hold the packed arena."""
        from . import synth
        r = synth.pack4(b)
        if r is None:
            return None
        arena, qo, to = r
        if arena_out is not None:
            dst = arena_out(arena.size)
            dst[:] = arena
            arena = dst
        return cls(b.tgt_tlen, b.tgt_aln_begin, b.aln_start, b.aln_len, qo, to, arena, None, None, packed4=True)

    def alg_bytes(self):
        """Algorithmic input bytes: sum(2*len_raw + 16) per alignment + 16 per target (+ backbone)."""
        b = int(2 * self.aln_len.astype(np.int64).sum() + 16 * self.n_alns + 16 * self.n_targets)
        if self.tgt_bb_off is not None:
            b += int(self.tgt_tlen.astype(np.int64).sum())
        return b


class Results:
    """Copy of a dgc_results (the handle's memory is only valid until the next submit, collect or run)."""

    def __init__(self, r: CResults):
        T, F = r.n_targets, r.n_frags
        self.status = np.ctypeslib.as_array(r.tgt_status, (T,)).copy() if T else np.zeros(0, np.uint32)
        self.frag_begin = np.ctypeslib.as_array(r.tgt_frag_begin, (T + 1,)).copy()
        if F:
            self.range0 = np.ctypeslib.as_array(r.frag_range0, (F,)).copy()
            self.range1 = np.ctypeslib.as_array(r.frag_range1, (F,)).copy()
            self.seq_off = np.ctypeslib.as_array(r.frag_seq_off, (F,)).copy()
            self.seq_len = np.ctypeslib.as_array(r.frag_seq_len, (F,)).copy()
        else:
            self.range0 = self.range1 = np.zeros(0, np.int32)
            self.seq_off = np.zeros(0, np.uint64)
            self.seq_len = np.zeros(0, np.uint32)
        self.seq = np.ctypeslib.as_array(r.seq_arena, (int(r.seq_bytes),)).copy() if r.seq_bytes else np.zeros(0, np.uint8)

    @property
    def n_targets(self):
        return int(self.status.shape[0])

    def frags(self, t):
        """[(range0, range1, seq bytes)] of target t -- the reference's std::vector<CnsResult>."""
        out = []
        for f in range(int(self.frag_begin[t]), int(self.frag_begin[t + 1])):
            o, n = int(self.seq_off[f]), int(self.seq_len[f])
            out.append((int(self.range0[f]), int(self.range1[f]), self.seq[o:o + n].tobytes()))
        return out

    def total_bases(self):
        return int(self.seq_len.astype(np.int64).sum())

    def fasta(self, ids, dazcon=False):
        """FASTA text in input-target order: main.cpp:139-145 / dazcon.cpp:92-99 header formats."""
        out = []
        well = 0
        for t in range(self.n_targets):
            for r0, r1, seq in self.frags(t):
                tid = ids[t].decode() if isinstance(ids[t], bytes) else ids[t]
                if dazcon:
                    out.append(f">{tid}/{well}/{r0}_{r1}\n{seq.decode()}\n")
                    well += 1
                else:
                    out.append(f">{tid}/{r0}_{r1}\n{seq.decode()}\n")
        return "".join(out)


class Consensus:
    """One GPU's consensus engine.  Mirrors ProgramOpts (-c min_cov, -m/-l min_len, -t trim)."""

    def __init__(self, min_cov=6, min_len=500, trim=50, flags=0, device=0, scratch_bytes=0):
        self.lib = load_library()
        o = Opts(min_cov, min_len, trim, flags, device, 0, scratch_bytes)
        h = C.c_void_p()
        rc = self.lib.dgc_create(C.byref(o), C.byref(h))
        if rc != 0:
            raise DagconError(f"dgc_create failed ({rc}): {self.lib.dgc_last_error(None).decode()}")
        self.h = h
        self._inflight = []

    def _check(self, rc, what):
        if rc != 0:
            raise DagconError(f"{what} failed ({rc}): {self.lib.dgc_last_error(self.h).decode()}")

    def close(self):
        if getattr(self, "h", None):
            self.lib.dgc_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # --- the reference-facing call: host buffers in, CnsResults out
    def __call__(self, batch: PackedBatch) -> Results:
        r = CResults()
        self._check(self.lib.dgc_consensus(self.h, C.byref(batch.c), C.byref(r)), "dgc_consensus")
        return Results(r)

    def submit(self, batch: PackedBatch):
        self._check(self.lib.dgc_submit(self.h, C.byref(batch.c)), "dgc_submit")
        self._inflight.append(batch)

    def collect(self, copy=True):
        r = CResults()
        self._check(self.lib.dgc_collect(self.h, C.byref(r)), "dgc_collect")
        self._inflight.pop(0)
        return Results(r) if copy else r

    # --- HBM-resident batches
    def upload(self, batch: PackedBatch):
        db = C.c_void_p()
        self._check(self.lib.dgc_upload(self.h, C.byref(batch.c), C.byref(db)), "dgc_upload")
        return db

    def run(self, db, copy=True):
        r = CResults()
        self._check(self.lib.dgc_run(self.h, db, C.byref(r)), "dgc_run")
        return Results(r) if copy else r

    def free(self, db):
        self.lib.dgc_dbatch_free(self.h, db)

    def timings(self):
        t = Timings()
        self.lib.dgc_get_timings(self.h, C.byref(t))
        return t.as_dict()

    # --- `-a`: ungapped 'pre' records, aligned on the device
    def submit_pre(self, batch: "PreBatch"):
        self._check(self.lib.dgc_submit_pre(self.h, C.byref(batch.c)), "dgc_submit_pre")
        self._inflight.append(batch)

    def debug_pre_strings(self, batch: "PreBatch"):
        """[(q, t, start1)] per record as the device aligned it ((b"", b"", start) when nothing aligns)."""
        R = batch.n_recs
        q, t = np.zeros(batch.cap + 16, np.uint8), np.zeros(batch.cap + 16, np.uint8)
        off, ln, st = np.zeros(R + 1, np.uint64), np.zeros(R + 1, np.uint32), np.zeros(R + 1, np.uint32)
        self._check(self.lib.dgc_debug_pre_strings(self.h, C.byref(batch.c), q.ctypes.data, t.ctypes.data, batch.cap + 16, off.ctypes.data,
                                                   ln.ctypes.data, st.ctypes.data), "dgc_debug_pre_strings")
        return [(q[int(off[k]):int(off[k]) + int(ln[k])].tobytes(), t[int(off[k]):int(off[k]) + int(ln[k])].tobytes(), int(st[k])) for k in range(R)]

    # --- test hooks
    def debug_normalize(self, batch: PackedBatch):
        """[(q, t, start)] after normalizeGaps+trimAln for every alignment (None for minLen-skipped ones)."""
        db = self.upload(batch)
        try:
            A = batch.n_alns
            cap = int(((2 * batch.aln_len.astype(np.int64) + 7) & ~7).sum()) + 8
            cols = np.zeros(cap, np.uint16)
            off = np.zeros(A + 1, np.uint64)
            ln = np.zeros(A + 1, np.uint32)
            st = np.zeros(A + 1, np.uint32)
            fl = np.zeros(A + 1, np.uint32)
            self._check(self.lib.dgc_debug_normalize(self.h, db, cols.ctypes.data, cap, off.ctypes.data, ln.ctypes.data,
                                                     st.ctypes.data, fl.ctypes.data), "dgc_debug_normalize")
            out = []
            for k in range(A):
                if fl[k] & 1:
                    out.append(None)
                    continue
                c = cols[int(off[k]):int(off[k]) + int(ln[k])]
                out.append(((c & 0xFF).astype(np.uint8).tobytes(), (c >> 8).astype(np.uint8).tobytes(), int(st[k])))
            return out
        finally:
            self.free(db)

    def debug_graph(self, batch: PackedBatch, target, stage):
        db = self.upload(batch)
        try:
            p, n = C.c_void_p(), C.c_uint64()
            self._check(self.lib.dgc_debug_graph(self.h, db, target, stage, C.byref(p), C.byref(n)), "dgc_debug_graph")
            s = C.string_at(p.value, n.value).decode()
            self.lib.dgc_debug_free(p)
            return s
        finally:
            self.free(db)
