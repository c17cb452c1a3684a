"""Seeded synthetic alignment generator; test/bench infrastructure.
Thin ctypes wrapper over synth.c; returns packed batches as numpy arrays in the layout the
C-ABI takes (include/dagcon_b200.h: dgc_batch)."""
import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libdgc_synth.so")


class _Params(C.Structure):
    _fields_ = [("seed", C.c_uint64), ("L", C.c_uint32), ("cov", C.c_double), ("err", C.c_double)]


class _CBatch(C.Structure):
    _fields_ = [("n_targets", C.c_uint32), ("n_alns", C.c_uint32),
                ("tgt_tlen", C.POINTER(C.c_uint32)), ("tgt_aln_begin", C.POINTER(C.c_uint32)),
                ("aln_start", C.POINTER(C.c_uint32)), ("aln_len", C.POINTER(C.c_uint32)),
                ("aln_q_off", C.POINTER(C.c_uint64)), ("aln_t_off", C.POINTER(C.c_uint64)),
                ("arena", C.POINTER(C.c_char)), ("arena_bytes", C.c_uint64), ("arena_cap", C.c_uint64),
                ("backbones", C.POINTER(C.c_char))]


_lib = None


def _load():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            raise ImportError(f"{_SO} missing: run `python __graft_entry__.py` (build) first")
        _lib = C.CDLL(_SO)
        _lib.dgs_batch_new.restype = C.POINTER(_CBatch)
        _lib.dgs_batch_new.argtypes = [C.POINTER(_Params), C.c_uint64, C.c_uint32]
        _lib.dgs_batch_free.argtypes = [C.POINTER(_CBatch)]
        _lib.dgs_write_m5.restype = C.c_int64
        _lib.dgs_write_m5.argtypes = [C.POINTER(_Params), C.c_uint64, C.c_uint32, C.c_char_p]
        _lib.dgs_alns_per_target.argtypes = [C.POINTER(_Params)]
        _lib.dgs_backbone.argtypes = [C.POINTER(_Params), C.c_uint64, C.c_char_p]
        _lib.dgs_pack4.restype = C.c_int64
        _lib.dgs_pack4.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    return _lib


@dataclass
class Batch:
    """Packed targets: CSR over alignments + one byte arena (raw q/t ASCII, as the AlnProvider
    delivers them).  backbones is None for the pbdagcon (size_t ctor) flavour."""
    tgt_tlen: np.ndarray        # u32 [T]
    tgt_aln_begin: np.ndarray   # u32 [T+1]
    aln_start: np.ndarray       # u32 [A]  1-based target start
    aln_len: np.ndarray         # u32 [A]  raw columns
    aln_q_off: np.ndarray       # u64 [A]
    aln_t_off: np.ndarray       # u64 [A]
    arena: np.ndarray           # u8
    tgt_bb_off: np.ndarray = None   # u64 [T] offsets of backbone strings in arena, or None
    backbones: np.ndarray = None    # u8 [T, L] truth (tests only)

    @property
    def n_targets(self):
        return int(self.tgt_tlen.shape[0])

    @property
    def n_alns(self):
        return int(self.aln_len.shape[0])

    def target_alns(self, t):
        """[(q, t, start1)] python view of one target (for the oracle)."""
        out = []
        a = self.arena
        for k in range(int(self.tgt_aln_begin[t]), int(self.tgt_aln_begin[t + 1])):
            n = int(self.aln_len[k])
            qo, to = int(self.aln_q_off[k]), int(self.aln_t_off[k])
            out.append((a[qo:qo + n].tobytes(), a[to:to + n].tobytes(), int(self.aln_start[k])))
        return out

    def alg_bytes(self):
        """Algorithmic input bytes: sum(2*len_raw + 16) + 16 per target (+ backbone)."""
        b = int(2 * self.aln_len.astype(np.int64).sum() + 16 * self.n_alns + 16 * self.n_targets)
        if self.tgt_bb_off is not None:
            b += int(self.tgt_tlen.astype(np.int64).sum())
        return b


def _np(ptr, n, dtype):
    return np.ctypeslib.as_array(ptr, shape=(n,)).view(dtype).copy() if n else np.zeros(0, dtype)


def gen_batch(seed, L, cov, t0, nt, err=0.15) -> Batch:
    lib = _load()
    p = _Params(seed, L, cov, err)
    cb = lib.dgs_batch_new(C.byref(p), t0, nt)
    try:
        b = cb.contents
        T, A = b.n_targets, b.n_alns
        arena = np.frombuffer(C.string_at(b.arena, b.arena_bytes), dtype=np.uint8).copy()
        bbs = np.frombuffer(C.string_at(b.backbones, T * L), dtype=np.uint8).reshape(T, L).copy()
        return Batch(_np(b.tgt_tlen, T, np.uint32), _np(b.tgt_aln_begin, T + 1, np.uint32),
                     _np(b.aln_start, A, np.uint32), _np(b.aln_len, A, np.uint32),
                     _np(b.aln_q_off, A, np.uint64), _np(b.aln_t_off, A, np.uint64), arena, None, bbs)
    finally:
        lib.dgs_batch_free(cb)


def write_m5(seed, L, cov, t0, nt, path, err=0.15) -> int:
    lib = _load()
    p = _Params(seed, L, cov, err)
    n = lib.dgs_write_m5(C.byref(p), t0, nt, path.encode())
    if n < 0:
        raise OSError(f"cannot write {path}")
    return n


def backbone(seed, L, cov, t, err=0.15) -> bytes:
    """The backbone target t's reads are drawn against (what gen_batch returns in Batch.backbones)."""
    lib = _load()
    p = _Params(seed, L, cov, err)
    buf = C.create_string_buffer(L + 1)
    lib.dgs_backbone(C.byref(p), t, buf)
    return buf.raw[:L]


def pack4(b: Batch):
    """(arena, aln_q_off, aln_t_off) of the batch with its strings as 4-bit codes, two columns per byte
    (DGC_BATCH_PACKED4), or None when a string holds a character outside "-ACGTN."; batches with backbone strings are
    not handled here."""
    lib = _load()
    assert b.tgt_bb_off is None
    ln = np.ascontiguousarray(b.aln_len, np.uint32)
    cap = int((((ln.astype(np.int64) + 1) // 2 + 15) & ~15).sum()) * 2 + 16
    out = np.zeros(cap, np.uint8)
    qo, to = np.zeros(b.n_alns, np.uint64), np.zeros(b.n_alns, np.uint64)
    arena = np.ascontiguousarray(b.arena, np.uint8)
    n = lib.dgs_pack4(arena.ctypes.data, b.n_alns, ln.ctypes.data, np.ascontiguousarray(b.aln_q_off, np.uint64).ctypes.data,
                      np.ascontiguousarray(b.aln_t_off, np.uint64).ctypes.data, out.ctypes.data, qo.ctypes.data, to.ctypes.data)
    if n < 0:
        return None
    return out[:n], qo, to


def alns_per_target(cov):
    p = _Params(0, 1000, cov, 0.15)
    return _load().dgs_alns_per_target(C.byref(p))


def concat(batches):
    """Concatenate batches (used to assemble multi-shape fuzz inputs)."""
    tl, ab, st, ln, qo, to, ar = [], [np.zeros(1, np.uint32)], [], [], [], [], []
    a_base, byte_base = 0, 0
    for b in batches:
        tl.append(b.tgt_tlen)
        ab.append(b.tgt_aln_begin[1:] + np.uint32(a_base))
        st.append(b.aln_start); ln.append(b.aln_len)
        qo.append(b.aln_q_off + np.uint64(byte_base)); to.append(b.aln_t_off + np.uint64(byte_base))
        pad = (-len(b.arena)) % 16
        ar.append(b.arena)
        if pad:
            ar.append(np.zeros(pad, np.uint8))
        a_base += b.n_alns
        byte_base += len(b.arena) + pad
    return Batch(np.concatenate(tl), np.concatenate(ab).astype(np.uint32), np.concatenate(st), np.concatenate(ln),
                 np.concatenate(qo), np.concatenate(to), np.concatenate(ar))


def from_python(targets, with_backbone=False) -> Batch:
    """targets: [(tlen, backbone bytes | None, [(q, t, start1)])] -> Batch."""
    tl, ab, st, ln, qo, to, bbo = [], [0], [], [], [], [], []
    chunks, pos = [], 0

    def put(bs):
        nonlocal pos
        off = pos
        chunks.append(bs)
        pos += len(bs)
        pad = (-pos) % 16
        if pad:
            chunks.append(b"\0" * pad)
            pos += pad
        return off

    for tlen, bb, alns in targets:
        tl.append(tlen)
        if with_backbone:
            bbo.append(put(bb))
        for q, t, s in alns:
            assert len(q) == len(t)
            st.append(s); ln.append(len(q))
            off = put(q + t)
            qo.append(off); to.append(off + len(q))
        ab.append(len(st))
    arena = np.frombuffer(b"".join(chunks), dtype=np.uint8).copy() if chunks else np.zeros(0, np.uint8)
    return Batch(np.array(tl, np.uint32), np.array(ab, np.uint32), np.array(st, np.uint32), np.array(ln, np.uint32),
                 np.array(qo, np.uint64), np.array(to, np.uint64), arena,
                 np.array(bbo, np.uint64) if with_backbone else None)
