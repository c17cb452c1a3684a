"""ctypes access to the checkers: the C oracle (oracle/libdagcon_oracle.so), the compiled
reference (oracle/_ref/libdagcon_ref.so, when built) and the synthetic generator.
Test infrastructure only."""
import ctypes as C
import os
import random

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SO = os.path.join(ROOT, "oracle", "libdagcon_oracle.so")
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libdagcon_ref.so")
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "pbdagcon_ref")

MODE_NORM, MODE_TRIM, MODE_MERGE, MODE_SINGLE, MODE_DUMP_PRE, MODE_DUMP_POST = 1, 2, 4, 8, 16, 32
MODE_PIPELINE = MODE_NORM | MODE_TRIM | MODE_MERGE


class _Checker:
    """Common wrapper over the orc_* / ref_* C entry points (same signatures by design)."""

    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.p = prefix
        f = getattr(self.lib, prefix + "_consensus")
        f.restype = C.c_void_p
        f.argtypes = [C.c_uint32, C.c_char_p, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p),
                      C.POINTER(C.c_int64), C.POINTER(C.c_uint32), C.c_int, C.c_int, C.c_int, C.c_int64]
        for name, res, args in [
            ("result_count", C.c_int, [C.c_void_p]),
            ("result_range", C.c_int, [C.c_void_p, C.c_int, C.c_int]),
            ("result_seq", C.c_int64, [C.c_void_p, C.c_int, C.POINTER(C.c_char_p)]),
            ("result_single", C.c_int64, [C.c_void_p, C.POINTER(C.c_char_p)]),
            ("result_dump", C.c_int64, [C.c_void_p, C.POINTER(C.c_char_p)]),
            ("result_free", None, [C.c_void_p]),
        ]:
            fn = getattr(self.lib, f"{prefix}_{name}")
            fn.restype = res
            fn.argtypes = args

    def _call(self, name, *a):
        return getattr(self.lib, f"{self.p}_{name}")(*a)

    def consensus(self, tlen, alns, backbone=None, mode=MODE_PIPELINE, trim=50, min_weight=0, min_len=0):
        """alns: list of (q: bytes, t: bytes, start1: int).  Returns dict(frags=[(r0, r1, seq)], single, dump, status)."""
        n = len(alns)
        qa = (C.c_char_p * n)(*[a[0] for a in alns])
        ta = (C.c_char_p * n)(*[a[1] for a in alns])
        la = (C.c_int64 * n)(*[len(a[0]) for a in alns])
        sa = (C.c_uint32 * n)(*[a[2] for a in alns])
        h = self._call("consensus", tlen, backbone, n, qa, ta, la, sa, mode, trim, min_weight, min_len)
        try:
            out = {"frags": [], "status": 0}
            if self.p == "orc":
                self.lib.orc_result_status.argtypes = [C.c_void_p]
                out["status"] = self.lib.orc_result_status(h)
            p = C.c_char_p()
            for i in range(self._call("result_count", h)):
                ln = self._call("result_seq", h, i, C.byref(p))
                seq = C.string_at(p, ln)
                out["frags"].append((self._call("result_range", h, i, 0), self._call("result_range", h, i, 1), seq))
            ln = self._call("result_single", h, C.byref(p))
            out["single"] = C.string_at(p, ln) if ln else b""
            ln = self._call("result_dump", h, C.byref(p))
            out["dump"] = C.string_at(p, ln).decode() if ln else ""
            return out
        finally:
            self._call("result_free", h)


class Oracle(_Checker):
    def __init__(self):
        super().__init__(ORACLE_SO, "orc")
        L = self.lib
        L.orc_normalize_gaps.restype = C.c_int64
        L.orc_normalize_gaps.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.c_char_p, C.c_char_p]
        L.orc_trim.restype = C.c_int64
        L.orc_trim.argtypes = [C.c_char_p, C.c_int64, C.c_int, C.POINTER(C.c_uint32), C.POINTER(C.c_int64)]

    def normalize(self, q, t):
        n = len(q)
        qo, to = C.create_string_buffer(2 * n + 1), C.create_string_buffer(2 * n + 1)
        m = self.lib.orc_normalize_gaps(q, t, n, qo, to)
        return qo.raw[:m], to.raw[:m]

    def trim(self, q, t, start, trim_len):
        st, lo = C.c_uint32(start), C.c_int64(0)
        m = self.lib.orc_trim(t, len(t), trim_len, C.byref(st), C.byref(lo))
        return q[lo.value:lo.value + m], t[lo.value:lo.value + m], st.value

    def normalize_trim(self, q, t, start, trim_len):
        qn, tn = self.normalize(q, t)
        return self.trim(qn, tn, start, trim_len)


class Ref(_Checker):
    """The unmodified reference sources, compiled against oracle/shim (only where /root/reference
    was available at build time; the built .so travels to the GPU box)."""

    def __init__(self):
        super().__init__(REF_SO, "ref")
        L = self.lib
        L.ref_normalize_trim.restype = C.c_int64
        L.ref_normalize_trim.argtypes = [C.c_char_p, C.c_char_p, C.c_int64, C.POINTER(C.c_uint32), C.c_int, C.c_int,
                                         C.c_char_p, C.c_char_p]
        L.ref_parse_line.restype = C.c_void_p
        L.ref_parse_line.argtypes = [C.c_char_p, C.c_int]
        L.ref_aln_str.restype = C.c_void_p
        L.ref_aln_str.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        L.ref_aln_u32.restype = C.c_uint32
        L.ref_aln_u32.argtypes = [C.c_void_p, C.c_int]
        L.ref_aln_free.argtypes = [C.c_void_p]

    def normalize_trim(self, q, t, start, trim_len, do_trim=True):
        n = len(q)
        qo, to = C.create_string_buffer(2 * n + 1), C.create_string_buffer(2 * n + 1)
        st = C.c_uint32(start)
        m = self.lib.ref_normalize_trim(q, t, n, C.byref(st), int(do_trim), trim_len, qo, to)
        return qo.raw[:m], to.raw[:m], st.value

    def parse_line(self, line, which):
        h = self.lib.ref_parse_line(line, which)
        try:
            n = C.c_int64()
            out = {}
            for i, k in enumerate(["id", "sid", "qstr", "tstr"]):
                p = self.lib.ref_aln_str(h, i, C.byref(n))
                out[k] = C.string_at(p, n.value)
            for i, k in enumerate(["tlen", "start", "end", "strand"]):
                out[k] = self.lib.ref_aln_u32(h, i)
            return out
        finally:
            self.lib.ref_aln_free(h)


def have_ref():
    return os.path.exists(REF_SO)


def checker():
    """The strongest checker available: the compiled reference (oracle/_ref travels to the GPU box) when it was
    built, else the C restatement.  Same consensus()/normalize_trim() interface either way."""
    return Ref() if have_ref() else Oracle()


# ------------------------------------------------------------------ python-side fuzz generators
def rand_target(rng: random.Random, L=None, n_aln=None, err=0.15, full=False, alphabet="ACGT", dots=False,
                min_span=20):
    """A small random target: (tlen, backbone bytes, [(q, t, start1)])."""
    L = L or rng.randint(30, 400)
    n_aln = n_aln if n_aln is not None else rng.randint(1, 12)
    bb = "".join(rng.choice(alphabet) for _ in range(L))
    alns = []
    for _ in range(n_aln):
        if full:
            s, e = 0, L
        else:
            span = rng.randint(min(min_span, L), L)
            s = rng.randint(0, L - span)
            e = s + span
        q, t = [], []
        pos = s
        while pos < e:
            u = rng.random()
            tb = bb[pos]
            if u < err * 0.5:
                q.append(rng.choice(alphabet)); t.append("." if dots and rng.random() < 0.2 else "-")
            elif u < err * 0.8:
                q.append("." if dots and rng.random() < 0.2 else "-"); t.append(tb); pos += 1
            elif u < err:
                q.append(rng.choice([c for c in alphabet if c != tb] or ["N"])); t.append(tb); pos += 1
            else:
                q.append(tb); t.append(tb); pos += 1
        alns.append(("".join(q).encode(), "".join(t).encode(), s + 1))
    return L, bb.encode(), alns
