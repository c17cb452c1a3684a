"""ctypes access to the dazcon front's test hooks (pbdagcon_b200/host/daz_capi.cpp) and, where it was built, the
reference's own DazAlnProvider (oracle/_ref/libdaz_ref.so).  Test infrastructure only."""
import ctypes as C
import hashlib
import json
import os

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HOST_SO = os.path.join(ROOT, "pbdagcon_b200", "host", "libdagcon_host.so")
DAZREF_SO = os.path.join(ROOT, "oracle", "_ref", "libdaz_ref.so")
DAZCON = os.path.join(ROOT, "pbdagcon_b200", "host", "dazcon")
DAZCON_REF = os.path.join(ROOT, "oracle", "_ref", "dazcon_ref")
FIX = os.path.join(ROOT, "tests", "golden", "daz")
_fixdir = None


def fixture_paths():
    """(db, las) of the committed fixture, laid out the way DAZZ_DB wants it: small.db next to its hidden companions
    .small.idx / .small.bps (committed under visible names, copied into a scratch directory once per process)."""
    global _fixdir
    if _fixdir is None:
        import atexit
        import shutil
        import tempfile
        d = tempfile.mkdtemp(prefix="dgc_dazfix_")
        atexit.register(shutil.rmtree, d, True)
        for f, g in (("small.db", "small.db"), ("small.idx", ".small.idx"), ("small.bps", ".small.bps"), ("small.las", "small.las")):
            shutil.copy(os.path.join(FIX, f), os.path.join(d, g))
        _fixdir = d
    return os.path.join(_fixdir, "small.db"), os.path.join(_fixdir, "small.las")


class _Lazy:
    def __init__(self, i):
        self.i = i

    def __fspath__(self):
        return fixture_paths()[self.i]

    def __str__(self):
        return fixture_paths()[self.i]

    def encode(self):
        return fixture_paths()[self.i].encode()


DB, LAS = _Lazy(0), _Lazy(1)


def gold():
    return json.load(open(os.path.join(ROOT, "tests", "golden", "golden_dazcon.json")))


class _Daz:
    def __init__(self, path, prefix):
        self.lib = C.CDLL(path)
        self.p = prefix
        f = getattr(self.lib, prefix + "dump")
        f.restype = C.c_void_p
        f.argtypes = [C.c_char_p, C.c_char_p, C.c_uint, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]
        t = getattr(self.lib, prefix + "text")
        t.restype = C.c_void_p
        t.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        getattr(self.lib, prefix + "free").argtypes = [C.c_void_p]
        s = getattr(self.lib, prefix + "ovl_score")
        s.restype = C.c_float
        s.argtypes = [C.POINTER(C.c_int), C.c_int]

    def dump(self, db, las, min_cov=6, max_hits=85, sort_cov=0, proper=0, targets=()):
        arr = (C.c_int * len(targets))(*targets)
        h = getattr(self.lib, self.p + "dump")(db.encode(), las.encode(), min_cov, max_hits, sort_cov, proper, arr, len(targets))
        n = C.c_uint64()
        s = C.string_at(getattr(self.lib, self.p + "text")(h, C.byref(n)), n.value)
        getattr(self.lib, self.p + "free")(h)
        return s

    def ovl_score(self, paths):
        flat = [x for p in paths for x in p]
        return float(getattr(self.lib, self.p + "ovl_score")((C.c_int * len(flat))(*flat), len(paths)))

    def ovl_scores(self, chains):
        return [self.ovl_score(c) for c in chains]


def host():
    d = _Daz(HOST_SO, "dgh_daz_")
    f = d.lib.dgh_daz_dump_device
    f.restype = C.c_void_p
    f.argtypes = [C.c_char_p, C.c_char_p, C.c_uint, C.c_uint, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]

    def dump_device(db, las, min_cov=6, max_hits=85, sort_cov=0, proper=0, targets=()):
        arr = (C.c_int * len(targets))(*targets)
        h = f(str(db).encode(), str(las).encode(), min_cov, max_hits, sort_cov, proper, arr, len(targets))
        n = C.c_uint64()
        s = C.string_at(d.lib.dgh_daz_text(h, C.byref(n)), n.value)
        d.lib.dgh_daz_free(h)
        return s
    d.dump_device = dump_device
    d.lib.dgh_daz_group_hits.restype = C.c_int
    d.lib.dgh_daz_group_hits.argtypes = [C.POINTER(C.c_int), C.c_int]
    return d


def have_dazref():
    return os.path.exists(DAZREF_SO)


class _RefProc:
    """The reference provider in a CHILD process.  oracle/_ref/libdaz_ref.so and the product's libdagcon_host.so each carry
    a statically linked libstdc++; loaded into one process, their GNU-unique locale symbols are merged and the
    reference's std::stringstream use crashes.  A process of its own keeps the checker untouched by the product."""

    def _call(self, what, payload):
        import pickle
        import subprocess
        import sys
        code = ("import pickle,sys; sys.path.insert(0,%r); import daz_helpers as d; r=d._Daz(d.DAZREF_SO,'dazref_'); "
                "w,p=pickle.load(sys.stdin.buffer); sys.stdout.buffer.write(pickle.dumps(getattr(r,w)(*p[0],**p[1])))" % os.path.dirname(__file__))
        out = subprocess.run([sys.executable, "-c", code], input=pickle.dumps((what, payload)), stdout=subprocess.PIPE, check=True).stdout
        return pickle.loads(out)

    def dump(self, *a, **kw):
        return self._call("dump", (a, kw))

    def ovl_score(self, *a, **kw):
        return self._call("ovl_score", (a, kw))

    def ovl_scores(self, chains):
        return self._call("ovl_scores", ((chains,), {}))


def ref():
    return _RefProc()


def parse_dump(text):
    """[(id, tlen, targSeq, [(sid, start, q, t)])] from a provider dump."""
    out = []
    for ln in text.split(b"\n"):
        if ln.startswith(b"T "):
            _, tid, tlen, n, seq = ln.split(b" ")
            out.append((tid.decode(), int(tlen), seq, []))
        elif ln.startswith(b"A "):
            _, sid, start, q, t = ln.split(b" ")
            out[-1][3].append((sid.decode(), int(start), q, t))
    return out


def md5(b):
    return hashlib.md5(b).hexdigest()
