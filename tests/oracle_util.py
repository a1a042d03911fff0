"""ctypes bindings of the ORACLE libraries (oracle/_ref/*.so). Test infrastructure only:
nothing under struspattern_b200/ may import this module."""
import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
REF_DIR = os.path.join(ORACLE_DIR, "_ref")

LEXEM_DT = np.dtype([("id", "<u4"), ("ordpos", "<u4"), ("origpos", "<u4"), ("origsize", "<u4")])
RESULT_DT = np.dtype([(n, "<u4") for n in ("name", "format", "ordpos", "ordend", "origseg", "origpos",
                                            "endseg", "endpos", "item_begin", "item_count")])
ITEM_DT = np.dtype([(n, "<u4") for n in ("variable", "format", "ordpos", "ordend", "origseg", "origpos",
                                          "endseg", "endpos")])
STATS_DT = np.dtype([(n, "<u8") for n in ("nof_programs_installed", "nof_altkey_programs_installed",
                                           "nof_signals_fired", "nof_open_patterns", "nof_events")])


def build_oracle():
    """Build the oracle libraries when a compiler is around (no-op if they are up to date)."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all"], check=True, stdout=subprocess.DEVNULL)


def _ptr(a, t=C.c_void_p):
    return a.ctypes.data_as(t) if a is not None else None


class RefMatcher:
    """The reference's own PatternMatcher (libsp_ref_matcher.so), driven like its tests do."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(REF_DIR, "libsp_ref_matcher.so")
            if not os.path.exists(path):
                build_oracle()
            L = C.CDLL(path)
            L.refm_create.restype = C.c_void_p
            L.refm_last_error.restype = C.c_char_p
            L.refm_name.restype = C.c_char_p
            L.refm_result_count.restype = C.c_uint64
            L.refm_item_count.restype = C.c_uint64
            L.refm_result_value.restype = C.c_int64
            L.refm_item_value.restype = C.c_int64
            for f in ("refm_destroy", "refm_last_error", "refm_define_option", "refm_define_term_frequency",
                      "refm_push_term", "refm_push_expression", "refm_push_pattern", "refm_attach_variable",
                      "refm_define_pattern", "refm_compile", "refm_run_batch", "refm_result_count",
                      "refm_item_count", "refm_fetch", "refm_name", "refm_result_value", "refm_item_value"):
                getattr(L, f).argtypes = None
            cls._lib = L
        return cls._lib

    def __init__(self):
        self.L = self.lib()
        self.h = C.c_void_p(self.L.refm_create())

    def __del__(self):
        if getattr(self, "h", None):
            self.L.refm_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.refm_last_error(self.h).decode())

    def define_option(self, name, value=0.0):
        self._ck(self.L.refm_define_option(self.h, name.encode(), C.c_double(value)))

    def define_term_frequency(self, termid, df):
        self._ck(self.L.refm_define_term_frequency(self.h, C.c_uint32(termid), C.c_double(df)))

    def push_term(self, termid):
        self._ck(self.L.refm_push_term(self.h, C.c_uint32(termid)))

    def push_expression(self, op, argc, rng, card):
        self._ck(self.L.refm_push_expression(self.h, C.c_int(op), C.c_size_t(argc), C.c_uint32(rng), C.c_uint32(card)))

    def push_pattern(self, name):
        self._ck(self.L.refm_push_pattern(self.h, name.encode()))

    def attach_variable(self, name):
        self._ck(self.L.refm_attach_variable(self.h, name.encode()))

    def define_pattern(self, name, fmt="", visible=True):
        self._ck(self.L.refm_define_pattern(self.h, name.encode(), fmt.encode(), C.c_int(1 if visible else 0)))

    def compile(self):
        self._ck(self.L.refm_compile(self.h))

    def _value(self, fn, idx):
        n = int(fn(self.h, C.c_uint64(idx), None, C.c_uint64(0)))
        if n == -100:       # SP_NO_VALUE: the reference delivered NULL
            return None
        assert n >= 0
        buf = C.create_string_buffer(n + 1)
        fn(self.h, C.c_uint64(idx), buf, C.c_uint64(n))
        return buf.raw[:n]

    def run(self, tokens, tok_offsets, origsegs=None, nthreads=1, with_results=True, values=False):
        """values=True adds "result_values" / "item_values": the value strings of PatternMatcherResult /
        PatternMatcherResultItem as bytes (None where the reference delivers NULL)."""
        tokens = np.ascontiguousarray(tokens, dtype=LEXEM_DT)
        tok_offsets = np.ascontiguousarray(tok_offsets, dtype=np.uint64)
        ndocs = len(tok_offsets) - 1
        if origsegs is not None:
            origsegs = np.ascontiguousarray(origsegs, dtype=np.uint32)
        self._ck(self.L.refm_run_batch(self.h, _ptr(tokens), _ptr(origsegs), _ptr(tok_offsets), C.c_uint32(ndocs),
                                       C.c_int(nthreads), C.c_int(1 if with_results else 0)))
        nres = int(self.L.refm_result_count(self.h))
        nitems = int(self.L.refm_item_count(self.h))
        out = {
            "results": np.zeros(nres if with_results else 0, dtype=RESULT_DT),
            "result_offsets": np.zeros(ndocs + 1, dtype=np.uint64),
            "items": np.zeros(nitems, dtype=ITEM_DT),
            "item_offsets": np.zeros(ndocs + 1, dtype=np.uint64),
            "stats": np.zeros(ndocs, dtype=STATS_DT),
            "status": np.zeros(ndocs, dtype=np.uint32),
        }
        self.L.refm_fetch(self.h, _ptr(out["results"]) if with_results and nres else None, _ptr(out["result_offsets"]),
                          _ptr(out["items"]) if nitems else None, _ptr(out["item_offsets"]), _ptr(out["stats"]),
                          _ptr(out["status"]))
        if values and with_results:
            out["result_values"] = [self._value(self.L.refm_result_value, i) for i in range(nres)]
            out["item_values"] = [self._value(self.L.refm_item_value, i) for i in range(nitems)]
        return out

    def name(self, idx):
        return self.L.refm_name(self.h, C.c_uint32(int(idx))).decode()

    result_name = name
    item_name = name


def canonical_results(out, result_name, item_name, sort=True):
    """Per document: sorted list of (name, ordpos, ordend, origseg, origpos, endseg, endpos, (items...))
    with items in emission order -- the multiset the parity contract compares."""
    docs = []
    ro, io = out["result_offsets"], out["item_offsets"]
    for d in range(len(ro) - 1):
        recs = []
        items = out["items"][int(io[d]):int(io[d + 1])]
        for r in out["results"][int(ro[d]):int(ro[d + 1])]:
            its = tuple((item_name(i["variable"]), int(i["ordpos"]), int(i["ordend"]), int(i["origseg"]), int(i["origpos"]),
                         int(i["endseg"]), int(i["endpos"]))
                        for i in items[int(r["item_begin"]):int(r["item_begin"]) + int(r["item_count"])])
            recs.append((result_name(r["name"]), int(r["ordpos"]), int(r["ordend"]), int(r["origseg"]), int(r["origpos"]),
                         int(r["endseg"]), int(r["endpos"]), its))
        if sort:
            recs.sort()
        docs.append(recs)
    return docs


class OracleLexer:
    """CPU restatement of level 1 (oracle/lexer_oracle.cpp -> libsp_oracle_lexer.so)."""

    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            path = os.path.join(REF_DIR, "libsp_oracle_lexer.so")
            if not os.path.exists(path):
                build_oracle()
            L = C.CDLL(path)
            L.lxo_create.restype = C.c_void_p
            L.lxo_last_error.restype = C.c_char_p
            L.lxo_result_count.restype = C.c_uint64
            L.lxo_raw_event_count.restype = C.c_uint64
            cls._lib = L
        return cls._lib

    def __init__(self):
        self.L = self.lib()
        self.h = C.c_void_p(self.L.lxo_create())

    def __del__(self):
        if getattr(self, "h", None):
            self.L.lxo_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.lxo_last_error(self.h).decode(errors="replace"))

    def define_option(self, name, value=0.0):
        self._ck(self.L.lxo_define_option(self.h, name.encode(), C.c_double(value)))

    def define_lexem(self, id, expr, result_index=0, level=0, posbind=0):
        e = expr if isinstance(expr, bytes) else expr.encode()
        self._ck(self.L.lxo_define_lexem(self.h, C.c_uint32(id), e, C.c_size_t(len(e)), C.c_uint32(result_index),
                                         C.c_uint32(level), C.c_uint32(posbind)))

    def define_symbol(self, symbolid, lexemid, name):
        n = name if isinstance(name, bytes) else name.encode()
        self._ck(self.L.lxo_define_symbol(self.h, C.c_uint32(symbolid), C.c_uint32(lexemid), n, C.c_size_t(len(n))))

    def compile(self):
        self._ck(self.L.lxo_compile(self.h))

    def scan(self, text, doc_offsets, nthreads=1, keep_raw=False):
        """text: bytes/np.uint8 array; doc_offsets: ndocs+1 byte offsets."""
        buf = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        offs = np.ascontiguousarray(doc_offsets, dtype=np.uint64)
        ndocs = len(offs) - 1
        self._ck(self.L.lxo_scan_batch(self.h, _ptr(buf), _ptr(offs), C.c_uint32(ndocs), C.c_int(nthreads),
                                       C.c_int(1 if keep_raw else 0)))
        n = int(self.L.lxo_result_count(self.h))
        out = {"lexems": np.zeros(n, dtype=LEXEM_DT), "offsets": np.zeros(ndocs + 1, dtype=np.uint64),
               "status": np.zeros(ndocs, dtype=np.uint32), "raw_event_count": int(self.L.lxo_raw_event_count(self.h))}
        self.L.lxo_fetch(self.h, _ptr(out["lexems"]) if n else None, _ptr(out["offsets"]), _ptr(out["status"]))
        if keep_raw:
            nr = out["raw_event_count"]
            out["raw"] = np.zeros((nr, 3), dtype=np.uint32)
            out["raw_offsets"] = np.zeros(ndocs + 1, dtype=np.uint64)
            self.L.lxo_fetch_raw(self.h, _ptr(out["raw"]) if nr else None, _ptr(out["raw_offsets"]))
        return out

    def match(self, src):
        """PatternLexerContext::match of one document -> list of (id, ordpos, origpos, origsize)."""
        b = src if isinstance(src, bytes) else src.encode()
        out = self.scan(b, [0, len(b)])
        return [tuple(int(x) for x in r) for r in out["lexems"]]
