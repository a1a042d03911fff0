"""ctypes binding of tests/hostsim (TEST INFRASTRUCTURE): the product's host compiler + the per-thread kernel
bodies run as CPU loops. Same interface as oracle_util.OracleLexer so tests can swap them."""
import ctypes as C
import os
import subprocess

import numpy as np

from oracle_util import LEXEM_DT, _ptr
from struspattern_b200.capi import ValueStrings

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HS_DIR = os.path.join(ROOT, "tests", "hostsim")
HS_LIB = os.path.join(HS_DIR, "_build", "libsp_hostsim.so")


def build_hostsim(force=False):
    srcs = [os.path.join(HS_DIR, "lexer_hostsim.cpp")] + [
        os.path.join(ROOT, "struspattern_b200", "csrc", f)
        for f in ("regex.hpp", "automaton.hpp", "lexer_compile.hpp", "lexer_device.cuh", "lexer_fused.cuh", "lexer_tables.h")]
    if not force and os.path.exists(HS_LIB) and all(os.path.getmtime(HS_LIB) >= os.path.getmtime(s) for s in srcs):
        return
    os.makedirs(os.path.dirname(HS_LIB), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I", os.path.join(ROOT, "include"),
                    "-I", os.path.join(ROOT, "struspattern_b200", "csrc"), "-o", HS_LIB, srcs[0]], check=True)


class HostSimLexer:
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            build_hostsim()
            L = C.CDLL(HS_LIB)
            L.hsim_create.restype = C.c_void_p
            L.hsim_last_error.restype = C.c_char_p
            L.hsim_result_count.restype = C.c_uint64
            L.hsim_raw_event_count.restype = C.c_uint64
            L.hsim_info.restype = C.c_int64
            L.hsim_serialize.restype = C.c_int64
            cls._lib = L
        return cls._lib

    def __init__(self, chunk=64, warmup=16, mode=1):
        self.L = self.lib()
        self.h = C.c_void_p(self.L.hsim_create())
        self.L.hsim_set_mode(self.h, C.c_int(mode))  # 0: per document loop, 1: island + slot formulation, 2: fused segments, 3: chunk resolution over completed events
        self.chunk, self.warmup = chunk, warmup

    def __del__(self):
        if getattr(self, "h", None):
            self.L.hsim_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.hsim_last_error(self.h).decode(errors="replace"))

    def define_option(self, name, value=0.0):
        self._ck(self.L.hsim_define_option(self.h, name.encode(), C.c_double(value)))

    def define_lexem(self, id, expr, result_index=0, level=0, posbind=0):
        e = expr if isinstance(expr, bytes) else expr.encode()
        self._ck(self.L.hsim_define_lexem(self.h, C.c_uint32(id), e, C.c_size_t(len(e)), C.c_uint32(result_index),
                                          C.c_uint32(level), C.c_uint32(posbind)))

    def define_symbol(self, symbolid, lexemid, name):
        n = name if isinstance(name, bytes) else name.encode()
        self._ck(self.L.hsim_define_symbol(self.h, C.c_uint32(symbolid), C.c_uint32(lexemid), n, C.c_size_t(len(n))))

    def compile(self):
        self._ck(self.L.hsim_compile(self.h))

    def info(self, key):
        return int(self.L.hsim_info(self.h, key.encode()))

    def serialize(self):
        n = int(self.L.hsim_serialize(self.h, None, C.c_uint64(0)))
        if n < 0:
            self._ck(n)
        buf = C.create_string_buffer(n)
        assert int(self.L.hsim_serialize(self.h, buf, C.c_uint64(n))) == n
        return buf.raw

    def deserialize(self, image):
        self._ck(self.L.hsim_deserialize(self.h, image, C.c_uint64(len(image))))

    def scan(self, text, doc_offsets, nthreads=1, keep_raw=False):
        buf = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
        # 16 byte aligned copy with slack so that the vector path is exercised
        store = np.zeros(len(buf) + 32, dtype=np.uint8)
        shift = (-store.ctypes.data) % 16
        abuf = store[shift:shift + len(buf)]
        abuf[:] = buf
        offs = np.ascontiguousarray(doc_offsets, dtype=np.uint64)
        ndocs = len(offs) - 1
        self._ck(self.L.hsim_scan_batch(self.h, _ptr(abuf), _ptr(offs), C.c_uint32(ndocs), C.c_uint32(self.chunk),
                                        C.c_uint32(self.warmup)))
        n = int(self.L.hsim_result_count(self.h))
        out = {"lexems": np.zeros(n, dtype=LEXEM_DT), "offsets": np.zeros(ndocs + 1, dtype=np.uint64),
               "status": np.zeros(ndocs, dtype=np.uint32), "raw_event_count": int(self.L.hsim_raw_event_count(self.h))}
        self.L.hsim_fetch(self.h, _ptr(out["lexems"]) if n else None, _ptr(out["offsets"]), _ptr(out["status"]))
        if keep_raw:
            nr = out["raw_event_count"]
            out["raw"] = np.zeros((nr, 3), dtype=np.uint32)
            out["raw_offsets"] = np.zeros(ndocs + 1, dtype=np.uint64)
            self.L.hsim_fetch_raw(self.h, _ptr(out["raw"]) if nr else _ptr(np.zeros(3, dtype=np.uint32)), _ptr(out["raw_offsets"]))
        return out

    def match(self, src):
        b = src if isinstance(src, bytes) else src.encode()
        out = self.scan(b, [0, len(b)])
        return [tuple(int(x) for x in r) for r in out["lexems"]]


HS_MATCHER_LIB = os.path.join(HS_DIR, "_build", "libsp_hostsim_matcher.so")


def build_hostsim_matcher(force=False):
    srcs = [os.path.join(HS_DIR, "matcher_hostsim.cpp")] + [
        os.path.join(ROOT, "struspattern_b200", "csrc", f)
        for f in ("matcher_program.hpp", "matcher_device.cuh", "matcher_tables.h", "matcher_eval.cuh")]
    if not force and os.path.exists(HS_MATCHER_LIB) and all(os.path.getmtime(HS_MATCHER_LIB) >= os.path.getmtime(s) for s in srcs):
        return
    os.makedirs(os.path.dirname(HS_MATCHER_LIB), exist_ok=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-w", "-I", os.path.join(ROOT, "include"),
                    "-I", os.path.join(ROOT, "struspattern_b200", "csrc"), "-o", HS_MATCHER_LIB, srcs[0]], check=True)


class HostSimMatcher(ValueStrings):
    """The product's rule-graph builder + per-document state machine run on the CPU (tests/hostsim/matcher_hostsim.cpp).
    Same interface as oracle_util.RefMatcher."""
    _lib = None

    @classmethod
    def lib(cls):
        if cls._lib is None:
            build_hostsim_matcher()
            L = C.CDLL(HS_MATCHER_LIB)
            L.msim_create.restype = C.c_void_p
            for f in ("msim_last_error", "msim_pattern_name", "msim_variable_name"):
                getattr(L, f).restype = C.c_char_p
            for f in ("msim_result_count", "msim_item_count", "msim_retries"):
                getattr(L, f).restype = C.c_uint64
            L.msim_info.restype = C.c_int64
            L.msim_result_value.restype = C.c_int64
            L.msim_item_value.restype = C.c_int64
            cls._lib = L
        return cls._lib

    def __init__(self, caps=None, mode=0):
        self.L = self.lib()
        self.h = C.c_void_p(self.L.msim_create())
        self.L.msim_set_mode(self.h, C.c_int(mode))
        if caps:
            self.L.msim_set_caps(self.h, *[C.c_uint32(c) for c in caps])

    def __del__(self):
        if getattr(self, "h", None):
            self.L.msim_destroy(self.h)
            self.h = None

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.L.msim_last_error(self.h).decode(errors="replace"))

    def define_option(self, name, value=0.0):
        self._ck(self.L.msim_define_option(self.h, name.encode(), C.c_double(value)))

    def define_term_frequency(self, termid, df):
        self._ck(self.L.msim_define_term_frequency(self.h, C.c_uint32(termid), C.c_double(df)))

    def push_term(self, termid):
        self._ck(self.L.msim_push_term(self.h, C.c_uint32(termid)))

    def push_expression(self, op, argc, rng, card):
        self._ck(self.L.msim_push_expression(self.h, C.c_int(op), C.c_size_t(argc), C.c_uint32(rng), C.c_uint32(card)))

    def push_pattern(self, name):
        self._ck(self.L.msim_push_pattern(self.h, name.encode()))

    def attach_variable(self, name):
        self._ck(self.L.msim_attach_variable(self.h, name.encode()))

    def define_pattern(self, name, fmt="", visible=True):
        self._ck(self.L.msim_define_pattern(self.h, name.encode(), fmt.encode(), C.c_int(1 if visible else 0)))

    def compile(self):
        self._ck(self.L.msim_compile(self.h))

    def info(self, key):
        return int(self.L.msim_info(self.h, key.encode()))

    def retries(self):
        return int(self.L.msim_retries(self.h))

    def run(self, tokens, tok_offsets, origsegs=None, nthreads=1, with_results=True):
        from oracle_util import ITEM_DT, RESULT_DT, STATS_DT
        tokens = np.ascontiguousarray(tokens, dtype=LEXEM_DT)
        tok_offsets = np.ascontiguousarray(tok_offsets, dtype=np.uint64)
        ndocs = len(tok_offsets) - 1
        if origsegs is not None:
            origsegs = np.ascontiguousarray(origsegs, dtype=np.uint32)
        self._ck(self.L.msim_run_batch(self.h, _ptr(tokens), _ptr(origsegs), _ptr(tok_offsets), C.c_uint32(ndocs), C.c_int(1), C.c_int(1)))
        nres, nitems = int(self.L.msim_result_count(self.h)), int(self.L.msim_item_count(self.h))
        out = {"results": np.zeros(nres, dtype=RESULT_DT), "result_offsets": np.zeros(ndocs + 1, dtype=np.uint64),
               "items": np.zeros(nitems, dtype=ITEM_DT), "item_offsets": np.zeros(ndocs + 1, dtype=np.uint64),
               "stats": np.zeros(ndocs, dtype=STATS_DT), "status": np.zeros(ndocs, dtype=np.uint32)}
        self.L.msim_fetch(self.h, _ptr(out["results"]) if nres else None, _ptr(out["result_offsets"]),
                          _ptr(out["items"]) if nitems else None, _ptr(out["item_offsets"]), _ptr(out["stats"]), _ptr(out["status"]))
        return out

    def _result_value(self, rp, itemsp, buf, n):
        return self.L.msim_result_value(self.h, rp, itemsp, buf, n)

    def _item_value(self, ip, buf, n):
        return self.L.msim_item_value(self.h, ip, buf, n)

    def result_name(self, idx):
        return self.L.msim_pattern_name(self.h, C.c_uint32(int(idx))).decode()

    def item_name(self, idx):
        return self.L.msim_variable_name(self.h, C.c_uint32(int(idx))).decode()
