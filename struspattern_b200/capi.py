"""ctypes binding of libstrus_pattern_b200.so (include/strus_pattern_b200.h) for the test and benchmark
harness. The classes mirror the reference's interface triple (instance -> context) for each level:

  Lexer        = PatternLexerInstanceInterface   (src/patternLexer.cpp:961-1151)
  LexerContext = PatternLexerContextInterface    (src/patternLexer.cpp:681-959), batched
  Matcher / MatcherContext likewise              (src/patternMatcher.cpp:345-704, 107-341)

There is no fallback: if the CUDA library is missing or no device is usable, calls raise."""
import ctypes as C
import os

import numpy as np

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libstrus_pattern_b200.so")

LEXEM_DT = np.dtype([("id", "<u4"), ("ordpos", "<u4"), ("origpos", "<u4"), ("origsize", "<u4")])
RESULT_PACKED_DT = np.dtype([(n, "<u4") for n in ("name_format", "ordpos", "origpos", "lens")])
ITEM_PACKED_DT = np.dtype([(n, "<u4") for n in ("var_lens", "ordpos", "origpos")])
RESULT_DT = np.dtype([(n, "<u4") for n in ("name", "format", "ordpos", "ordend", "origseg", "origpos",
                                            "endseg", "endpos", "item_begin", "item_count")])
ITEM_DT = np.dtype([(n, "<u4") for n in ("variable", "format", "ordpos", "ordend", "origseg", "origpos",
                                          "endseg", "endpos")])
STATS_DT = np.dtype([(n, "<u8") for n in ("nof_programs_installed", "nof_altkey_programs_installed",
                                           "nof_signals_fired", "nof_open_patterns", "nof_events")])

NO_VALUE = -100     # SP_NO_VALUE

BIND_CONTENT, BIND_SUCCESSOR, BIND_PREDECESSOR, BIND_UNIQUE = 0, 1, 2, 3
OP_SEQUENCE, OP_SEQUENCE_IMM, OP_SEQUENCE_STRUCT, OP_WITHIN, OP_WITHIN_STRUCT, OP_ANY, OP_AND = range(7)

_lib = None


class PatternError(RuntimeError):
    def __init__(self, code, msg):
        RuntimeError.__init__(self, "%s (code %d)" % (msg, code))
        self.code = code


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError("CUDA library %s is not built (run __graft_entry__.build()); there is no CPU fallback" % LIB_PATH)
        L = C.CDLL(LIB_PATH)
        vp, cp, u32, u64p = C.c_void_p, C.c_char_p, C.c_uint32, C.c_void_p
        sigs = {
            "sp_lexer_create": (vp, []), "sp_lexer_destroy": (None, [vp]), "sp_lexer_last_error": (cp, [vp]),
            "sp_lexer_define_option": (C.c_int, [vp, cp, C.c_double]),
            "sp_lexer_define_lexem": (C.c_int, [vp, u32, cp, C.c_size_t, u32, u32, u32]),
            "sp_lexer_define_symbol": (C.c_int, [vp, u32, u32, cp, C.c_size_t]),
            "sp_lexer_get_symbol": (u32, [vp, u32, cp, C.c_size_t]),
            "sp_lexer_compile": (C.c_int, [vp]), "sp_lexer_table_info": (C.c_int64, [vp, cp]),
            "sp_lexer_serialize": (C.c_int64, [vp, vp, C.c_uint64]), "sp_lexer_deserialize": (C.c_int, [vp, vp, C.c_uint64]),
            "sp_lexer_ctx_create": (vp, [vp, vp]), "sp_lexer_ctx_destroy": (None, [vp]),
            "sp_lexer_ctx_last_error": (cp, [vp]),
            "sp_lexer_scan": (C.c_int, [vp, vp, C.c_int, u64p, u32]),
            "sp_lexer_result_count": (C.c_int64, [vp]),
            "sp_lexer_result_lexems_device": (vp, [vp]), "sp_lexer_result_offsets_device": (vp, [vp]),
            "sp_lexer_result_status_device": (vp, [vp]),
            "sp_lexer_fetch": (C.c_int, [vp, vp, vp, vp]), "sp_lexer_ctx_counter": (C.c_int64, [vp, cp]),
            "sp_lexer_fetch_range": (C.c_int, [vp, u32, u32, vp, C.c_uint64, vp, vp]),
            "sp_version": (cp, []), "sp_device_count": (C.c_int, []),
        }
        msigs = {
            "sp_matcher_create": (vp, []), "sp_matcher_destroy": (None, [vp]), "sp_matcher_last_error": (cp, [vp]),
            "sp_matcher_define_option": (C.c_int, [vp, cp, C.c_double]),
            "sp_matcher_define_term_frequency": (C.c_int, [vp, u32, C.c_double]),
            "sp_matcher_push_term": (C.c_int, [vp, u32]),
            "sp_matcher_push_expression": (C.c_int, [vp, C.c_int, C.c_size_t, u32, u32]),
            "sp_matcher_push_pattern": (C.c_int, [vp, cp]), "sp_matcher_attach_variable": (C.c_int, [vp, cp]),
            "sp_matcher_define_pattern": (C.c_int, [vp, cp, cp, C.c_int]),
            "sp_matcher_compile": (C.c_int, [vp]), "sp_matcher_upload": (C.c_int, [vp]),
            "sp_matcher_pattern_name": (cp, [vp, u32]), "sp_matcher_variable_name": (cp, [vp, u32]),
            "sp_matcher_table_info": (C.c_int64, [vp, cp]), "sp_matcher_format_string": (cp, [vp, u32]),
            "sp_matcher_ctx_counter": (C.c_int64, [vp, cp]),
            "sp_matcher_ctx_create": (vp, [vp, vp]), "sp_matcher_ctx_destroy": (None, [vp]),
            "sp_matcher_ctx_last_error": (cp, [vp]),
            "sp_matcher_run": (C.c_int, [vp, vp, vp, vp, u32, C.c_int]),
            "sp_matcher_result_count": (C.c_int64, [vp]), "sp_matcher_item_count": (C.c_int64, [vp]),
            "sp_matcher_results_device": (vp, [vp]), "sp_matcher_items_device": (vp, [vp]),
            "sp_matcher_fetch": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
            "sp_matcher_run_lexer_output": (C.c_int, [vp, vp]),
            "sp_matcher_fetch_range": (C.c_int, [vp, u32, u32, vp, C.c_uint64, vp, vp, C.c_uint64, vp, vp, vp]),
            "sp_matcher_result_value": (C.c_int64, [vp, vp, vp, vp, C.c_uint64]),
            "sp_matcher_item_value": (C.c_int64, [vp, vp, vp, C.c_uint64]),
            "sp_resolve_value": (C.c_int64, [vp, C.c_uint64, vp, C.c_uint64, vp, C.c_uint64]),
            "sp_matcher_fetch_packed": (C.c_int, [vp, vp, vp, vp, vp, vp, vp]),
            "sp_unpack_results": (None, [vp, vp, vp, vp, u32, vp, vp]),
            "sp_matcher_count_noalloc": (C.c_int, [vp, vp, vp, u32, C.c_uint64, vp, vp, vp, vp, vp]),
        }
        for table in (sigs, msigs):
            for name, (res, args) in table.items():
                f = getattr(L, name)
                f.restype, f.argtypes = res, args
        _lib = L
    return _lib


def resolve_value(value, source):
    """sp_resolve_value: the span references of a value string replaced by the bytes of `source` (segment 0)."""
    if value is None:
        return None
    src = np.ascontiguousarray(np.frombuffer(bytes(source), dtype=np.uint8)) if not isinstance(source, np.ndarray) else np.ascontiguousarray(source, dtype=np.uint8)
    L = lib()
    n = int(L.sp_resolve_value(value, len(value), _ptr(src), len(src), None, 0))
    if n < 0:
        raise PatternError(n, "corrupt value string")
    buf = C.create_string_buffer(n + 1)
    L.sp_resolve_value(value, len(value), _ptr(src), len(src), buf, n)
    return buf.raw[:n]


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Lexer:
    def __init__(self):
        self.L = lib()
        self.h = self.L.sp_lexer_create()

    def close(self):
        if getattr(self, "h", None):
            self.L.sp_lexer_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise PatternError(rc, self.L.sp_lexer_last_error(self.h).decode(errors="replace"))

    def define_option(self, name, value=0.0):
        self._ck(self.L.sp_lexer_define_option(self.h, name.encode(), value))

    def define_lexem(self, id, expr, result_index=0, level=0, posbind=BIND_CONTENT):
        e = expr if isinstance(expr, bytes) else expr.encode()
        self._ck(self.L.sp_lexer_define_lexem(self.h, id, e, len(e), result_index, level, posbind))

    def define_symbol(self, symbolid, lexemid, name):
        n = name if isinstance(name, bytes) else name.encode()
        self._ck(self.L.sp_lexer_define_symbol(self.h, symbolid, lexemid, n, len(n)))

    def get_symbol(self, lexemid, name):
        n = name if isinstance(name, bytes) else name.encode()
        return int(self.L.sp_lexer_get_symbol(self.h, lexemid, n, len(n)))

    def compile(self):
        self._ck(self.L.sp_lexer_compile(self.h))

    def table_info(self, key):
        return int(self.L.sp_lexer_table_info(self.h, key.encode()))

    def serialize(self):
        """-> bytes: image of the compiled tables (sp_lexer_serialize)"""
        n = int(self.L.sp_lexer_serialize(self.h, None, 0))
        if n < 0:
            self._ck(n)
        buf = C.create_string_buffer(n)
        m = int(self.L.sp_lexer_serialize(self.h, buf, n))
        if m != n:
            self._ck(m if m < 0 else -1)
        return buf.raw

    def deserialize(self, image):
        """loads an image made by serialize() into this (empty) lexer instead of define_* + compile"""
        self._ck(self.L.sp_lexer_deserialize(self.h, image, len(image)))

    def create_context(self, stream=None):
        return LexerContext(self, stream)

    # convenience with the interface of the oracle classes (tests swap them)
    def scan(self, text, doc_offsets, **_):
        ctx = self.create_context()
        try:
            return ctx.scan(text, doc_offsets)
        finally:
            ctx.close()

    def match(self, src):
        b = src if isinstance(src, bytes) else src.encode()
        out = self.scan(b, [0, len(b)])
        return [tuple(int(x) for x in r) for r in out["lexems"]]


class LexerContext:
    def __init__(self, lexer, stream=None):
        self.L = lexer.L
        self.lexer = lexer  # keep the instance alive (src/patternLexer.cpp:685: raw pointer into the instance)
        self.h = self.L.sp_lexer_ctx_create(lexer.h, stream)
        if not self.h:
            raise PatternError(-4, self.L.sp_lexer_last_error(lexer.h).decode(errors="replace"))
        self.ndocs = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.sp_lexer_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc < 0:
            raise PatternError(rc, self.L.sp_lexer_ctx_last_error(self.h).decode(errors="replace"))
        return rc

    def launch(self, text, doc_offsets, on_device=False):
        """text: host uint8 array (or int device pointer when on_device); doc_offsets: host uint64[ndocs+1]."""
        offs = np.ascontiguousarray(doc_offsets, dtype=np.uint64)
        self.ndocs = len(offs) - 1
        if on_device:
            tp = C.c_void_p(int(text))
        else:
            self._text = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray)) else np.ascontiguousarray(text, dtype=np.uint8)
            tp = _ptr(self._text)
        self._ck(self.L.sp_lexer_scan(self.h, tp, 1 if on_device else 0, _ptr(offs), self.ndocs))

    def count(self):
        return self._ck(self.L.sp_lexer_result_count(self.h))

    def counter(self, key):
        return int(self.L.sp_lexer_ctx_counter(self.h, key.encode()))

    def fetch(self):
        n = self.count()
        out = {"lexems": np.zeros(n, dtype=LEXEM_DT), "offsets": np.zeros(self.ndocs + 1, dtype=np.uint64),
               "status": np.zeros(self.ndocs, dtype=np.uint32)}
        self._ck(self.L.sp_lexer_fetch(self.h, _ptr(out["lexems"]) if n else None, _ptr(out["offsets"]),
                                       _ptr(out["status"]) if self.ndocs else None))
        out["raw_event_count"] = self.counter("raw_events")
        return out

    def fetch_docs(self, first, ndocs):
        """Lexems of the documents [first, first+ndocs) of the last scan; offsets rebased to the slice."""
        offs = np.zeros(ndocs + 1, dtype=np.uint64)
        self._ck(self.L.sp_lexer_fetch_range(self.h, first, ndocs, None, C.c_uint64(0), _ptr(offs), None))
        n = int(offs[-1] - offs[0])
        out = {"lexems": np.zeros(n, dtype=LEXEM_DT), "offsets": offs, "status": np.zeros(ndocs, dtype=np.uint32)}
        self._ck(self.L.sp_lexer_fetch_range(self.h, first, ndocs, _ptr(out["lexems"]) if n else None, C.c_uint64(n), _ptr(offs), _ptr(out["status"]) if ndocs else None))
        out["offsets"] = offs - offs[0]
        return out

    def fetch_into(self, lexems, offsets, status):
        """Copy the results into caller provided (e.g. pinned) arrays; returns the number of lexems."""
        n = self.count()
        if n > len(lexems):
            raise PatternError(-1, "result buffer too small: %d lexems" % n)
        self._ck(self.L.sp_lexer_fetch(self.h, _ptr(lexems), _ptr(offsets), _ptr(status)))
        return n

    def scan(self, text, doc_offsets, on_device=False):
        self.launch(text, doc_offsets, on_device)
        return self.fetch()


class ValueStrings:
    """visible_results for classes that provide _result_value( result_ptr, doc_items_ptr, buf, n) and
    _item_value( item_ptr, buf, n) with the return convention of sp_matcher_result_value."""

    def _value(self, call):
        n = int(call(None, 0))
        if n == NO_VALUE:
            return None
        if n < 0:
            raise PatternError(n, "value string: corrupt item list or unknown format handle")
        buf = C.create_string_buffer(n + 1)
        call(buf, n)
        return buf.raw[:n]

    def visible_results(self, out):
        """The output of a run (fetch / run) as a caller of the reference's fetchResults sees it (src/patternMatcher.cpp:164-190,
        248-269): per document the same arrays without the nested records (the items of a result with a format string and of
        an item that refers to a formatted sub expression are the material of the value string, not items), plus
        "result_values" / "item_values": bytes, or None where the reference delivers NULL. Span references inside the values
        are resolved with resolve_value."""
        item_size = ITEM_DT.itemsize
        results, items = out["results"].copy(), out["items"]
        base = items.ctypes.data
        keep, rvals, ivals = [], [], []
        io = out["item_offsets"]
        ro = out["result_offsets"]
        new_io = [0]
        fmt = items["format"]
        for d in range(len(ro) - 1):
            ib = int(io[d])
            for ri in range(int(ro[d]), int(ro[d + 1])):
                r = results[ri]
                first = len(keep) - new_io[-1]
                if r["format"]:
                    rp = C.c_void_p(results.ctypes.data + ri * RESULT_DT.itemsize)
                    rvals.append(self._value(lambda b, n: self._result_value(rp, C.c_void_p(base + ib * item_size), b, C.c_uint64(n))))
                else:
                    rvals.append(None)
                    k, e = ib + int(r["item_begin"]), ib + int(r["item_begin"]) + int(r["item_count"])
                    while k < e:
                        f = int(fmt[k])
                        keep.append(k)
                        if f & 0xFFFF:
                            ip = C.c_void_p(base + k * item_size)
                            ivals.append(self._value(lambda b, n: self._item_value(ip, b, C.c_uint64(n))))
                        else:
                            ivals.append(None)
                        k += 1 + (f >> 16)
                results[ri]["item_begin"] = first
                results[ri]["item_count"] = len(keep) - new_io[-1] - first
            new_io.append(len(keep))
        vis = dict(out)
        vis["results"] = results
        vis["items"] = items[np.array(keep, dtype=np.int64)] if keep else items[:0]
        vis["items"]["format"] &= 0xFFFF
        vis["item_offsets"] = np.array(new_io, dtype=np.uint64)
        vis["result_values"], vis["item_values"] = rvals, ivals
        return vis



class Matcher(ValueStrings):
    """PatternMatcherInstanceInterface over the C ABI; same method names as oracle_util.RefMatcher."""

    def __init__(self):
        self.L = lib()
        self.h = self.L.sp_matcher_create()

    def close(self):
        if getattr(self, "h", None):
            self.L.sp_matcher_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc != 0:
            raise PatternError(rc, self.L.sp_matcher_last_error(self.h).decode(errors="replace"))

    def define_option(self, name, value=0.0):
        self._ck(self.L.sp_matcher_define_option(self.h, name.encode(), value))

    def define_term_frequency(self, termid, df):
        self._ck(self.L.sp_matcher_define_term_frequency(self.h, termid, df))

    def push_term(self, termid):
        self._ck(self.L.sp_matcher_push_term(self.h, termid))

    def push_expression(self, op, argc, rng, card):
        self._ck(self.L.sp_matcher_push_expression(self.h, op, argc, rng, card))

    def push_pattern(self, name):
        self._ck(self.L.sp_matcher_push_pattern(self.h, name.encode()))

    def attach_variable(self, name):
        self._ck(self.L.sp_matcher_attach_variable(self.h, name.encode()))

    def define_pattern(self, name, fmt="", visible=True):
        self._ck(self.L.sp_matcher_define_pattern(self.h, name.encode(), fmt.encode(), 1 if visible else 0))

    def compile(self):
        self._ck(self.L.sp_matcher_compile(self.h))

    def table_info(self, key):
        return int(self.L.sp_matcher_table_info(self.h, key.encode()))

    def result_name(self, handle):
        s = self.L.sp_matcher_pattern_name(self.h, int(handle))
        return s.decode() if s else ""

    def item_name(self, handle):
        s = self.L.sp_matcher_variable_name(self.h, int(handle))
        return s.decode() if s else ""

    def create_context(self, stream=None):
        return MatcherContext(self, stream)

    def _result_value(self, rp, itemsp, buf, n):
        return self.L.sp_matcher_result_value(self.h, rp, itemsp, buf, n)

    def _item_value(self, ip, buf, n):
        return self.L.sp_matcher_item_value(self.h, ip, buf, n)

    def upload(self):
        self._ck(self.L.sp_matcher_upload(self.h))

    def count_noalloc(self, tokens_dev, tok_offsets_dev, ndocs, ntokens, doc_results_dev, doc_items_dev, doc_stats_dev, doc_flags_dev, stream=None):
        """sp_matcher_count_noalloc: all arguments but the sizes are int device pointers of the caller."""
        self._ck(self.L.sp_matcher_count_noalloc(self.h, C.c_void_p(int(tokens_dev)), C.c_void_p(int(tok_offsets_dev)), int(ndocs), C.c_uint64(int(ntokens)),
                                                 C.c_void_p(int(doc_results_dev)), C.c_void_p(int(doc_items_dev)), C.c_void_p(int(doc_stats_dev)),
                                                 C.c_void_p(int(doc_flags_dev)), stream))

    def run(self, tokens, tok_offsets, origsegs=None, **_):
        ctx = self.create_context()
        try:
            return ctx.run(tokens, tok_offsets, origsegs)
        finally:
            ctx.close()


class MatcherContext:
    def __init__(self, matcher, stream=None):
        self.L = matcher.L
        self.matcher = matcher
        self.h = self.L.sp_matcher_ctx_create(matcher.h, stream)
        if not self.h:
            raise PatternError(-4, self.L.sp_matcher_last_error(matcher.h).decode(errors="replace"))
        self.ndocs = 0

    def close(self):
        if getattr(self, "h", None):
            self.L.sp_matcher_ctx_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc < 0:
            raise PatternError(rc, self.L.sp_matcher_ctx_last_error(self.h).decode(errors="replace"))
        return rc

    def counter(self, key):
        return int(self.L.sp_matcher_ctx_counter(self.h, key.encode()))

    def launch(self, tokens, tok_offsets, origsegs=None, on_device=False, ndocs=None):
        """Host arrays (LEXEM_DT tokens, uint64 offsets, optional uint32 segments) or, with on_device, int device pointers."""
        if on_device:
            self.ndocs = int(ndocs)
            self._ck(self.L.sp_matcher_run(self.h, C.c_void_p(int(tokens)), C.c_void_p(int(origsegs)) if origsegs else None,
                                           C.c_void_p(int(tok_offsets)), self.ndocs, 1))
            return
        self._tok = np.ascontiguousarray(tokens, dtype=LEXEM_DT)
        self._off = np.ascontiguousarray(tok_offsets, dtype=np.uint64)
        self._seg = np.ascontiguousarray(origsegs, dtype=np.uint32) if origsegs is not None else None
        self.ndocs = len(self._off) - 1
        self._ck(self.L.sp_matcher_run(self.h, _ptr(self._tok), _ptr(self._seg), _ptr(self._off), self.ndocs, 0))

    def launch_lexer_output(self, lexer_ctx):
        self.ndocs = lexer_ctx.ndocs
        self._ck(self.L.sp_matcher_run_lexer_output(self.h, lexer_ctx.h))

    def count(self):
        return self._ck(self.L.sp_matcher_result_count(self.h)), self._ck(self.L.sp_matcher_item_count(self.h))

    def fetch(self):
        nres, nitems = self.count()
        nd = self.ndocs
        out = {"results": np.zeros(nres, dtype=RESULT_DT), "result_offsets": np.zeros(nd + 1, dtype=np.uint64),
               "items": np.zeros(nitems, dtype=ITEM_DT), "item_offsets": np.zeros(nd + 1, dtype=np.uint64),
               "stats": np.zeros(nd, dtype=STATS_DT), "status": np.zeros(nd, dtype=np.uint32)}
        self._ck(self.L.sp_matcher_fetch(self.h, _ptr(out["results"]) if nres else None, _ptr(out["result_offsets"]),
                                         _ptr(out["items"]) if nitems else None, _ptr(out["item_offsets"]),
                                         _ptr(out["stats"]) if nd else None, _ptr(out["status"]) if nd else None))
        return out

    def fetch_docs(self, first, ndocs):
        """Results of the documents [first, first+ndocs) of the last run; offsets rebased to the slice."""
        ro, io = np.zeros(ndocs + 1, dtype=np.uint64), np.zeros(ndocs + 1, dtype=np.uint64)
        self._ck(self.L.sp_matcher_fetch_range(self.h, first, ndocs, None, C.c_uint64(0), _ptr(ro), None, C.c_uint64(0), _ptr(io), None, None))
        nres, nitems = int(ro[-1] - ro[0]), int(io[-1] - io[0])
        out = {"results": np.zeros(nres, dtype=RESULT_DT), "items": np.zeros(nitems, dtype=ITEM_DT),
               "stats": np.zeros(ndocs, dtype=STATS_DT), "status": np.zeros(ndocs, dtype=np.uint32)}
        self._ck(self.L.sp_matcher_fetch_range(self.h, first, ndocs, _ptr(out["results"]) if nres else None, C.c_uint64(nres), _ptr(ro),
                                               _ptr(out["items"]) if nitems else None, C.c_uint64(nitems), _ptr(io),
                                               _ptr(out["stats"]) if ndocs else None, _ptr(out["status"]) if ndocs else None))
        out["result_offsets"], out["item_offsets"] = ro - ro[0], io - io[0]
        return out

    def fetch_into(self, results, result_offsets, items, item_offsets, stats, status):
        nres, nitems = self.count()
        if nres > len(results) or nitems > len(items):
            raise PatternError(-1, "result buffers too small: %d results, %d items" % (nres, nitems))
        self._ck(self.L.sp_matcher_fetch(self.h, _ptr(results), _ptr(result_offsets), _ptr(items), _ptr(item_offsets), _ptr(stats), _ptr(status)))
        return nres, nitems

    def fetch_packed_into(self, results, result_offsets, items, item_offsets, stats, status):
        """sp_matcher_fetch_packed into caller provided (e.g. pinned) arrays of RESULT_PACKED_DT / ITEM_PACKED_DT; raises
        PatternError (code -6) if the run is not packable."""
        nres, nitems = self.count()
        if nres > len(results) or nitems > len(items):
            raise PatternError(-1, "result buffers too small: %d results, %d items" % (nres, nitems))
        self._ck(self.L.sp_matcher_fetch_packed(self.h, _ptr(results), _ptr(result_offsets), _ptr(items), _ptr(item_offsets), _ptr(stats), _ptr(status)))
        return nres, nitems

    def run(self, tokens, tok_offsets, origsegs=None):
        self.launch(tokens, tok_offsets, origsegs)
        return self.fetch()


def unpack_results(results, result_offsets, items, item_offsets, out_results, out_items):
    """sp_unpack_results (host code): packed records of the documents described by the offsets -> sp_result / sp_item records."""
    lib().sp_unpack_results(_ptr(results), _ptr(result_offsets), _ptr(items), _ptr(item_offsets), len(result_offsets) - 1, _ptr(out_results), _ptr(out_items))
