"""BENCH/TEST INFRASTRUCTURE: python front of bench_util/corpus_gen.c (synthetic corpora of SURVEY.md 8d)."""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libcorpus.so")
SRC = os.path.join(HERE, "corpus_gen.c")
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB) or os.path.getmtime(LIB) < os.path.getmtime(SRC):
            os.makedirs(os.path.dirname(LIB), exist_ok=True)
            subprocess.run(["gcc", "-O2", "-fPIC", "-shared", "-pthread", "-o", LIB, SRC, "-lm"], check=True)
        L = C.CDLL(LIB)
        L.corpus_layout.restype = C.c_uint32
        L.corpus_layout.argtypes = [C.c_uint64, C.c_uint64, C.c_uint32, C.c_uint32, C.c_void_p, C.c_uint32]
        L.corpus_fill.argtypes = [C.c_uint64, C.c_char_p, C.c_void_p, C.c_uint32, C.c_double, C.c_void_p, C.c_void_p,
                                  C.c_uint32, C.c_uint32]
        L.corpus_vocab_word.restype = C.c_int
        L.corpus_vocab_word.argtypes = [C.c_uint64, C.c_uint32, C.c_double, C.c_char_p]
        _lib = L
    return _lib


def vocabulary(seed, size, fixed_words=(), nonascii=0.15):
    """-> list of bytes: the fixed words first (rank order), then generated ones."""
    L = lib()
    words = [w.encode() if isinstance(w, str) else w for w in fixed_words]
    seen = set(words)
    buf = C.create_string_buffer(80)
    rank = 0
    while len(words) < size:
        n = L.corpus_vocab_word(seed, rank, nonascii, buf)
        w = buf.raw[:n]
        rank += 1
        if w in seen:
            continue
        seen.add(w)
        words.append(w)
    return words


def layout(seed, target_bytes, min_doc=4096, max_doc=65536):
    L = lib()
    max_docs = int(target_bytes // max(1, min_doc)) + 2
    offs = np.zeros(max_docs + 1, dtype=np.uint64)
    nd = L.corpus_layout(seed, int(target_bytes), min_doc, max_doc, offs.ctypes.data, max_docs)
    return offs[:nd + 1].copy()


def fill(seed, vocab, text, offs, zipf_s=1.0, nthreads=None):
    """Fills the uint8 array `text` (len >= offs[-1]) with the documents described by offs."""
    L = lib()
    blob = b"".join(vocab)
    voff = np.zeros(len(vocab) + 1, dtype=np.uint32)
    voff[1:] = np.cumsum([len(w) for w in vocab])
    L.corpus_fill(seed, blob, voff.ctypes.data, len(vocab), zipf_s, text.ctypes.data, offs.ctypes.data, len(offs) - 1,
                  nthreads or min(64, os.cpu_count() or 1))


def generate(seed, target_bytes, vocab, min_doc=4096, max_doc=65536, zipf_s=1.0, nthreads=None):
    """-> (text uint8[target_bytes], doc_offsets uint64[ndocs+1])"""
    offs = layout(seed, target_bytes, min_doc, max_doc)
    text = np.zeros(int(target_bytes), dtype=np.uint8)
    fill(seed, vocab, text, offs, zipf_s, nthreads)
    return text, offs
