"""Parity of the CUDA lexer (through the C ABI) with the level-1 oracle and the reference's golden vector."""
import os

import numpy as np
import pytest

from lexer_cases import LEXEM_SETS, define_set, make_batch
from oracle_util import OracleLexer
from test_oracle_lexer import _golden_cases, build_case

pytestmark = pytest.mark.gpu


def _gpu_lexer():
    from struspattern_b200 import Lexer
    return Lexer


def test_golden_case1_exact():
    case = _golden_cases()[0]
    lx = build_case(_gpu_lexer(), case)
    lx.compile()
    assert lx.match(case["src_latin1"].encode("latin-1")) == [tuple(r) for r in case["results"]]


def _same(ro, rg):
    assert (ro["status"] == rg["status"]).all()
    assert (ro["offsets"] == rg["offsets"]).all()
    assert len(ro["lexems"]) == len(rg["lexems"]) and (ro["lexems"] == rg["lexems"]).all()
    assert ro["raw_event_count"] == rg["raw_event_count"]


@pytest.mark.parametrize("name", sorted(LEXEM_SETS))
def test_random_text_matches_oracle(name):
    o = define_set(OracleLexer(), name)
    g = define_set(_gpu_lexer()(), name)
    text, offs = make_batch(seed=77 + len(name), ndocs=300, mean_bytes=2000)
    _same(o.scan(text, offs, nthreads=8), g.scan(text, offs))


def test_device_resident_text_and_context_reuse():
    import torch
    o = define_set(OracleLexer(), "phrases_dates")
    g = define_set(_gpu_lexer()(), "phrases_dates")
    ctx = g.create_context()
    for seed in (1, 2, 3):
        text, offs = make_batch(seed=seed, ndocs=50 * seed, mean_bytes=700)
        dev = torch.from_numpy(text).cuda()
        _same(o.scan(text, offs, nthreads=8), ctx.scan(dev.data_ptr(), offs, on_device=True))
        if seed == 2:
            assert ctx.counter("sync:blocking") == 1     # waits sleep on an event from here on (many contexts per host)


def test_speculation_is_repaired():
    defs = [(1, r"x.*y", 0, 1, 0), (2, r"[a-z]+", 0, 0, 0)]
    o, g = OracleLexer(), _gpu_lexer()()
    for lx in (o, g):
        lx.define_option("DOTALL")
        for d in defs:
            lx.define_lexem(*d)
        lx.compile()
    text = (b"x" + b"ab " * 1000 + b"y " + b"cd " * 500) * 2
    offs = [0, len(text) // 2, len(text)]
    ctx = g.create_context()
    _same(o.scan(text, offs), ctx.scan(text, offs))
    assert ctx.counter("fixup_chunks") > 0


def test_edge_cases():
    o = define_set(OracleLexer(), "words_numbers")
    g = define_set(_gpu_lexer()(), "words_numbers")
    docs = [b"", b"the", b"", b"", b"new york 42", b"a", b"", b"3.14, x."]
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    _same(o.scan(b"".join(docs), offs), g.scan(b"".join(docs), offs))
    for offs in ([0, 0, 0], [0]):
        r = g.scan(b"", offs)
        assert len(r["lexems"]) == 0 and (r["offsets"] == 0).all()
    # token size limit: error of that document only (src/patternLexer.cpp:727-730)
    g2 = _gpu_lexer()()
    g2.define_lexem(1, "\\ba+\\b", 0, 0, 0)
    g2.compile()
    text = b"a" * 65535 + b"a" * 65534
    out = g2.scan(text, [0, 65535, len(text)])
    assert out["status"][0] != 0 and out["status"][1] == 0
    assert [tuple(int(x) for x in r) for r in out["lexems"]] == [(1, 1, 0, 65534)]


def test_double_insert_quirk_and_symbols():
    lx = _gpu_lexer()()
    lx.define_lexem(1, "[a-z]\\b", 0, 0, 0)
    lx.define_lexem(2, "a b c\\b", 0, 0, 0)
    lx.define_symbol(9, 2, "a b c")
    lx.compile()
    assert lx.match("a b c") == [(1, 1, 0, 1), (2, 1, 0, 5), (9, 1, 0, 5), (0, 1, 0, 0), (1, 2, 4, 1)]


def test_unique_elements_across_islands_and_all_three_formulations(monkeypatch):
    """BindUnique behind BindUnique across island borders; the three second halves of the pipeline (per document
    loop, dense island pipeline, slot pipeline = default) must give the oracle's output."""
    defs = [(1, r"x+", 0, 0, 3), (2, r"[a-w]+", 0, 0, 0), (3, r"#", 0, 0, 1), (4, r"[.,]", 0, 0, 2), (5, r"y", 0, 1, 3)]
    rng = np.random.default_rng(3)
    docs = []
    for _ in range(400):
        toks = rng.choice(["x", "xx", "y", "ab", "#", ".", ",", "q"], size=int(rng.integers(0, 200)), p=[0.3, 0.1, 0.15, 0.15, 0.1, 0.08, 0.07, 0.05])
        docs.append(" ".join(toks).encode())
    docs += [b"x x x", b". x", b"# . # x x y y ab", b"x", b""]
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    text = np.frombuffer(b"".join(docs), dtype=np.uint8)
    o, g = OracleLexer(), _gpu_lexer()()
    for lx in (o, g):
        for d in defs:
            lx.define_lexem(*d)
        lx.compile()
    ro = o.scan(text, offs, nthreads=8)
    for env in (None, "SP_LEXER_HYBRID"):   # default (chunk resolution) and the island/slot pipeline
        if env:
            monkeypatch.setenv(env, "0")
        ctx = g.create_context()
        _same(ro, ctx.scan(text, offs))
        ctx.close()
        if env:
            monkeypatch.delenv(env)


def test_large_islands_use_their_slot_range():
    # one long chain of overlapping matches = one island far larger than the thread local array
    defs = [(1, r"[a-z]+ [a-z]+", 0, 1, 0), (2, r"[a-z]+", 0, 0, 0), (3, r"[a-z]+ [a-z]+ [a-z]+", 0, 2, 0)]
    o, g = OracleLexer(), _gpu_lexer()()
    for lx in (o, g):
        for d in defs:
            lx.define_lexem(*d)
        lx.define_symbol(77, 2, "ab")
        lx.compile()
    text = (b"ab cd ef gh " * 40 + b". ") * 20
    offs = [0, len(text) // 2 + 3, len(text)]
    _same(o.scan(text, offs), g.scan(text, offs))


# ---------------------------------------------------------------------------------------------------------
# Segment formulations: chunk resolution over the completed events (k_lex_resolve_chunks, the default second half,
# counter second_half == 4) and the fully fused scan kernel (k_lex_fused, SP_LEXER_FUSED=1, second_half == 3); both
# end in k_fused_fix / AggScan / k_fused_docs / k_fused_write.
@pytest.mark.parametrize("fused", [0, 1])
@pytest.mark.parametrize("name", sorted(LEXEM_SETS))
@pytest.mark.parametrize("chunk", [16, 48, 512])
def test_fused_kernel_matches_oracle(name, chunk, fused, monkeypatch):
    monkeypatch.setenv("SP_LEXER_FUSED", str(fused))
    monkeypatch.setenv("SP_LEXER_CHUNK", str(chunk))
    monkeypatch.setenv("SP_LEXER_WARMUP", str(max(4, chunk // 8)))
    o = define_set(OracleLexer(), name)
    g = define_set(_gpu_lexer()(), name)
    ctx = g.create_context()
    text, offs = make_batch(seed=chunk + 31 * len(name), ndocs=200, mean_bytes=1500)
    _same(o.scan(text, offs, nthreads=8), ctx.scan(text, offs))
    assert ctx.counter("chunk_bytes") == chunk
    if name != "no_boundaries":
        assert ctx.counter("second_half") == (3 if fused else 4) and ctx.counter("fused_fallbacks") == 0
    ctx.close()


def test_fused_kernel_equals_island_pipeline_on_the_bench_lexem_set(monkeypatch):
    import sys, os
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from struspattern_b200 import workloads as W
    from bench_util import corpus
    g = W.define_lexer(_gpu_lexer()())
    vocab = corpus.vocabulary(11, 5000, W.PHRASE_FIRST + W.PHRASE_SECOND + W.ABBREVIATIONS)
    nbytes = 8 << 20
    offs = corpus.layout(12, nbytes)
    text = np.zeros(nbytes, dtype=np.uint8)
    corpus.fill(12, vocab, text, offs)
    ctx = g.create_context()
    rf = ctx.scan(text, offs)
    assert ctx.counter("second_half") == 4
    ctx.close()
    monkeypatch.setenv("SP_LEXER_FUSED", "1")
    ctx = g.create_context()
    rl = ctx.scan(text, offs)
    assert ctx.counter("second_half") == 3
    ctx.close()
    _same(rl, rf)
    monkeypatch.setenv("SP_LEXER_FUSED", "0")
    monkeypatch.setenv("SP_LEXER_HYBRID", "0")
    ctx = g.create_context()
    rl = ctx.scan(text, offs)
    assert ctx.counter("second_half") == 2
    ctx.close()
    _same(rl, rf)
    o = W.define_lexer(OracleLexer())
    nd = 40
    ro = o.scan(text[:int(offs[nd])], offs[:nd + 1], nthreads=8)
    assert (ro["lexems"] == rf["lexems"][:len(ro["lexems"])]).all() and (ro["offsets"] == rf["offsets"][:nd + 1]).all()


def test_fused_kernel_falls_back_to_the_island_pipeline():
    cases = [
        ([(1, r"[a-z]+", 0, 0, 0), (3, r"#", 0, 0, 1)], [b"# # #", b"ab # cd"]),
        ([(1, r"x([0-9]*)y", 1, 0, 0), (2, r"[a-z]+", 0, 0, 0)], [b"ab xy x12y cd xy"]),
    ]
    for defs, docs in cases:
        offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
        o, g = OracleLexer(), _gpu_lexer()()
        for lx in (o, g):
            for d in defs:
                lx.define_lexem(*d)
            lx.compile()
        for fused in ("0", "1"):
            os.environ["SP_LEXER_FUSED"] = fused
            try:
                ctx = g.create_context()
                _same(o.scan(b"".join(docs), offs), ctx.scan(b"".join(docs), offs))
                assert ctx.counter("second_half") == 2 and ctx.counter("fused_fallbacks") == 1
                ctx.close()
            finally:
                del os.environ["SP_LEXER_FUSED"]


def test_full_size_corpus_properties(monkeypatch):
    """BASELINE.json configs[1] at its full size (1 GiB of text, the bench corpus): too large for the oracle as a whole, so the
    output is checked through properties that do not depend on the size -- documents are closed units
    (src/patternLexer.cpp:858-950: all state is per match() call), so any split of the batch gives the same lexems; a sample
    of documents spread over the batch is compared with the oracle bit by bit; the slot pipeline gives the same output."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from struspattern_b200 import workloads as W
    nbytes = int(os.environ.get("SP_TEST_FULL_BYTES", 1 << 30))
    text, offs, _ = bench.make_corpus(nbytes, 0, False, bench.vocabulary())
    offs = np.ascontiguousarray(offs, dtype=np.uint64)
    ndocs = len(offs) - 1
    g = W.define_lexer(_gpu_lexer()())
    ctx = g.create_context()
    whole = ctx.scan(text, offs)
    assert ctx.counter("second_half") == 4 and ctx.counter("fused_fallbacks") == 0
    lex, lo = whole["lexems"], whole["offsets"]
    assert (whole["status"] == 0).all() and len(lo) == ndocs + 1 and int(lo[-1]) == len(lex)
    # structure: positions inside the document, ordinal positions start at 1 and never fall, start offsets never fall
    doc_of = np.repeat(np.arange(ndocs, dtype=np.int64), np.diff(lo).astype(np.int64))
    doclen = np.diff(offs).astype(np.int64)
    assert (lex["origpos"].astype(np.int64) + lex["origsize"] <= doclen[doc_of]).all()
    first = lo[:-1][np.diff(lo) > 0].astype(np.int64)
    assert (lex["ordpos"][first] == 1).all()
    inner = np.ones(len(lex), dtype=bool); inner[first] = False
    step = np.diff(lex["ordpos"].astype(np.int64), prepend=0)
    assert ((step == 0) | (step == 1))[inner].all()
    assert (np.diff(lex["origpos"].astype(np.int64), prepend=0)[inner] >= 0).all()
    del doc_of, inner, step
    # any split of the batch at document borders gives the same lexems
    for cut in (ndocs // 3, ndocs // 2 + 1):
        a = ctx.scan(text[:int(offs[cut])], offs[:cut + 1])
        b = ctx.scan(text[int(offs[cut]):], offs[cut:] - offs[cut])
        assert int(a["offsets"][-1]) + int(b["offsets"][-1]) == len(lex)
        assert (a["lexems"] == lex[:len(a["lexems"])]).all() and (b["lexems"] == lex[len(a["lexems"]):]).all()
        assert (a["offsets"] == lo[:cut + 1]).all() and (b["offsets"] + lo[cut] == lo[cut:]).all()
        assert a["raw_event_count"] + b["raw_event_count"] == whole["raw_event_count"]
        del a, b
    ctx.close()
    # a sample of documents from all over the batch against the oracle
    o = W.define_lexer(OracleLexer())
    pick = np.unique(np.linspace(0, ndocs - 1, 48).astype(np.int64))
    stext = np.concatenate([text[int(offs[d]):int(offs[d + 1])] for d in pick])
    soffs = np.concatenate([[0], np.cumsum(doclen[pick])]).astype(np.uint64)
    ro = o.scan(stext, soffs, nthreads=8)
    for k, d in enumerate(pick):
        mine = lex[int(lo[d]):int(lo[d + 1])]
        ref = ro["lexems"][int(ro["offsets"][k]):int(ro["offsets"][k + 1])]
        assert len(mine) == len(ref) and (mine == ref).all(), "document %d differs from the oracle" % d
    # the island/slot formulation gives the same bytes
    monkeypatch.setenv("SP_LEXER_HYBRID", "0")
    ctx = g.create_context()
    other = ctx.scan(text, offs)
    assert ctx.counter("second_half") == 2
    ctx.close()
    assert (other["offsets"] == lo).all() and (other["lexems"] == lex).all() and other["raw_event_count"] == whole["raw_event_count"]


def test_compiled_image_gives_the_same_lexer():
    """sp_lexer_serialize / sp_lexer_deserialize: the lexer loaded from an image scans like the one that was compiled."""
    import time
    from struspattern_b200 import workloads as W
    t0 = time.perf_counter()
    a = W.define_lexer(_gpu_lexer()(), symbols=[(900 + i, W.WORD, w) for i, w in enumerate(("the", "of", "and"))])
    t_compile = time.perf_counter() - t0
    image = a.serialize()
    t0 = time.perf_counter()
    b = _gpu_lexer()()
    b.deserialize(image)
    t_load = time.perf_counter() - t0
    assert b.serialize() == image and b.table_info("fwd_states") == a.table_info("fwd_states")
    assert b.get_symbol(W.WORD, "of") == 901 and t_load < t_compile
    text, offs = make_batch(seed=9, ndocs=200, mean_bytes=3000)
    _same(a.scan(text, offs), b.scan(text, offs))
    with pytest.raises(Exception):
        _gpu_lexer()().deserialize(image[:-3])


def test_one_large_document(monkeypatch):
    """A single document is scanned by as many threads as it has chunks (the reference scans one block
    below 4 GiB sequentially, src/patternLexer.cpp:866): 4 MiB against the oracle, 128 MiB through two formulations."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    from struspattern_b200 import workloads as W
    text, _, _ = bench.make_corpus(128 << 20, 0, False, bench.vocabulary())
    g = W.define_lexer(_gpu_lexer()())
    o = W.define_lexer(OracleLexer())
    small = np.array([0, 4 << 20], dtype=np.uint64)
    _same(o.scan(text[:4 << 20], small), g.scan(text[:4 << 20], small))
    one = np.array([0, len(text)], dtype=np.uint64)
    ctx = g.create_context()
    a = ctx.scan(text, one)
    assert ctx.counter("second_half") == 4 and int(a["status"][0]) == 0
    ctx.close()
    lex = a["lexems"]
    assert len(lex) > 10_000_000 and int(lex["ordpos"][0]) == 1
    step = np.diff(lex["ordpos"].astype(np.int64))
    assert ((step == 0) | (step == 1)).all() and (np.diff(lex["origpos"].astype(np.int64)) >= 0).all()
    assert int((lex["origpos"].astype(np.int64) + lex["origsize"]).max()) <= len(text)
    # the first 4 MiB of the large document give the same lexems as the small one up to the last complete match
    b = g.scan(text[:4 << 20], small)["lexems"]
    keep = b["origpos"].astype(np.int64) + b["origsize"] < (4 << 20) - 4096
    n = int(keep.sum())
    assert keep[:n].all() and (lex[:n] == b[:n]).all()
    monkeypatch.setenv("SP_LEXER_HYBRID", "0")
    ctx = g.create_context()
    c = ctx.scan(text, one)
    assert ctx.counter("second_half") == 2
    ctx.close()
    _same(a, c)


@pytest.mark.parametrize("name", ["words_numbers", "unicode_case"])
def test_tma_staged_text_tiles_give_the_same_output(name):
    """k_lex_count<true> / k_lex_emit<true> read their chunks from shared memory tiles filled by the bulk copy engine
    (cp.async.bulk + mbarrier) with SP_LEXER_TMA=1 / counter "tma=1"; by default they read global memory. Same output, also for batches
    whose last chunk is partial, for text at an unaligned device address (no staging) and for tiny batches."""
    import torch
    o = define_set(OracleLexer(), name)
    g = define_set(_gpu_lexer()(), name)
    for seed, ndocs, mean in ((5, 400, 3000), (6, 3, 90), (7, 1, 511), (8, 1, 512), (9, 700, 700)):
        text, offs = make_batch(seed=seed, ndocs=ndocs, mean_bytes=mean)
        ro = o.scan(text, offs, nthreads=8)
        outs = []
        for key in ("tma=1", "tma=0"):
            ctx = g.create_context()
            assert ctx.counter(key) == int(key[-1])
            outs.append(ctx.scan(text, offs))
        _same(ro, outs[0]); _same(ro, outs[1])
        dev = torch.zeros(len(text) + 64, dtype=torch.uint8, device="cuda")
        dev[3:3 + len(text)] = torch.from_numpy(text).cuda()
        ctx = g.create_context()
        _same(ro, ctx.scan(dev.data_ptr() + 3, offs, on_device=True))      # unaligned: falls back to global reads
