"""Product host compiler (regex -> DFA tables) + per-thread kernel bodies, run as CPU loops by tests/hostsim,
against the level-1 oracle and the reference's golden vector. No GPU needed; the GPU tests (test_lexer_gpu.py)
run the same cases through the C ABI and the CUDA kernels."""
import numpy as np
import pytest

from hostsim_util import HostSimLexer
from lexer_cases import LEXEM_SETS, define_set, make_batch
from oracle_util import OracleLexer
from test_oracle_lexer import _golden_cases, build_case


def test_golden_case1_exact():
    case = _golden_cases()[0]
    lx = build_case(HostSimLexer, case)
    lx.compile()
    assert lx.match(case["src_latin1"].encode("latin-1")) == [tuple(r) for r in case["results"]]


@pytest.mark.parametrize("idx", [1, 2])
def test_edit_distance_is_declared_unsupported(idx):
    lx = build_case(HostSimLexer, _golden_cases()[idx])
    with pytest.raises(RuntimeError, match="unsupported"):
        lx.compile()


def _same(ro, rh, captures):
    assert (ro["status"] == rh["status"]).all()
    assert (ro["offsets"] == rh["offsets"]).all()
    assert len(ro["lexems"]) == len(rh["lexems"]) and (ro["lexems"] == rh["lexems"]).all()
    # raw engine events: the oracle lists them before, the product after the sub expression selection
    keep_o = ~np.isin(ro["raw"][:, 0], captures)
    keep_h = ~np.isin(rh["raw"][:, 0], captures)
    assert (ro["raw"][keep_o] == rh["raw"][keep_h]).all()


@pytest.mark.parametrize("name", sorted(LEXEM_SETS))
@pytest.mark.parametrize("chunk,warmup", [(16, 4), (64, 16), (512, 64)])
def test_random_text_matches_oracle(name, chunk, warmup):
    o = define_set(OracleLexer(), name)
    h = define_set(HostSimLexer(chunk=chunk, warmup=warmup), name)
    captures = [i + 1 for i, d in enumerate(LEXEM_SETS[name]) if d[2]]
    text, offs = make_batch(seed=chunk * 1000 + len(name), ndocs=30, mean_bytes=250)
    _same(o.scan(text, offs, nthreads=4, keep_raw=True), h.scan(text, offs, keep_raw=True), captures)


def test_speculation_is_repaired():
    """A lexem that carries state over any distance ('x.*y' with DOTALL) defeats the warm-up guess of the
    chunk start states; the verify/fix-up rounds must still give the exact result."""
    defs = [(1, r"x.*y", 0, 1, 0), (2, r"[a-z]+", 0, 0, 0)]
    o, h = OracleLexer(), HostSimLexer(chunk=32, warmup=8)
    for lx in (o, h):
        lx.define_option("DOTALL")
        for d in defs:
            lx.define_lexem(*d)
        lx.compile()
    text = (b"x" + b"ab " * 100 + b"y " + b"cd " * 50) * 2
    offs = [0, len(text) // 2, len(text)]
    ro, rh = o.scan(text, offs, keep_raw=True), h.scan(text, offs, keep_raw=True)
    _same(ro, rh, [])
    assert h.info("fixups") > 0 and h.info("rounds") > 0


def test_documents_smaller_than_a_chunk_and_empty_documents():
    o = define_set(OracleLexer(), "words_numbers")
    h = define_set(HostSimLexer(chunk=256, warmup=32), "words_numbers")
    docs = [b"", b"the", b"", b"", b"new york 42", b"a", b"", b"3.14, x."]
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    text = b"".join(docs)
    _same(o.scan(text, offs, keep_raw=True), h.scan(text, offs, keep_raw=True), [])
    # all documents empty, and no documents at all
    for offs in ([0, 0, 0], [0]):
        r = h.scan(b"", offs)
        assert len(r["lexems"]) == 0 and (r["offsets"] == 0).all()


def test_double_insert_quirk_and_symbols():
    for cls in (OracleLexer, HostSimLexer):
        lx = cls()
        lx.define_lexem(1, "[a-z]\\b", 0, 0, 0)
        lx.define_lexem(2, "a b c\\b", 0, 0, 0)
        lx.define_symbol(9, 2, "a b c")
        lx.compile()
        assert lx.match("a b c") == [(1, 1, 0, 1), (2, 1, 0, 5), (9, 1, 0, 5), (0, 1, 0, 0), (1, 2, 4, 1)]


def test_token_size_limit_is_an_error_of_the_document():
    h = HostSimLexer(chunk=4096, warmup=64)
    h.define_lexem(1, "\\ba+\\b", 0, 0, 0)
    h.compile()
    text = b"a" * 65535 + b"a" * 65534
    out = h.scan(text, [0, 65535, len(text)])
    assert out["status"][0] != 0 and out["status"][1] == 0
    assert [tuple(int(x) for x in r) for r in out["lexems"]] == [(1, 1, 0, 65534)]


def test_compile_errors_and_state_errors():
    for bad in ["a(", "[a-", "a**", "(?=x)", "\\1", "x*"]:
        lx = HostSimLexer()
        lx.define_lexem(1, bad)
        with pytest.raises(RuntimeError):
            lx.compile()
    lx = HostSimLexer()
    lx.define_lexem(1, "a")
    lx.compile()
    with pytest.raises(RuntimeError):
        lx.define_lexem(2, "b")  # src/patternLexer.cpp:999-1002


def test_unique_elements_across_islands_and_both_formulations():
    """BindUnique behind BindUnique is dropped (src/patternLexer.cpp:926-929) even when the two sit in different
    islands; the per-document loop (mode 0) and the island/slot formulation (mode 1) must agree with the oracle."""
    defs = [(1, r"x+", 0, 0, 3), (2, r"[a-w]+", 0, 0, 0), (3, r"#", 0, 0, 1), (4, r"[.,]", 0, 0, 2), (5, r"y", 0, 1, 3)]
    rng = np.random.default_rng(3)
    docs = []
    for _ in range(40):
        toks = rng.choice(["x", "xx", "y", "ab", "#", ".", ",", "q"], size=int(rng.integers(0, 60)), p=[0.3, 0.1, 0.15, 0.15, 0.1, 0.08, 0.07, 0.05])
        docs.append(" ".join(toks).encode())
    docs += [b"x x x", b". x", b"# . # x x y y ab", b"x", b""]
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    text = b"".join(docs)
    outs = []
    for lx in (OracleLexer(), HostSimLexer(chunk=32, warmup=8, mode=0), HostSimLexer(chunk=32, warmup=8, mode=1)):
        for d in defs:
            lx.define_lexem(*d)
        lx.compile()
        outs.append((lx, lx.scan(text, offs)))
    ref = outs[0][1]
    for lx, out in outs[1:]:
        assert (out["offsets"] == ref["offsets"]).all() and (out["lexems"] == ref["lexems"]).all()
    assert outs[2][0].info("fixed_islands") > 0 and outs[2][0].info("islands") > 100


@pytest.mark.parametrize("name", ["words_numbers", "no_boundaries"])
def test_per_document_loop_still_matches(name):
    o = define_set(OracleLexer(), name)
    h = define_set(HostSimLexer(chunk=64, warmup=16, mode=0), name)
    text, offs = make_batch(seed=77, ndocs=20, mean_bytes=300)
    ro, rh = o.scan(text, offs), h.scan(text, offs)
    assert (ro["offsets"] == rh["offsets"]).all() and (ro["lexems"] == rh["lexems"]).all()


# ---------------------------------------------------------------------------------------------------------
# Fused segment formulation (lexer_fused.cuh, mode 2): the same per-thread functions the kernels k_lex_fused,
# k_fused_fix, k_fused_docs and k_fused_write run.
@pytest.mark.parametrize("mode", [2, 3])
@pytest.mark.parametrize("name", sorted(LEXEM_SETS))
@pytest.mark.parametrize("chunk,warmup", [(16, 4), (48, 16), (512, 64)])
def test_fused_segments_match_oracle(name, chunk, warmup, mode):
    """mode 2: scan + completion + resolution in one thread per chunk; mode 3: chunk resolution over the completed events"""
    o = define_set(OracleLexer(), name)
    h = define_set(HostSimLexer(chunk=chunk, warmup=warmup, mode=mode), name)
    assert h.info("cut_states") > 0
    for seed in (1, 2):
        text, offs = make_batch(seed=seed * 7919 + chunk + len(name), ndocs=40, mean_bytes=300)
        ro, rh = o.scan(text, offs, nthreads=4), h.scan(text, offs)
        assert (ro["status"] == rh["status"]).all() and (ro["offsets"] == rh["offsets"]).all()
        assert len(ro["lexems"]) == len(rh["lexems"]) and (ro["lexems"] == rh["lexems"]).all()
        # every set except the one with BindUnique lexems stays on the fused path
        if name != "no_boundaries":
            assert h.info("fused") == 1 and h.info("fused_legacy") == 0


def test_fused_leading_predecessor_elements_are_decided_by_the_ranges_in_front():
    """Punctuation (BindPredecessor) at the begin of a thread's range is emitted iff its document has a content
    element before: documents that start with punctuation runs longer than a chunk exercise the look back."""
    defs = [(1, r"[a-z]+", 0, 0, 0), (2, r"[.,;]", 0, 0, 2), (3, r"#", 0, 0, 1)]
    docs = [b". , ; . , ; . , ; . , ; . , ; ab . , ; . , ; . , ; . , ; cd .", b", , , , , , , , , , , , , , , , ,",
            b"ab" + b" ." * 40, b"# # . . ab , #", b"", b". ." * 30 + b" x" + b" ;" * 30]
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    text = b"".join(docs)
    o = OracleLexer()
    for d in defs:
        o.define_lexem(*d)
    o.compile()
    ro = o.scan(text, offs)
    for chunk, mode in [(8, 2), (16, 2), (32, 2), (512, 2), (8, 3), (16, 3), (32, 3), (512, 3)]:
        h = HostSimLexer(chunk=chunk, warmup=4, mode=mode)
        for d in defs:
            h.define_lexem(*d)
        h.compile()
        rh = h.scan(text, offs)
        assert (ro["offsets"] == rh["offsets"]).all() and (ro["lexems"] == rh["lexems"]).all()
        assert h.info("fused") == 1
        if chunk <= 32:
            assert h.info("lead_fixed") > 0


def test_fused_falls_back_where_it_must():
    """Successor elements in a document without content and empty selected sub expressions are answered by the island
    pipeline (fused_legacy) with the oracle's output; long chains of overlapping matches (one segment with far more
    open elements than the fast work memory holds) stay on the segment path."""
    cases = [
        ([(1, r"[a-z]+", 0, 0, 0), (3, r"#", 0, 0, 1)], [b"# # #", b"ab # cd"], True),
        ([(1, r"[a-z]+ [a-z]+", 0, 1, 0), (2, r"[a-z]+", 0, 0, 0), (3, r"[a-z]+ [a-z]+ [a-z]+", 0, 2, 0)], [(b"ab cd ef gh " * 40 + b". ") * 3], False),
        ([(1, r"x([0-9]*)y", 1, 0, 0), (2, r"[a-z]+", 0, 0, 0)], [b"ab xy x12y cd xy"], True),
    ]
    for defs, docs, legacy in cases:
        offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
        text = b"".join(docs)
        for mode in (2, 3):
            o, h = OracleLexer(), HostSimLexer(chunk=32, warmup=8, mode=mode)
            for lx in (o, h):
                for d in defs:
                    lx.define_lexem(*d)
                lx.compile()
            ro, rh = o.scan(text, offs), h.scan(text, offs)
            assert (ro["offsets"] == rh["offsets"]).all() and (ro["lexems"] == rh["lexems"]).all()
            assert h.info("fused_legacy") == (1 if legacy else 0) and h.info("fused") == (0 if legacy else 1)


def test_fused_golden_case_and_symbol_quirk():
    case = _golden_cases()[0]
    for mode in (2, 3):
        lx = build_case(lambda: HostSimLexer(chunk=16, warmup=8, mode=mode), case)
        lx.compile()
        assert lx.match(case["src_latin1"].encode("latin-1")) == [tuple(r) for r in case["results"]]
    lx = HostSimLexer(chunk=512, warmup=8, mode=2)
    lx.define_lexem(1, "[a-z]\\b", 0, 0, 0)
    lx.define_lexem(2, "a b c\\b", 0, 0, 0)
    lx.define_symbol(9, 2, "a b c")
    lx.compile()
    assert lx.match("a b c") == [(1, 1, 0, 1), (2, 1, 0, 5), (9, 1, 0, 5), (0, 1, 0, 0), (1, 2, 4, 1)]


def test_compiled_image_round_trip():
    """LexerProgram::serialize / deserialize (sp_lexer_serialize / sp_lexer_deserialize): a lexer loaded from the image of a
    compiled one has the same tables (image of the image is identical) and gives the oracle's output."""
    for name in sorted(LEXEM_SETS):
        a = define_set(HostSimLexer(chunk=64, warmup=16, mode=3), name)
        image = a.serialize()
        assert image[:4] == b"SPLX" and image == a.serialize()
        b = HostSimLexer(chunk=64, warmup=16, mode=3)
        b.deserialize(image)
        assert b.serialize() == image
        for key in ("fwd_states", "fwd_stride", "rev_states", "cut_states", "acc_base"):
            assert a.info(key) == b.info(key)
        text, offs = make_batch(seed=5, ndocs=40, mean_bytes=600)
        o = define_set(OracleLexer(), name)
        ro, rb = o.scan(text, offs), b.scan(text, offs)
        assert (ro["offsets"] == rb["offsets"]).all() and (ro["lexems"] == rb["lexems"]).all()


def test_compiled_image_is_checked():
    a = define_set(HostSimLexer(), sorted(LEXEM_SETS)[0])
    image = a.serialize()
    for bad, what in ((image[:-1], "truncated"), (b"XXXX" + image[4:], "not a lexer image"),
                      (image[:200] + bytes([image[200] ^ 1]) + image[201:], "checksum"), (image + b"\0", "truncated")):
        b = HostSimLexer()
        with pytest.raises(Exception) as err:
            b.deserialize(bad)
        assert what in str(err.value)
    # only an empty lexer takes an image; an uncompiled one has none
    with pytest.raises(Exception):
        a.deserialize(image)
    c = HostSimLexer()
    c.define_lexem(1, "[a-z]+")
    with pytest.raises(Exception):
        c.serialize()


def test_too_large_definition_sets_are_reported_as_unsupported():
    """One forward automaton for all definitions, 16 bit state numbers: a set that needs more states is refused with a message
    that says so (it used to surface as a bare "automaton has too many states" argument error)."""
    from hostsim_util import HostSimLexer
    lx = HostSimLexer()
    for i in range(12):
        lx.define_lexem(i + 1, r"[a-c]*%s[a-c]{%d}x" % ("abc"[i % 3], 9 + i), 0, 1, 0)
    with pytest.raises(RuntimeError) as exc:
        lx.compile()
    assert "automaton states" in str(exc.value) and "12 lexem definitions" in str(exc.value)
