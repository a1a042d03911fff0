"""BASELINE configs 3 and 4 at sizes the oracles finish in seconds: the full two-level pipeline (lexer output stays in HBM
and feeds the matcher) and the dictionary-style high fan-out rule set, CUDA path through the C ABI against the level-1
oracle + the reference's own automaton on the same seeded inputs."""
import os
import sys

import numpy as np
import pytest

import matcher_cases as MC
from oracle_util import OracleLexer, RefMatcher

pytestmark = pytest.mark.gpu


@pytest.fixture(params=["statemachine", "auto"], autouse=True)
def mode(request, monkeypatch):
    monkeypatch.setenv("SP_MATCHER_MODE", request.param)
    return request.param


ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SEED = 0x5354525553000003


def corpus(nbytes, nsymbols):
    sys.path.insert(0, ROOT)
    from bench_util import corpus as CG
    from struspattern_b200 import workloads as W
    fixed = W.PHRASE_FIRST + W.PHRASE_SECOND + W.ABBREVIATIONS + ["the", "of", "and", "to", "in", "a", "is", "that", "for", "it"]
    vocab = CG.vocabulary(SEED, 5000, fixed)
    text, offs = CG.generate(SEED + 1, nbytes, vocab, min_doc=2048, max_doc=16384)
    return text, offs, W.symbol_table(vocab, nsymbols)


def test_c3_full_pipeline_small(mode):
    from struspattern_b200 import Lexer, Matcher
    from struspattern_b200 import workloads as W
    text, offs, symbols = corpus(3 << 20, 800)
    olx = W.define_lexer(OracleLexer(), nphrases=140, symbols=symbols)
    glx = W.define_lexer(Lexer(), nphrases=140, symbols=symbols)
    assert glx.table_info("definitions") >= 180
    ro = olx.scan(text, offs, nthreads=8)
    lctx = glx.create_context()
    go = lctx.scan(text, offs)
    assert (ro["offsets"] == go["offsets"]).all() and (ro["lexems"] == go["lexems"]).all()
    assert int((go["lexems"]["id"] >= W.SYMBOL_BASE).sum()) > 10000     # symbol lexems are produced
    ref, gpu = RefMatcher(), Matcher()
    for m in (ref, gpu):
        W.define_c3_patterns(m, 33, 3000, 800)
        m.compile()
    rm = ref.run(ro["lexems"], ro["offsets"], nthreads=8)
    mctx = gpu.create_context()
    mctx.launch_lexer_output(lctx)          # lexems never leave the device
    gm = mctx.fetch()
    n = MC.compare_outputs(ref, rm, gpu, gm, check_order=(mode == "statemachine"))
    assert n > 500 and len(gm["items"]) > n
    if mode == "auto":      # the C3 rule set is one of the window evaluator; few documents may need the state machine
        assert mctx.counter("eval_capable") == 1 and mctx.counter("eval_docs") == len(offs) - 1
        assert mctx.counter("eval_flagged") * 10 <= len(offs) - 1


@pytest.mark.parametrize("compile_", [False, True])
def test_c4_dictionary_patterns_small(compile_, mode):
    from struspattern_b200 import Matcher
    from struspattern_b200 import workloads as W
    ref, gpu = RefMatcher(), Matcher()
    for m in (ref, gpu):
        W.define_c4_patterns(m, 44, 20000, 3000)
        if compile_:
            m.compile()
    toks, offs = W.c4_token_documents(45, 48, 3000, 300, 900)
    rm = ref.run(toks, offs, nthreads=8)
    gm = gpu.run(toks, offs)
    n = MC.compare_outputs(ref, rm, gpu, gm, check_order=(mode == "statemachine"))
    assert n > 1000
    if compile_:
        assert int(gm["stats"]["nof_altkey_programs_installed"].sum()) > 0
