"""Parity of the CUDA matcher (through the C ABI) with the reference's own level-2 sources
(oracle/_ref/libsp_ref_matcher.so) on the same seeded inputs: results with their items (in emission order for the
state machine, as the multiset per document for the window evaluator), document status and the event cascade counters."""
import json
import os

import numpy as np
import pytest

import matcher_cases as MC
from oracle_util import LEXEM_DT, RefMatcher
from test_oracle_matcher import define_simple_patterns, simple_document

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(params=["statemachine", "auto"], autouse=True)
def mode(request, monkeypatch):
    """Both formulations of level 2: the per-document state machine (results in the reference's emission order) and the
    automatic choice (window evaluator where the rule set allows it: the same results per document, ordered by key token)."""
    monkeypatch.setenv("SP_MATCHER_MODE", request.param)
    return request.param


def both(define, compile_, options=()):
    from struspattern_b200 import Matcher
    ms = []
    for cls in (RefMatcher, Matcher):
        m = cls()
        for name, value in options:
            m.define_option(name, value)
        define(m)
        if compile_:
            m.compile()
        ms.append(m)
    return ms


@pytest.mark.parametrize("compile_", [False, True])
def test_simple_token_pattern_match_kat(compile_, mode):
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    ref, gpu = both(lambda m: define_simple_patterns(m, kat["patterns"]), compile_)
    doc = simple_document()
    ro, go = ref.run(doc, [0, len(doc)]), gpu.run(doc, [0, len(doc)])
    got = {(gpu.result_name(r["name"]), int(r["ordpos"])) for r in go["results"]}
    exp = {(p["name"], pos) for p in kat["patterns"] for pos in p["results"]}
    assert got == exp
    MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed,nrules,nfeatures", [(1, 100, 50), (3, 1000, 1000)])
def test_c1_style_rules(seed, nrules, nfeatures, compile_, mode):
    ref, gpu = both(lambda m: MC.define_c1_rules(m, seed, nrules, nfeatures), compile_)
    toks, offs = MC.random_documents(seed + 100, 40, 300, nfeatures)
    ro, go = ref.run(toks, offs, nthreads=4), gpu.run(toks, offs)
    assert MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine")) > 0


@pytest.mark.parametrize("compile_", [False, True])
def test_stopword_alt_keys(compile_, mode):
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 7, 400, 30, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_STRUCT, MC.OP_WITHIN_STRUCT, MC.OP_SEQUENCE_IMM),
                                                nterms=(2, 4)), compile_)
    toks, offs = MC.random_documents(77, 20, 250, 30)
    ro, go = ref.run(toks, offs, nthreads=4), gpu.run(toks, offs)
    MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))
    if compile_:
        assert int(go["stats"]["nof_altkey_programs_installed"].sum()) > 0


@pytest.mark.parametrize("seed", [11, 12])
def test_expression_trees(seed, mode):
    ref, gpu = both(lambda m: MC.define_tree_rules(m, seed, 120, 40, depth=3), True)
    toks, offs = MC.random_documents(seed + 500, 16, 200, 40)
    ro, go = ref.run(toks, offs, nthreads=4), gpu.run(toks, offs)
    MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))
    assert len(go["items"]) > 0


def test_cardinality_long_ranges_and_exclusive(mode):
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 21, 300, 25, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_ANY, MC.OP_SEQUENCE_IMM, MC.OP_AND),
                                                nterms=(2, 5), card_prob=0.5), True, options=(("exclusive", 0.0), ("maxResultSize", 40.0)))
    toks, offs = MC.random_documents(22, 10, 200, 25, tokens_per_pos=0.7)
    MC.compare_outputs(ref, ref.run(toks, offs), gpu, gpu.run(toks, offs), check_order=(mode == "statemachine"))

    def define(m):
        rng = np.random.default_rng(5)
        for r in range(150):
            terms = rng.integers(1, 20, size=int(rng.integers(2, 4)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_WITHIN, terms, int(rng.integers(60, 200)), 0)
            m.define_pattern("L%u" % r, "", True)
    ref, gpu = both(define, False)
    toks, offs = MC.random_documents(6, 6, 500, 20)
    toks["ordpos"] = toks["ordpos"] * 3 + (toks["ordpos"] // 50) * 100
    MC.compare_outputs(ref, ref.run(toks, offs), gpu, gpu.run(toks, offs), check_order=(mode == "statemachine"))


def test_errors_empty_documents_and_segments(mode):
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 41, 100, 20), True)
    toks, offs = MC.random_documents(42, 5, 50, 20)
    toks["ordpos"][int(offs[1]) + 10] = 1
    offs2 = np.array([offs[0], offs[1], offs[2], offs[3], offs[3], offs[4], offs[5]], dtype=np.uint64)
    ro, go = ref.run(toks, offs2), gpu.run(toks, offs2)
    assert ro["status"][1] != 0 and go["status"][1] != 0
    MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))
    segs = (np.arange(len(toks)) // 17).astype(np.uint32)
    toks["id"][::9] = 0
    MC.compare_outputs(ref, ref.run(toks, offs, origsegs=segs), gpu, gpu.run(toks, offs, origsegs=segs), check_order=(mode == "statemachine"))
    empty = gpu.run(np.zeros(0, dtype=LEXEM_DT), [0])
    assert len(empty["results"]) == 0 and len(empty["status"]) == 0


def test_arena_and_pool_overflow_are_retried(mode):
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 51, 500, 40), True)
    toks, offs = MC.random_documents(52, 64, 300, 40)
    ro = ref.run(toks, offs, nthreads=4)
    ctx = gpu.create_context()
    ctx.counter("arena=8")
    ctx.counter("mode=statemachine")    # arenas only exist on the state machine path
    go = ctx.run(toks, offs)
    assert ctx.counter("arena_retries") > 0
    MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))
    go2 = ctx.run(toks, offs)   # context reuse
    MC.compare_outputs(ref, ro, gpu, go2, check_order=(mode == "statemachine"))


def test_many_documents_c1_shape(mode):
    # BASELINE config 1 shape at reduced size: 10k features, 200 docs x 1000 tokens, 1k rules, optimiser on
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 4711, 1000, 10000), True)
    toks, offs = MC.random_documents(4712, 200, 1000, 10000)
    ro, go = ref.run(toks, offs, nthreads=8), gpu.run(toks, offs)
    n = MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine"))
    assert n > 1000


def test_lexer_output_feeds_the_matcher_on_device(mode):
    from struspattern_b200 import Lexer, Matcher
    from lexer_cases import define_set, make_batch
    lx = define_set(Lexer(), "words_numbers", with_symbols=False)
    text, doffs = make_batch(seed=9, ndocs=32, mean_bytes=2000)
    lctx = lx.create_context()
    lout = lctx.scan(text, doffs)

    def define(m):
        # lexem ids of the set: 1 word, 2 number, 4 '.', 5 punctuation
        for k, (op, terms, rng) in enumerate([(MC.OP_SEQUENCE, (1, 2), 2), (MC.OP_WITHIN, (2, 1), 4), (MC.OP_SEQUENCE_IMM, (1, 1, 5), 3), (MC.OP_ANY, (2, 4), 1)]):
            for i, t in enumerate(terms):
                m.push_term(t); m.attach_variable("v%d" % i)
            m.push_expression(op, len(terms), rng, 0)
            m.define_pattern("p%d" % k, "", True)
    ref, gpu = both(define, True)
    ro = ref.run(lout["lexems"], lout["offsets"], nthreads=4)
    mctx = gpu.create_context()
    mctx.launch_lexer_output(lctx)
    go = mctx.fetch()
    assert MC.compare_outputs(ref, ro, gpu, go, check_order=(mode == "statemachine")) > 0
