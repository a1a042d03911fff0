"""The window evaluator (struspattern_b200/csrc/matcher_eval.cuh: one fold of fireSignal per (key token, program) pair,
no per-document state) run on the CPU by tests/hostsim, against the reference's own level-2 sources
(oracle/_ref/libsp_ref_matcher.so). Contract: the multiset of complete result records per document (ordered items) and
the exact cascade counters; documents the evaluator flags as order dependent run on the state machine."""
import json
import os

import numpy as np
import pytest

import matcher_cases as MC
from hostsim_util import HostSimMatcher
from oracle_util import RefMatcher
from test_oracle_matcher import define_simple_patterns, simple_document

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def both(define, compile_, options=()):
    ms = []
    for cls in (RefMatcher, HostSimMatcher):
        m = cls(mode=1) if cls is HostSimMatcher else cls()
        for name, value in options:
            m.define_option(name, value)
        define(m)
        if compile_:
            m.compile()
        ms.append(m)
    return ms


def check(ref, sim, toks, offs, origsegs=None, expect_eval=True, max_flagged=None):
    ro, so = ref.run(toks, offs, origsegs=origsegs), sim.run(toks, offs, origsegs=origsegs)
    n = MC.compare_outputs(ref, ro, sim, so, check_order=False)
    if expect_eval:
        assert sim.info("eval_capable") == 1
        assert sim.info("eval_docs") == len(offs) - 1
        if max_flagged is not None:
            assert sim.info("eval_flagged") <= max_flagged
    return n, so


@pytest.mark.parametrize("compile_", [False, True])
def test_simple_token_pattern_match_kat(compile_):
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    ref, sim = both(lambda m: define_simple_patterns(m, kat["patterns"]), compile_)
    doc = simple_document()
    n, so = check(ref, sim, doc, [0, len(doc)], max_flagged=0)
    got = {(sim.result_name(r["name"]), int(r["ordpos"])) for r in so["results"]}
    assert got == {(p["name"], pos) for p in kat["patterns"] for pos in p["results"]}


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed,nrules,nfeatures", [(1, 100, 50), (2, 300, 200), (3, 1000, 1000)])
def test_c1_style_rules(seed, nrules, nfeatures, compile_):
    ref, sim = both(lambda m: MC.define_c1_rules(m, seed, nrules, nfeatures), compile_)
    toks, offs = MC.random_documents(seed + 100, 12, 300, nfeatures)
    n, _ = check(ref, sim, toks, offs, max_flagged=0)
    assert n > 0


@pytest.mark.parametrize("compile_", [False, True])
def test_stopword_alt_keys(compile_):
    ref, sim = both(lambda m: MC.define_c1_rules(m, 7, 400, 30, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_STRUCT, MC.OP_WITHIN_STRUCT, MC.OP_SEQUENCE_IMM),
                                                nterms=(2, 4)), compile_)
    toks, offs = MC.random_documents(77, 10, 250, 30)
    n, so = check(ref, sim, toks, offs)
    if compile_:
        assert int(so["stats"]["nof_altkey_programs_installed"].sum()) > 0
    # 30 features and up to 4 terms: "within" rules that name a term twice are common here, and a token that hits two
    # live triggers of one "within" rule binds the variable of whichever comes first in the reference's bucket array:
    # such documents are handed to the state machine, the comparison above covers both paths


@pytest.mark.parametrize("compile_", [False, True])
def test_cardinality_and_several_tokens_per_position(compile_):
    ref, sim = both(lambda m: MC.define_c1_rules(m, 21, 300, 25, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_ANY, MC.OP_SEQUENCE_IMM),
                                                nterms=(2, 5), card_prob=0.5), compile_)
    toks, offs = MC.random_documents(22, 10, 200, 25, tokens_per_pos=0.7)
    check(ref, sim, toks, offs)


def test_long_ranges():
    def define(m):
        rng = np.random.default_rng(5)
        for r in range(150):
            terms = rng.integers(1, 20, size=int(rng.integers(2, 4)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_WITHIN, terms, int(rng.integers(60, 200)), 0)
            m.define_pattern("L%u" % r, "", True)
    ref, sim = both(define, False)
    toks, offs = MC.random_documents(6, 6, 500, 20)
    toks["ordpos"] = toks["ordpos"] * 3 + (toks["ordpos"] // 50) * 100
    check(ref, sim, toks, offs)


def define_any_nested(m, seed, nrules, nfeatures):
    """The C3 shape: any(t..) with a variable as first or second argument of a sequence / within, with and without
    cardinality 2 on the inner expression, with and without variables on the inner terms."""
    rng = np.random.default_rng(seed)
    feat = MC.Zipf(nfeatures, 0.8)
    for r in range(nrules):
        n = int(rng.integers(2, 5))
        inner_first = rng.random() < 0.5
        if not inner_first:
            m.push_term(int(feat.draw(rng)))
            if rng.random() < 0.7:
                m.attach_variable("head")
        inner = []
        while len(inner) < n:       # distinct alternatives (a term named twice makes the binding order dependent)
            t = int(feat.draw(rng))
            if t not in inner or r % 10 == 0:
                inner.append(t)
        for i, t in enumerate(inner):
            m.push_term(t)
            if rng.random() < 0.7:
                m.attach_variable("v%d" % i)
        m.push_expression(MC.OP_ANY, n, 1, 2 if rng.random() < 0.3 else 0)
        if rng.random() < 0.6:
            m.attach_variable("alt")
        if inner_first:
            m.push_term(int(feat.draw(rng)))
            if rng.random() < 0.7:
                m.attach_variable("tail")
        op = (MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_IMM)[int(rng.integers(0, 3))]
        m.push_expression(op, 2, int(rng.integers(2, 8)), 0)
        m.define_pattern("N%u" % (r % 13), "", True)


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_nested_any_expressions(seed, compile_):
    ref, sim = both(lambda m: (define_any_nested(m, seed, 150, 30), MC.define_c1_rules(m, seed, 100, 30)), compile_)
    toks, offs = MC.random_documents(seed + 40, 8, 250, 30)
    n, so = check(ref, sim, toks, offs)
    assert sim.info("eval_follows") > 0
    assert len(so["items"]) > 0
    assert sim.info("eval_flagged") < len(offs) - 1


def test_references_to_term_patterns():
    """definePattern on a bare term wraps it into an "any" program (src/patternMatcher.cpp:567-575); other patterns
    refer to it by name (pushPattern)."""
    def define(m):
        for t in range(1, 8):
            m.push_term(t)
            m.attach_variable("w")
            m.define_pattern("T%u" % t, "", t % 2 == 0)
        rng = np.random.default_rng(3)
        for r in range(60):
            m.push_pattern("T%u" % int(rng.integers(1, 8)))
            if r % 3:
                m.attach_variable("a")
            m.push_term(int(rng.integers(1, 12)))
            m.attach_variable("b")
            m.push_expression((MC.OP_SEQUENCE, MC.OP_WITHIN)[r % 2], 2, int(rng.integers(1, 6)), 0)
            m.define_pattern("R%u" % r, "", True)
    for compile_ in (False, True):
        ref, sim = both(define, compile_)
        toks, offs = MC.random_documents(9, 6, 200, 12)
        check(ref, sim, toks, offs)


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [11, 12, 13, 14])
def test_expression_trees_use_the_state_machine(seed, compile_):
    # nested sequences / withins emit their events late: not a rule set of the evaluator, all documents fall back
    ref, sim = both(lambda m: MC.define_tree_rules(m, seed, 120, 40, depth=3), compile_)
    toks, offs = MC.random_documents(seed + 500, 8, 200, 40)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)


def test_irregular_documents_fall_back():
    ref, sim = both(lambda m: MC.define_c1_rules(m, 41, 100, 20), True)
    toks, offs = MC.random_documents(42, 5, 50, 20)
    b = int(offs[1])
    toks["ordpos"][b + 10] = 1
    offs2 = np.array([offs[0], offs[1], offs[2], offs[3], offs[3], offs[4], offs[5]], dtype=np.uint64)
    ro, so = ref.run(toks, offs2), sim.run(toks, offs2)
    assert ro["status"][1] != 0 and so["status"][1] != 0
    MC.compare_outputs(ref, ro, sim, so, check_order=False)
    assert sim.info("eval_flagged") == 1
    segs = (np.arange(len(toks)) // 17).astype(np.uint32)
    toks["id"][::9] = 0
    check(ref, sim, toks, offs, origsegs=segs, expect_eval=False)


def test_c4_style_dictionary():
    from struspattern_b200 import workloads as W
    ref, sim = both(lambda m: W.define_c4_patterns(m, 5, 3000, 400), True)
    toks, offs = W.c4_token_documents(6, 6, 400, 150, 300)
    n, so = check(ref, sim, toks, offs)
    assert int(so["stats"]["nof_altkey_programs_installed"].sum()) > 0
    assert n > 0


def test_c3_style_patterns():
    from struspattern_b200 import workloads as W
    ref, sim = both(lambda m: W.define_c3_patterns(m, 33, 1500, 60), True)
    rng = np.random.default_rng(8)
    ndocs, n = 6, 400
    toks = np.zeros(ndocs * n, dtype=MC.LEXEM_DT)
    ids = np.where(rng.random(ndocs * n) < 0.25, rng.integers(1, 14, ndocs * n), W.SYMBOL_BASE + MC.Zipf(60, 1.0).draw(rng, ndocs * n) - 1)
    toks["id"] = ids
    toks["ordpos"] = np.tile(np.arange(1, n + 1), ndocs) - (np.tile(np.arange(n), ndocs) % 5 == 4)  # some tokens share a position
    toks["origpos"] = np.tile(np.arange(n) * 4, ndocs)
    toks["origsize"] = 3
    offs = np.arange(ndocs + 1, dtype=np.uint64) * n
    cnt, so = check(ref, sim, toks, offs)
    assert cnt > 0 and len(so["items"]) > 0
    assert sim.info("eval_flagged") < ndocs


def test_sequences_that_name_a_term_several_times_stay_on_the_evaluator():
    """sequence(a, b, a, a): one token hits up to three live triggers of a rule, but only the one whose signal value is due
    accepts it, whatever the order of the reference's trigger array: no document is handed to the state machine."""
    def define(m):
        rng = np.random.default_rng(17)
        for r in range(200):
            terms = rng.integers(1, 6, size=int(rng.integers(3, 6)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_SEQUENCE_IMM, terms, int(rng.integers(3, 9)), 0)
            m.define_pattern("S%u" % r, "", True)
    for compile_ in (False, True):
        ref, sim = both(define, compile_)
        toks, offs = MC.random_documents(18, 8, 300, 6)
        n, _ = check(ref, sim, toks, offs, max_flagged=0)
        assert n > 0
