"""The product's rule-graph builder (matcher_program.hpp) and per-document state machine (matcher_device.cuh),
run on the CPU by tests/hostsim, against the reference's own level-2 sources (oracle/_ref/libsp_ref_matcher.so)."""
import json
import os

import numpy as np
import pytest

import matcher_cases as MC
from hostsim_util import HostSimMatcher
from oracle_util import LEXEM_DT, RefMatcher
from test_oracle_matcher import define_simple_patterns, simple_document

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def both(define, compile_, caps=None, options=()):
    ms = []
    for cls in (RefMatcher, HostSimMatcher):
        m = cls(caps) if (caps and cls is HostSimMatcher) else cls()
        for name, value in options:
            m.define_option(name, value)
        define(m)
        if compile_:
            m.compile()
        ms.append(m)
    return ms


@pytest.mark.parametrize("compile_", [False, True])
def test_simple_token_pattern_match_kat(compile_):
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    ref, sim = both(lambda m: define_simple_patterns(m, kat["patterns"]), compile_)
    doc = simple_document()
    ro, so = ref.run(doc, [0, len(doc)]), sim.run(doc, [0, len(doc)])
    got = {(sim.result_name(r["name"]), int(r["ordpos"])) for r in so["results"]}
    exp = {(p["name"], pos) for p in kat["patterns"] for pos in p["results"]}
    assert got == exp
    MC.compare_outputs(ref, ro, sim, so, check_order=True)


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed,nrules,nfeatures", [(1, 100, 50), (2, 300, 200), (3, 1000, 1000)])
def test_c1_style_rules(seed, nrules, nfeatures, compile_):
    ref, sim = both(lambda m: MC.define_c1_rules(m, seed, nrules, nfeatures), compile_)
    toks, offs = MC.random_documents(seed + 100, 12, 300, nfeatures)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    n = MC.compare_outputs(ref, ro, sim, so, check_order=True)
    assert n > 0


@pytest.mark.parametrize("compile_", [False, True])
def test_stopword_alt_keys_are_exercised(compile_):
    # few features -> every key event is a stop word candidate; ranges <= maxRange(5) move programs to alt keys
    ref, sim = both(lambda m: MC.define_c1_rules(m, 7, 400, 30, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_STRUCT, MC.OP_WITHIN_STRUCT, MC.OP_SEQUENCE_IMM),
                                                nterms=(2, 4)), compile_)
    toks, offs = MC.random_documents(77, 10, 250, 30)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)
    if compile_:
        assert int(so["stats"]["nof_altkey_programs_installed"].sum()) > 0
        assert sim.info("stopwords") > 0


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [11, 12, 13, 14])
def test_expression_trees(seed, compile_):
    ref, sim = both(lambda m: MC.define_tree_rules(m, seed, 120, 40, depth=3), compile_)
    toks, offs = MC.random_documents(seed + 500, 8, 200, 40)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)
    assert len(so["items"]) > 0


@pytest.mark.parametrize("compile_", [False, True])
def test_cardinality_and_long_terms(compile_):
    ref, sim = both(lambda m: MC.define_c1_rules(m, 21, 300, 25, ops=(MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_ANY, MC.OP_SEQUENCE_IMM, MC.OP_AND),
                                                nterms=(2, 5), card_prob=0.5), compile_)
    toks, offs = MC.random_documents(22, 10, 200, 25, tokens_per_pos=0.7)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)


def test_long_ranges_use_the_expiry_heap():
    def define(m):
        rng = np.random.default_rng(5)
        for r in range(150):
            terms = rng.integers(1, 20, size=int(rng.integers(2, 4)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_WITHIN, terms, int(rng.integers(60, 200)), 0)
            m.define_pattern("L%u" % r, "", True)
    ref, sim = both(define, False)
    toks, offs = MC.random_documents(6, 6, 500, 20)
    # sparse positions: jumps over the 64 position window
    toks["ordpos"] = toks["ordpos"] * 3 + (toks["ordpos"] // 50) * 100
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)


def test_exclusive_option():
    opts = (("exclusive", 0.0), ("maxResultSize", 40.0))
    ref, sim = both(lambda m: MC.define_c1_rules(m, 31, 200, 20, nterms=(2, 4)), True, options=opts)
    toks, offs = MC.random_documents(32, 6, 200, 20)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)
    plain = HostSimMatcher()
    MC.define_c1_rules(plain, 31, 200, 20, nterms=(2, 4))
    plain.compile()
    assert len(plain.run(toks, offs)["results"]) > len(so["results"])


def test_errors_and_edge_documents():
    ref, sim = both(lambda m: MC.define_c1_rules(m, 41, 100, 20), True)
    toks, offs = MC.random_documents(42, 5, 50, 20)
    # document 1: positions not ascending; document 3: empty
    b, e = int(offs[1]), int(offs[2])
    toks["ordpos"][b + 10] = 1
    offs2 = np.array([offs[0], offs[1], offs[2], offs[3], offs[3], offs[4], offs[5]], dtype=np.uint64)
    ro, so = ref.run(toks, offs2), sim.run(toks, offs2)
    assert ro["status"][1] != 0 and so["status"][1] != 0
    MC.compare_outputs(ref, ro, sim, so, check_order=True)
    # term id 0 and segments
    segs = (np.arange(len(toks)) // 17).astype(np.uint32)
    toks["id"][::9] = 0
    ro, so = ref.run(toks, offs, origsegs=segs), sim.run(toks, offs, origsegs=segs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)


def test_definition_errors_match_the_reference():
    for bad in (lambda m: m.push_expression(MC.OP_SEQUENCE, 2, 1, 0),                      # empty stack
                lambda m: (m.push_term(1), m.attach_variable("a"), m.attach_variable("b")),  # two variables
                lambda m: m.define_pattern("x", "", True),                                  # nothing on the stack
                lambda m: ([m.push_term(i + 1) for i in range(33)], m.push_expression(MC.OP_WITHIN, 33, 1, 0)),
                lambda m: m.define_option("nonsense", 1.0),
                lambda m: m.push_term(1 << 29)):
        for cls in (RefMatcher, HostSimMatcher):
            with pytest.raises(RuntimeError):
                bad(cls())


def test_arena_overflow_is_retried_with_larger_arenas():
    ref, sim = both(lambda m: MC.define_c1_rules(m, 51, 500, 40), True, caps=(8, 8, 8, 8, 8, 4, 8, 4))
    toks, offs = MC.random_documents(52, 4, 300, 40)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    MC.compare_outputs(ref, ro, sim, so, check_order=True)
    assert sim.retries() > 0
