"""The window evaluator on the GPU (k_eval_pass, struspattern_b200/csrc/matcher_eval_kernels.cuh) through the C ABI against the
reference's own automaton: result multisets per document with ordered items, exact cascade counters, and the hand-over
of flagged documents to the state machine inside one batch."""
import numpy as np
import pytest

import matcher_cases as MC
from oracle_util import LEXEM_DT, RefMatcher
from test_matcher_eval_hostsim import define_any_nested

pytestmark = pytest.mark.gpu


def both(define, compile_):
    from struspattern_b200 import Matcher
    ms = []
    for cls in (RefMatcher, Matcher):
        m = cls()
        define(m)
        if compile_:
            m.compile()
        ms.append(m)
    return ms


def run_eval(gpu, toks, offs, origsegs=None):
    ctx = gpu.create_context()
    ctx.counter("mode=eval")
    out = ctx.run(toks, offs, origsegs)
    return ctx, out


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [1, 2])
def test_nested_any_expressions(seed, compile_):
    ref, gpu = both(lambda m: (define_any_nested(m, seed, 150, 30), MC.define_c1_rules(m, seed, 100, 30)), compile_)
    toks, offs = MC.random_documents(seed + 40, 40, 250, 30)
    ctx, go = run_eval(gpu, toks, offs)
    assert MC.compare_outputs(ref, ref.run(toks, offs, nthreads=4), gpu, go, check_order=False) > 0
    assert gpu.table_info("eval_follows") > 0
    assert ctx.counter("eval_flagged") < len(offs) - 1


def test_no_flags_on_plain_c1_rules_and_context_reuse():
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 3, 1000, 1000), True)
    toks, offs = MC.random_documents(103, 300, 300, 1000)
    ro = ref.run(toks, offs, nthreads=8)
    ctx, go = run_eval(gpu, toks, offs)
    assert ctx.counter("eval_flagged") == 0
    n = MC.compare_outputs(ref, ro, gpu, go, check_order=False)
    go2 = ctx.run(toks, offs)
    assert MC.compare_outputs(ref, ro, gpu, go2, check_order=False) == n
    # the order inside a document is the order of the key tokens: deterministic from run to run
    assert (go["results"] == go2["results"]).all() and (go["items"] == go2["items"]).all()


def test_references_and_private_patterns():
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
        ref, gpu = both(define, compile_)
        toks, offs = MC.random_documents(9, 30, 200, 12)
        ctx, go = run_eval(gpu, toks, offs)
        assert MC.compare_outputs(ref, ref.run(toks, offs, nthreads=4), gpu, go, check_order=False) > 0


def test_irregular_documents_offsets_and_segments():
    ref, gpu = both(lambda m: MC.define_c1_rules(m, 41, 100, 20), True)
    toks, offs = MC.random_documents(42, 5, 50, 20)
    toks["ordpos"][int(offs[1]) + 10] = 1
    offs2 = np.array([offs[0], offs[1], offs[2], offs[3], offs[3], offs[4], offs[5]], dtype=np.uint64)
    ctx, go = run_eval(gpu, toks, offs2)
    ro = ref.run(toks, offs2)
    assert ro["status"][1] != 0 and go["status"][1] != 0
    MC.compare_outputs(ref, ro, gpu, go, check_order=False)
    assert ctx.counter("eval_flagged") == 1
    segs = (np.arange(len(toks)) // 17).astype(np.uint32)
    toks["id"][::9] = 0
    ctx, go = run_eval(gpu, toks, offs, origsegs=segs)
    MC.compare_outputs(ref, ref.run(toks, offs, origsegs=segs), gpu, go, check_order=False)
    # documents that do not start at token 0
    sub = offs[1:]
    ctx, go = run_eval(gpu, toks, sub, origsegs=segs)
    MC.compare_outputs(ref, ref.run(toks, sub, origsegs=segs), gpu, go, check_order=False)


def test_rule_sets_outside_of_the_evaluator_are_refused_in_eval_mode():
    from struspattern_b200 import PatternError
    ref, gpu = both(lambda m: MC.define_tree_rules(m, 11, 120, 40, depth=3), True)
    toks, offs = MC.random_documents(511, 4, 100, 40)
    assert gpu.table_info("eval_capable") == 0
    with pytest.raises(PatternError):
        run_eval(gpu, toks, offs)
    go = gpu.run(toks, offs)        # automatic mode: state machine
    MC.compare_outputs(ref, ref.run(toks, offs), gpu, go, check_order=True)


def test_long_ranges_and_many_tokens_per_position():
    def define(m):
        rng = np.random.default_rng(5)
        for r in range(150):
            terms = rng.integers(1, 20, size=int(rng.integers(2, 4)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_WITHIN, terms, int(rng.integers(60, 200)), 0)
            m.define_pattern("L%u" % r, "", True)
    ref, gpu = both(define, False)
    toks, offs = MC.random_documents(6, 12, 500, 20, tokens_per_pos=0.8)
    toks["ordpos"] = toks["ordpos"] * 3 + (toks["ordpos"] // 50) * 100
    ctx, go = run_eval(gpu, toks, offs)
    MC.compare_outputs(ref, ref.run(toks, offs, nthreads=4), gpu, go, check_order=False)


def test_both_count_formulations_give_the_same_output():
    """k_eval_count_join (hash join over the window of a key token) and k_eval_count_step (every rule followed token by
    token) must agree bit for bit, also for long ranges, many tokens per position and windows that overflow the hash table."""
    def define(m):
        MC.define_c1_rules(m, 9, 600, 60, nterms=(2, 4))
        rng = np.random.default_rng(5)
        for r in range(60):
            terms = rng.integers(1, 40, size=int(rng.integers(2, 4)))
            MC.define_term_rule(m, MC.OP_SEQUENCE if r % 2 else MC.OP_WITHIN, terms, int(rng.integers(12, 90)), 0)
            m.define_pattern("L%u" % r, "", True)
    ref, gpu = both(define, True)
    toks, offs = MC.random_documents(77, 60, 400, 60, tokens_per_pos=0.15)     # about 7 tokens per position: big windows
    outs = []
    for key in ("count=join", "count=step"):
        ctx = gpu.create_context()
        ctx.counter("mode=eval")
        ctx.counter(key)
        outs.append(ctx.run(toks, offs))
    for f in ("results", "items", "result_offsets", "item_offsets", "stats", "status"):
        assert (outs[0][f] == outs[1][f]).all(), f
    MC.compare_outputs(ref, ref.run(toks, offs, nthreads=4), gpu, outs[0], check_order=False)


def define_distinct_term_rules(m, seed, nrules, nfeatures):
    rng = np.random.default_rng(seed)
    ops = (MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_IMM)
    for r in range(nrules):
        terms = rng.choice(nfeatures, size=int(rng.integers(2, 4)), replace=False) + 1
        MC.define_term_rule(m, ops[r % 3], terms, int(rng.integers(1, 11)), 0)
        m.define_pattern("D%u" % r, "", True)


def test_crowded_positions_windows_beyond_the_join_table():
    """Six to seven tokens per ordinal position: the windows of the hash join hold more than 64 tokens or more than 32
    distinct ids, so their rules are followed one by one (no aggregate, no pair lookups); mixed with windows that fit."""
    ref, gpu = both(lambda m: define_distinct_term_rules(m, 21, 400, 120), True)
    crowded, offs1 = MC.random_documents(23, 12, 400, 120, delim_prob=0.0, tokens_per_pos=0.15)
    sparse, offs2 = MC.random_documents(24, 12, 400, 120, delim_prob=0.0)
    toks = np.concatenate([crowded, sparse])
    offs = np.concatenate([offs1, offs2[1:] + offs1[-1]])
    ctx, go = run_eval(gpu, toks, offs)
    assert MC.compare_outputs(ref, ref.run(toks, offs, nthreads=4), gpu, go, check_order=False) > 0
    assert ctx.counter("eval_flagged") <= (len(offs) - 1) // 2
