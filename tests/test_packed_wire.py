"""Packed wire format of the level-2 output (include/strus_pattern_b200.h: sp_matcher_fetch_packed, sp_unpack_results):
packed on the device, unpacked on the host, equal to what sp_matcher_fetch delivers; runs that do not fit are refused."""
import numpy as np
import pytest

import matcher_cases as MC

pytestmark = pytest.mark.gpu


def fetch_both(ctx):
    from struspattern_b200.capi import ITEM_DT, ITEM_PACKED_DT, RESULT_DT, RESULT_PACKED_DT, STATS_DT, unpack_results
    out = ctx.fetch()
    nres, nitems, nd = len(out["results"]), len(out["items"]), ctx.ndocs
    pr, pi = np.zeros(nres + 1, dtype=RESULT_PACKED_DT), np.zeros(nitems + 1, dtype=ITEM_PACKED_DT)
    ro, io = np.zeros(nd + 1, dtype=np.uint64), np.zeros(nd + 1, dtype=np.uint64)
    stats, status = np.zeros(nd, dtype=STATS_DT), np.zeros(nd, dtype=np.uint32)
    ctx.fetch_packed_into(pr, ro, pi, io, stats, status)
    res, items = np.zeros(nres + 1, dtype=RESULT_DT), np.zeros(nitems + 1, dtype=ITEM_DT)
    unpack_results(pr, ro, pi, io, res, items)
    return out, {"results": res[:nres], "items": items[:nitems], "result_offsets": ro, "item_offsets": io, "stats": stats, "status": status}


@pytest.mark.parametrize("mode", ["statemachine", "auto"])
def test_packed_equals_records(mode):
    from struspattern_b200 import Matcher
    from test_matcher_eval_hostsim import define_any_nested
    m = Matcher()
    define_any_nested(m, 2, 150, 30)
    MC.define_c1_rules(m, 3, 400, 60)
    m.compile()
    ctx = m.create_context()
    ctx.counter("mode=" + mode)
    toks, offs = MC.random_documents(11, 60, 200, 60)
    ctx.launch(toks, offs)
    a, b = fetch_both(ctx)
    assert len(a["results"]) > 0 and len(a["items"]) > 0
    for key in ("results", "items", "result_offsets", "item_offsets", "stats", "status"):
        assert (a[key] == b[key]).all(), key
    # an empty batch and a document without results
    ctx.launch(toks[:0], np.array([0, 0], dtype=np.uint64))
    a, b = fetch_both(ctx)
    assert len(b["results"]) == 0 and b["result_offsets"].tolist() == [0, 0]


def test_runs_that_do_not_fit_are_refused():
    from struspattern_b200 import Matcher
    from struspattern_b200.capi import PatternError
    m = Matcher()
    MC.define_c1_rules(m, 3, 200, 30)
    m.compile()
    ctx = m.create_context()
    toks, offs = MC.random_documents(12, 10, 100, 30)
    segs = np.ones(len(toks), dtype=np.uint32)       # segment 1: the packed records have no room for segments
    ctx.launch(toks, offs, origsegs=segs)
    with pytest.raises(PatternError) as err:
        fetch_both(ctx)
    assert err.value.code == -6
    assert len(ctx.fetch()["results"]) > 0             # the records are still there
