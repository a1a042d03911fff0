"""Packed wire format of the level-2 output (include/strus_pattern_b200.h: sp_matcher_fetch_packed, sp_unpack_results):
packed on the device, unpacked on the host, equal to what sp_matcher_fetch delivers; runs that do not fit are refused."""
import numpy as np
import pytest

import matcher_cases as MC



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


@pytest.mark.gpu
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


@pytest.mark.gpu
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


def test_unpack_is_host_code_and_restores_the_records():
    """sp_unpack_results needs no device: records packed with numpy by the rule of the header come back field by field."""
    from struspattern_b200.capi import ITEM_DT, ITEM_PACKED_DT, RESULT_DT, RESULT_PACKED_DT, unpack_results
    rng = np.random.default_rng(3)
    ndocs, nres = 5, 40
    counts = rng.integers(0, 4, nres).astype(np.uint32)
    res = np.zeros(nres, dtype=RESULT_DT)
    res["name"] = rng.integers(1, 1 << 18, nres)
    res["format"] = rng.integers(0, 1 << 14, nres)
    res["ordpos"] = rng.integers(1, 1 << 20, nres)
    res["ordend"] = res["ordpos"] + rng.integers(0, 256, nres).astype(np.uint32)
    res["origpos"] = rng.integers(0, 1 << 30, nres)
    res["endpos"] = res["origpos"] + rng.integers(0, 1 << 16, nres).astype(np.uint32)
    res["item_count"] = counts
    ro = np.array([0, 7, 7, 20, 33, 40], dtype=np.uint64)
    nitems = int(counts.sum())
    io = np.array([int(counts[:int(r)].sum()) for r in ro], dtype=np.uint64)
    for d in range(ndocs):
        b, e = int(ro[d]), int(ro[d + 1])
        res["item_begin"][b:e] = np.concatenate(([0], np.cumsum(counts[b:e])[:-1])) if e > b else []
    items = np.zeros(nitems, dtype=ITEM_DT)
    items["variable"] = rng.integers(1, 1 << 12, nitems)
    items["ordpos"] = rng.integers(1, 1 << 20, nitems)
    items["ordend"] = items["ordpos"] + rng.integers(0, 16, nitems).astype(np.uint32)
    items["origpos"] = rng.integers(0, 1 << 30, nitems)
    items["endpos"] = items["origpos"] + rng.integers(0, 1 << 16, nitems).astype(np.uint32)
    pr, pi = np.zeros(nres, dtype=RESULT_PACKED_DT), np.zeros(nitems, dtype=ITEM_PACKED_DT)
    pr["name_format"] = res["name"] | (res["format"] << 18)
    pr["ordpos"], pr["origpos"] = res["ordpos"], res["origpos"]
    pr["lens"] = (res["ordend"] - res["ordpos"]) | (res["item_count"] << 8) | ((res["endpos"] - res["origpos"]) << 16)
    pi["var_lens"] = items["variable"] | ((items["ordend"] - items["ordpos"]) << 12) | ((items["endpos"] - items["origpos"]) << 16)
    pi["ordpos"], pi["origpos"] = items["ordpos"], items["origpos"]
    out_res, out_items = np.zeros(nres, dtype=RESULT_DT), np.zeros(nitems, dtype=ITEM_DT)
    try:
        unpack_results(pr, ro, pi, io, out_res, out_items)
    except RuntimeError as e:       # library not built on this machine
        pytest.skip(str(e))
    assert (out_res == res).all() and (out_items == items).all()
