"""sp_matcher_count_noalloc (include/strus_pattern_b200.h): the count pass of the window evaluator with its scratch arrays in
one static device array instead of cudaMalloc'ed buffers. Called through the C ABI and checked against the existing entry
points (sp_matcher_run + sp_matcher_fetch): results / items per document and the statistics counters, bit exact, on two
calls in a row that share the scratch array (the second batch is smaller, so stale contents of the first would show).
This file is named to be collected last."""
import numpy as np
import pytest

import matcher_cases as MC

pytestmark = pytest.mark.gpu


def device_copy(torch, a):
    return torch.from_numpy(np.ascontiguousarray(a).view(np.uint8).reshape(-1).copy()).cuda()


def count_noalloc(torch, m, toks, offs):
    nd = len(offs) - 1
    d_tok = device_copy(torch, toks) if len(toks) else torch.zeros(16, dtype=torch.uint8, device="cuda")
    d_off = device_copy(torch, np.asarray(offs, dtype=np.uint64))
    # poisoned outputs: whatever the call leaves untouched shows up
    d_res = torch.full((max(nd, 1),), -1, dtype=torch.int32, device="cuda")
    d_items = torch.full((max(nd, 1),), -1, dtype=torch.int32, device="cuda")
    d_flags = torch.full((max(nd, 1),), -1, dtype=torch.int32, device="cuda")
    d_stats = torch.full((max(nd, 1) * 5,), -1, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    m.count_noalloc(d_tok.data_ptr(), d_off.data_ptr(), nd, len(toks), d_res.data_ptr(), d_items.data_ptr(), d_stats.data_ptr(), d_flags.data_ptr())
    torch.cuda.synchronize()
    from struspattern_b200.capi import STATS_DT
    return (d_res.cpu().numpy().view(np.uint32)[:nd], d_items.cpu().numpy().view(np.uint32)[:nd],
            d_stats.cpu().numpy().view(STATS_DT)[:nd], d_flags.cpu().numpy().view(np.uint32)[:nd])


def check_against_run(torch, m, ctx, toks, offs):
    out = ctx.run(toks, offs)
    nres, nitems, stats, flags = count_noalloc(torch, m, toks, offs)
    assert set(np.unique(flags)) <= {0, 1}
    keep = flags == 0
    assert (np.diff(out["result_offsets"]).astype(np.uint32)[keep] == nres[keep]).all()
    assert (np.diff(out["item_offsets"]).astype(np.uint32)[keep] == nitems[keep]).all()
    for name in stats.dtype.names:
        assert (out["stats"][name][keep] == stats[name][keep]).all(), name
    assert (nres[~keep] == 0).all() and (nitems[~keep] == 0).all()
    return int(nres.sum()), int(keep.sum())


def test_count_noalloc_matches_run_on_two_calls_in_a_row():
    import torch
    from struspattern_b200 import Matcher
    from struspattern_b200.capi import PatternError
    from test_matcher_eval_hostsim import define_any_nested
    m = Matcher()
    define_any_nested(m, 2, 150, 30)
    MC.define_c1_rules(m, 3, 400, 60)
    m.compile()
    assert m.table_info("eval_capable") == 1
    m.upload()
    ctx = m.create_context()
    toks1, offs1 = MC.random_documents(11, 40, 150, 60)
    toks2, offs2 = MC.random_documents(12, 10, 100, 60)
    n1, kept1 = check_against_run(torch, m, ctx, toks1, offs1)
    n2, kept2 = check_against_run(torch, m, ctx, toks2, offs2)      # smaller batch on the scratch the first one left
    assert n1 > 0 and n2 > 0 and kept1 > 0 and kept2 > 0
    assert check_against_run(torch, m, ctx, toks1, offs1) == (n1, kept1)
    # a document whose tokens are out of order is flagged (sp_matcher_run hands it to the state machine), the others count
    toks4 = toks2.copy()
    toks4["ordpos"][int(offs2[1]) + 10] = 1
    assert check_against_run(torch, m, ctx, toks4, offs2)[1] == len(offs2) - 2
    # a document without tokens and an empty batch
    offs3 = np.array([0, 0, len(toks2)], dtype=np.uint64)
    check_against_run(torch, m, ctx, toks2, offs3)
    nres, nitems, stats, flags = count_noalloc(torch, m, toks2[:0], np.array([0, 0], dtype=np.uint64))
    assert nres.tolist() == [0] and nitems.tolist() == [0] and flags.tolist() == [0]
    # sizes beyond the stated maximum are refused with a telling error, nothing is launched
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    with pytest.raises(PatternError, match="SP_NOALLOC_MAX_TOKENS"):
        m.count_noalloc(d.data_ptr(), d.data_ptr(), 1, 262144 + 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
    with pytest.raises(PatternError, match="SP_NOALLOC_MAX_DOCS"):
        m.count_noalloc(d.data_ptr(), d.data_ptr(), 65536 + 1, 10, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
    # a batch that may need more of the scratch array than there is (rules queued for the full fold): counted or refused
    toks5, offs5 = MC.random_documents(13, 400, 400, 60)
    try:
        assert check_against_run(torch, m, ctx, toks5, offs5)[0] > n1
    except PatternError as e:
        assert "SP_NOALLOC_QUEUE" in str(e) or "SP_NOALLOC_MAX_RULES" in str(e)
    # and the call works after a refusal
    assert check_against_run(torch, m, ctx, toks2, offs2) == (n2, kept2)


def test_count_noalloc_refuses_what_it_cannot_do():
    import torch
    from struspattern_b200 import Matcher
    from struspattern_b200.capi import PatternError
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    m = Matcher()
    MC.define_c1_rules(m, 5, 50, 20)
    m.compile()
    with pytest.raises(PatternError, match="sp_matcher_upload"):     # tables not on the device: the call does not allocate them
        m.count_noalloc(d.data_ptr(), d.data_ptr(), 1, 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
    t = Matcher()
    MC.define_tree_rules(t, 11, 40, 20, depth=3)      # nested sequences: state machine only
    t.compile()
    t.upload()
    assert t.table_info("eval_capable") == 0
    with pytest.raises(PatternError, match="state machine"):
        t.count_noalloc(d.data_ptr(), d.data_ptr(), 1, 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), d.data_ptr())
