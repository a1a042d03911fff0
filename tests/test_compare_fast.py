"""The vectorised parity comparison used by bench.py for samples of millions of results (matcher_cases.compare_outputs_fast)
accepts equal outputs in any result order and detects a changed item, two swapped items and a changed result record."""
import numpy as np
import pytest

import matcher_cases as MC
from hostsim_util import HostSimMatcher
from oracle_util import RefMatcher


def _copy(o):
    return {k: (v.copy() if hasattr(v, "copy") else v) for k, v in o.items()}


def test_fast_comparison_detects_differences():
    ref, sim = RefMatcher(), HostSimMatcher(mode=1)
    for m in (ref, sim):
        MC.define_c1_rules(m, 3, 500, 300)
        m.compile()
    toks, offs = MC.random_documents(5, 40, 300, 300)
    ro, so = ref.run(toks, offs), sim.run(toks, offs)
    n = MC.compare_outputs_fast(ref, ro, sim, so, exact_docs=2)
    assert n == len(ro["results"]) > 1000
    bad = _copy(so)
    bad["items"]["ordpos"][len(bad["items"]) - 3] += 1
    with pytest.raises(AssertionError):
        MC.compare_outputs_fast(ref, ro, sim, bad, exact_docs=2)
    bad = _copy(so)
    for ri in range(len(bad["results"]) - 1, 0, -1):
        r = bad["results"][ri]
        if r["item_count"] >= 2:
            d = int(np.searchsorted(bad["result_offsets"], ri, side="right")) - 1
            b = int(bad["item_offsets"][d]) + int(r["item_begin"])
            if tuple(bad["items"][b]) != tuple(bad["items"][b + 1]):
                tmp = bad["items"][b].copy(); bad["items"][b] = bad["items"][b + 1]; bad["items"][b + 1] = tmp
                break
    with pytest.raises(AssertionError):
        MC.compare_outputs_fast(ref, ro, sim, bad, exact_docs=2)
    bad = _copy(so)
    bad["results"]["endpos"][len(bad["results"]) - 1] += 1
    with pytest.raises(AssertionError):
        MC.compare_outputs_fast(ref, ro, sim, bad, exact_docs=2)
    bad = _copy(so)
    bad["stats"]["nof_signals_fired"][7] += 1
    with pytest.raises(AssertionError):
        MC.compare_outputs_fast(ref, ro, sim, bad, exact_docs=2)
