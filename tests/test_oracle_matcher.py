"""The level-2 oracle = the reference's own sources (oracle/_ref/libsp_ref_matcher.so) against the
reference's known-answer test."""
import json
import os

import numpy as np

from oracle_util import LEXEM_DT, RefMatcher, canonical_results

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
OPS = {"OpSequence": 0, "OpSequenceImm": 1, "OpSequenceStruct": 2, "OpWithin": 3, "OpWithinStruct": 4, "OpAny": 5, "OpAnd": 6}
DELIM = 1 << 24


def simple_document():
    # tests/simpleTokenPatternMatch/src/testSimpleTokenPatternMatch.cpp:79-97
    toks = []
    for ii in range(100):
        toks.append((ii + 1, ii + 1))
        if (ii + 1) % 10 == 0:
            toks.append((DELIM, ii + 2))
    toks += [(1, 101), (1, 102)]
    arr = np.zeros(len(toks), dtype=LEXEM_DT)
    for i, (t, p) in enumerate(toks):
        arr[i] = (t, p, i, 1)
    return arr


def define_simple_patterns(m, patterns):
    for p in patterns:
        for termid, var in p["terms"]:
            m.push_term(termid)
            if var:
                m.attach_variable("A%u" % var)
        m.push_expression(OPS[p["op"]], p["argc"], p["range"], p["cardinality"])
        m.define_pattern(p["name"], "", not p["name"].startswith("_"))


def test_simple_token_pattern_match_kat():
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    m = RefMatcher()
    define_simple_patterns(m, kat["patterns"])
    m.compile()
    doc = simple_document()
    out = m.run(doc, [0, len(doc)])
    got = {(m.name(r["name"]), int(r["ordpos"])) for r in out["results"]}
    exp = {(p["name"], pos) for p in kat["patterns"] for pos in p["results"]}
    assert got == exp  # compared as a set, like the reference (:261-301)
    # items come out last-bound first
    recs = canonical_results(out, m.name, m.name)[0]
    assert all([i[0] for i in r[7]] == ["A2", "A1"] for r in recs)


def test_threads_share_one_instance():
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    m = RefMatcher()
    define_simple_patterns(m, kat["patterns"])
    m.compile()
    doc = simple_document()
    toks = np.concatenate([doc] * 8)
    offs = np.arange(9, dtype=np.uint64) * len(doc)
    one = m.run(toks, offs, nthreads=1)
    four = m.run(toks, offs, nthreads=4)
    assert canonical_results(one, m.name, m.name) == canonical_results(four, m.name, m.name)
    assert (one["stats"] == four["stats"]).all()
    assert int(one["stats"]["nof_programs_installed"][0]) == 18
