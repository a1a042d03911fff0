"""The level-1 oracle (oracle/lexer_oracle.cpp) against the reference's golden vectors and against an
independent cross-check (python `re`) of the restated engine contract."""
import json
import os
import re

import numpy as np
import pytest

from oracle_util import OracleLexer

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")
BIND_CONTENT, BIND_SUCCESSOR, BIND_PREDECESSOR, BIND_UNIQUE = 0, 1, 2, 3


def _golden_cases():
    return json.load(open(os.path.join(GOLDEN, "charRegexMatch.json")))["cases"]


def build_case(lexer_cls, case):
    lx = lexer_cls()
    lx.define_option("DOTALL")  # tests/charRegexMatch/src/testCharRegexMatch.cpp:236
    for p in case["patterns"]:
        lx.define_lexem(p["id"], p["expression_latin1"].encode("latin-1"), p["resultIndex"], p["level"],
                        BIND_CONTENT if p["haspos"] else BIND_PREDECESSOR)
    for s in case["symbols"]:
        lx.define_symbol(s["id"], s["patternid"], s["name"])
    return lx


def test_golden_case1_exact():
    """36 lexems, element by element in order (testCharRegexMatch.cpp:107-160, compare loop :247-261)."""
    case = _golden_cases()[0]
    lx = build_case(OracleLexer, case)
    lx.compile()
    got = lx.match(case["src_latin1"].encode("latin-1"))
    assert got == [tuple(r) for r in case["results"]]


@pytest.mark.parametrize("idx", [1, 2])
def test_golden_edit_distance_cases_are_declared_unsupported(idx):
    """Cases 2/3 use 'expr ~1' (Hyperscan edit distance + libtre approximate re-match): a 'next' row.
    The oracle must say so instead of returning something."""
    lx = build_case(OracleLexer, _golden_cases()[idx])
    with pytest.raises(RuntimeError, match="unsupported"):
        lx.compile()


def test_edit_distance_suffix_is_stripped_like_the_reference():
    # src/patternLexer.cpp:605-626: 'abc ~1' -> expression 'abc', distance 1; 'abc~' / 'abc ~x' stay literal
    lx = OracleLexer()
    lx.define_lexem(1, "abc ~0")  # distance 0 == exact
    lx.compile()
    assert lx.match("xabcx") == [(1, 1, 1, 3)]


def _brute_events(patterns, text, flags=0):
    """Independent statement of the engine contract with python's re on code points:
    event (idx, from, to) iff text[from:to] matches in context; from = leftmost per (idx, to)."""
    n = len(text)
    ev = []
    for idx, pat in enumerate(patterns, 1):
        rx = re.compile("(?:%s)" % pat, flags | re.ASCII)
        best = {}
        for s in range(n):
            for e in range(s + 1, n + 1):
                # match text[s:e] exactly, but let assertions see the real neighbours
                m = rx.match(text, s)
                if m is None:
                    break
                mm = re.compile("(?:%s)(?=.{%d}\\Z)" % (pat, n - e), flags | re.ASCII | re.DOTALL).match(text, s)
                if mm is not None and mm.end() == e and e not in best:
                    best[e] = s
        ev += [(idx, s, e) for e, s in best.items()]
    ev.sort(key=lambda t: (t[2], t[0]))
    return ev


@pytest.mark.parametrize("patterns,text", [
    ([r"[0-9]+\b", r"[a-zA-Z0-9]+\b", r"\w+\s\w+\b"], "ab 12 c3d  x9 77"),
    ([r"\bfoo\b", r"fo+", r"o\b"], "foo fooo xfoo foo"),
    ([r"a|ab|abc", r"b+c?"], "abcabbbc ab"),
    ([r"[^a ]+", r"x.y", r"\Bb"], "abxqy x\ny bb"),
    ([r"(?:ab)+", r"a{2,3}", r"(a|b){2}"], "ababab aaaa"),
    ([r"\d+\.\d+", r"\d+", r"[.,]"], "3.14, 42. 0.5"),
])
def test_raw_events_against_python_re(patterns, text):
    lx = OracleLexer()
    for i, p in enumerate(patterns, 1):
        lx.define_lexem(i, p, 0, 0, BIND_CONTENT)
    lx.compile()
    out = lx.scan(text.encode(), [0, len(text.encode())], keep_raw=True)
    got = [tuple(int(x) for x in r) for r in out["raw"]]
    assert got == _brute_events(patterns, text)


def test_utf8_offsets_are_bytes_and_classes_consume_code_points():
    lx = OracleLexer()
    lx.define_lexem(1, "[^ ]+", 0, 0, BIND_CONTENT)  # negated class: whole code points
    lx.define_lexem(2, "\\w+", 0, 1, BIND_CONTENT)   # ASCII \w without UCP
    lx.define_lexem(3, "ö.", 0, 2, BIND_CONTENT)
    lx.compile()
    text = "aö€ bc".encode()  # a(1) ö(2) €(3) space b c
    got = lx.match(text)
    # level 2 'ö€' (1..6) covers nothing fully: [^ ]+ spans 0..6 level 0, \w+ 'a' 0..1 level 1
    assert (3, 2, 1, 5) in got or (3, 1, 1, 5) in got
    assert all(o + s <= len(text) for _, _, o, s in got)
    # no lexem starts or ends inside a multi byte sequence
    bounds = {0, 1, 3, 6, 7, 8, 9}
    assert all(o in bounds and o + s in bounds for _, _, o, s in got)


def test_ucp_word_class():
    lx = OracleLexer()
    lx.define_option("UCP")
    lx.define_lexem(1, "\\b\\w+\\b", 0, 0, BIND_CONTENT)
    lx.compile()
    text = "größe naïve".encode()
    assert lx.match(text) == [(1, 1, 0, 7), (1, 2, 8, 6)]


def test_symbol_event_follows_its_lexem_and_is_not_replaced():
    # [a-z]+ with symbol "the" on "then": (lexem 0..3) is replaced by (lexem 0..4), the symbol stays
    lx = OracleLexer()
    lx.define_lexem(1, "[a-z]+", 0, 0, BIND_CONTENT)
    lx.define_symbol(7, 1, "the")
    lx.compile()
    assert lx.match("then") == [(7, 1, 0, 3), (1, 1, 0, 4)]


def test_double_insert_quirk_is_restated():
    """src/patternLexer.cpp:807-821: inserting lexem+symbol in front of >= 2 later-starting elements loses
    one of them and leaves a zero-initialised element (probe in SURVEY.md section 10)."""
    lx = OracleLexer()
    lx.define_lexem(1, "[a-z]\\b", 0, 0, BIND_CONTENT)      # single letters a, b, c
    lx.define_lexem(2, "a b c\\b", 0, 0, BIND_CONTENT)      # arrives last, starts first, has a symbol
    lx.define_symbol(9, 2, "a b c")
    lx.compile()
    got = lx.match("a b c")
    # array before the last insert: [a@0, b@2, c@4]; two elements start later than the new one:
    # c moves by two places, b by one onto ... and is overwritten by the symbol; a zero element remains
    assert got == [(1, 1, 0, 1), (2, 1, 0, 5), (9, 1, 0, 5), (0, 1, 0, 0), (1, 2, 4, 1)]


def test_position_bind_rules():
    # src/patternLexer.cpp:893-945
    lx = OracleLexer()
    lx.define_lexem(1, "[a-z]+", 0, 0, BIND_CONTENT)
    lx.define_lexem(2, "[.]", 0, 0, BIND_PREDECESSOR)
    lx.define_lexem(3, "#", 0, 0, BIND_SUCCESSOR)
    lx.define_lexem(4, "!", 0, 0, BIND_UNIQUE)
    lx.compile()
    assert lx.match(". #ab. #cd !! ef") == [
        (3, 1, 2, 1),            # leading successor: ordpos 1; leading predecessor '.' dropped
        (1, 1, 3, 2), (2, 1, 5, 1), (3, 2, 7, 1), (1, 2, 8, 2),
        (4, 3, 11, 1),           # first '!' unique: own position; the second one directly after is dropped
        (1, 4, 14, 2)]
    assert lx.match(". . #") == []  # no content element at all -> empty result (:919-922)


def test_token_size_limit_is_an_error():
    lx = OracleLexer()
    lx.define_lexem(1, "\\ba+\\b", 0, 0, BIND_CONTENT)  # \\b in front: one start only, keeps the oracle linear
    lx.compile()
    text = b"a" * 65535
    out = lx.scan(text, [0, len(text)])
    assert out["status"][0] != 0 and len(out["lexems"]) == 0
    text = b"a" * 65534
    out = lx.scan(text, [0, len(text)])
    assert out["status"][0] == 0 and [tuple(int(x) for x in r) for r in out["lexems"]] == [(1, 1, 0, 65534)]


def test_compile_errors():
    for bad in ["a(", "[a-", "a**", "(?=x)", "\\1", "x*"]:
        lx = OracleLexer()
        lx.define_lexem(1, bad)
        with pytest.raises(RuntimeError):
            lx.compile()
