"""The published two-level worked example of the reference (doc/webpage/introduction_struspattern.htm:82-102 rules,
:171-176 input, :188-226 token listing, :232-236 result listing; fixture tests/golden/worked_example.json) as a known-answer
test: second pin of level 1 (UTF-8 classes \\p{Lu} / \\p{Ll}, levels ^1..^7, supersede logic, ordinal positions) and of the
rule file front end, level 2 with nested pattern references and the item order of gatherResultItems. CPU: level-1 oracle
+ the reference's own automaton; GPU: the CUDA path."""
import json
import os

import numpy as np
import pytest

GOLDEN = json.load(open(os.path.join(os.path.dirname(__file__), "golden", "worked_example.json"), encoding="utf-8"))


def run_example(lexer, matcher, order_exact):
    from struspattern_b200 import rulefile
    prog = rulefile.load(GOLDEN["rules"], lexer, matcher)
    text = np.frombuffer(GOLDEN["text"].encode("utf-8"), dtype=np.uint8)
    offs = np.array([0, len(text)], dtype=np.uint64)
    lo = lexer.scan(text, offs)
    assert rulefile.format_tokens(prog, text, lo["lexems"]) == GOLDEN["tokens"]
    mo = matcher.run(lo["lexems"], lo["offsets"])
    got = rulefile.format_results(text, mo, matcher.result_name, matcher.item_name, base_offset=GOLDEN["text_offset"])
    if order_exact:
        assert got == GOLDEN["results"]
    else:
        assert sorted(got) == sorted(GOLDEN["results"])
    return prog


def test_oracle_reproduces_the_published_example():
    from oracle_util import OracleLexer, RefMatcher
    prog = run_example(OracleLexer(), RefMatcher(), True)
    assert prog.lexem_ids == {"WORD": 1, "SENT": 2, "ABBREV": 3, "URL": 4, "EMAIL": 5, "CAPWORD": 6, "LOWORD": 7}


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["statemachine", "auto"])
def test_cuda_path_reproduces_the_published_example(mode, monkeypatch):
    from struspattern_b200 import Lexer, Matcher
    monkeypatch.setenv("SP_MATCHER_MODE", mode)
    run_example(Lexer(), Matcher(), True)   # "Name" refers to "Name": not a rule set of the evaluator, emission order in both modes


def test_rule_file_syntax_errors():
    from oracle_util import OracleLexer, RefMatcher
    from struspattern_b200 import rulefile
    for bad in ("WORD : /a/", "X = sequence( WORD | 2 ;", "WORD ^ : /a/;", "A = [\"x\"];"):
        with pytest.raises((rulefile.RuleSyntaxError, ValueError, RuntimeError)):
            rulefile.load("WORD ^1 : /\\w+/;\n" + bad, OracleLexer(), RefMatcher())
