"""Result format strings (src/patternMatcher.cpp:172-181, 253-262, 560-566): value strings of results
and items, and which items a caller still sees, against the reference's own matcher compiled around the oracle's
formatter shim (oracle/shim/strus/lib/pattern_resultformat.hpp, written independently of csrc/result_format.hpp; the real
formatter is strusAnalyzer code outside the reference tree: unpinned against it, pinned on the published example
doc/webpage/introduction_struspattern.htm:143-161).
CPU: the product's host code + device code compiled for the host (tests/hostsim). GPU: the C ABI."""
import numpy as np
import pytest

import matcher_cases as MC
from hostsim_util import HostSimMatcher
from oracle_util import LEXEM_DT, RefMatcher

AMOUNT, CHF, EUR, WORD = 1, 2, 3, 4


def define_money(m):
    """doc/webpage/introduction_struspattern.htm:148-157"""
    m.push_term(CHF)
    m.define_pattern("Currency", "CHF", True)
    m.push_term(EUR)
    m.define_pattern("Currency", "EUR", True)
    m.push_term(AMOUNT)
    m.attach_variable("value")
    m.push_pattern("Currency")
    m.attach_variable("currency")
    m.push_expression(MC.OP_SEQUENCE_IMM, 2, 1, 0)
    m.define_pattern("MoneyAmount", "{currency} {value}", True)


MONEY_TEXT = b"pay 3'645 Eur or 12 sFr. now"
MONEY_TOKENS = [(WORD, 0, 3), (AMOUNT, 4, 5), (EUR, 10, 3), (WORD, 14, 2), (AMOUNT, 17, 2), (CHF, 20, 4), (WORD, 25, 3)]


def money_tokens():
    toks = np.zeros(len(MONEY_TOKENS), dtype=LEXEM_DT)
    for i, (tid, pos, size) in enumerate(MONEY_TOKENS):
        toks[i] = (tid, i + 1, pos, size)
    return toks


def canon(m, vis):
    """per document: sorted records (name, value, span, visible items with names and values)"""
    docs = []
    ro, io = vis["result_offsets"], vis["item_offsets"]
    for d in range(len(ro) - 1):
        recs = []
        ib = int(io[d])
        for ri in range(int(ro[d]), int(ro[d + 1])):
            r = vis["results"][ri]
            items = []
            for k in range(ib + int(r["item_begin"]), ib + int(r["item_begin"]) + int(r["item_count"])):
                it = vis["items"][k]
                items.append((m.item_name(it["variable"]), vis["item_values"][k]) + tuple(int(it[f]) for f in ("ordpos", "ordend", "origseg", "origpos", "endseg", "endpos")))
            recs.append((m.result_name(r["name"]), vis["result_values"][ri]) + tuple(int(r[f]) for f in ("ordpos", "ordend", "origseg", "origpos", "endseg", "endpos")) + (tuple(items),))
        docs.append(sorted(recs, key=repr))
    return docs


def check(ref, prod, toks, offs, in_order):
    ro = ref.run(toks, offs, values=True)
    po = prod.visible_results(prod.run(toks, offs))
    a, b = canon(ref, ro), canon(prod, po)
    assert a == b
    if in_order:    # the state machine keeps the emission order of the reference
        assert ro["result_values"] == po["result_values"] and ro["item_values"] == po["item_values"]
    for name in ro["stats"].dtype.names:
        assert (ro["stats"][name] == po["stats"][name]).all()
    return a


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("mode", [0, 1])
def test_money_amount_example(compile_, mode):
    from struspattern_b200.capi import resolve_value    # noqa: F401  (host code of the library; the GPU variant is below)
    ref, sim = RefMatcher(), HostSimMatcher(mode=mode)
    for m in (ref, sim):
        define_money(m)
        if compile_:
            m.compile()
    toks = money_tokens()
    docs = check(ref, sim, toks, [0, len(toks)], in_order=(mode == 0))
    values = {(r[0], r[1]) for r in docs[0]}
    assert ("MoneyAmount", b"EUR \x010,4,0,9\x02") in values and ("MoneyAmount", b"CHF \x010,17,0,19\x02") in values
    assert ("Currency", b"EUR") in values and ("Currency", b"CHF") in values
    assert all(r[-1] == () for r in docs[0])    # results with a format string have no items


def define_formatted_rules(m, seed, nrules, nfeatures, nested):
    """Random patterns with format strings: literal text, variables that occur once, several times or never in a match,
    separators, references to formatted patterns with a variable on them (their value is made of THEIR items: nested
    records, `nested` only), formatted and plain patterns mixed."""
    rng = np.random.default_rng(seed)
    feat = MC.Zipf(nfeatures, 0.8)
    for i, v in enumerate("abcwx"):     # a private pattern that makes the variables known to every format below
        m.push_term(i + 1)
        m.attach_variable(v)
    m.push_expression(MC.OP_WITHIN, 5, 30, 0)
    m.define_pattern("KnownVars", "", False)
    subs = []
    for s in range(6):
        n = int(rng.integers(1, 4))
        if n == 1:
            m.push_term(int(feat.draw(rng)))
            if nested and s % 2 == 0:
                m.attach_variable("w")
        else:
            for i in range(n):
                m.push_term(int(feat.draw(rng)))
                if nested:
                    m.attach_variable(("w", "x")[i % 2])
            m.push_expression(MC.OP_ANY, n, 1, 0)
        fmt = ("S%d" % s, "<{w}>", "{w|+}/{x}", "")[s % 4] if nested else ("S%d" % s, "const", "", "c\\{%d\\}" % s)[s % 4]
        if not nested and "{w" in fmt:
            fmt = "S"
        m.define_pattern("Sub%d" % s, fmt, s % 3 == 0)
        subs.append("Sub%d" % s)
    formats = ["", "{a}", "[{a}|{b}]", "{a|, } and {b|;}", "lit", "{c}{c}", "x\\{{a}\\}", "{b} {a} {c|-}"]
    for r in range(nrules):
        n = int(rng.integers(2, 4))
        names = ["a", "b", "c"]
        for i in range(n):
            if rng.random() < 0.35:
                m.push_pattern(subs[int(rng.integers(0, len(subs)))])
            else:
                m.push_term(int(feat.draw(rng)))
            if rng.random() < 0.8:
                m.attach_variable(names[int(rng.integers(0, 3))])
        op = (MC.OP_SEQUENCE, MC.OP_WITHIN, MC.OP_SEQUENCE_IMM)[int(rng.integers(0, 3))]
        m.push_expression(op, n, int(rng.integers(2, 7)), 0)
        m.define_pattern("F%d" % (r % 11), formats[int(rng.integers(0, len(formats)))], True)


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [1, 2, 3])
def test_formats_with_nested_items_on_the_state_machine(seed, compile_):
    ref, sim = RefMatcher(), HostSimMatcher(mode=0)
    for m in (ref, sim):
        define_formatted_rules(m, seed, 80, 12, nested=True)
        if compile_:
            m.compile()
    assert sim.info("eval_capable") == 0       # items of formatted sub expressions travel as nested records
    toks, offs = MC.random_documents(seed + 70, 6, 200, 12)
    docs = check(ref, sim, toks, offs, in_order=True)
    vals = [r[1] for d in docs for r in d if r[1] is not None]
    itemvals = [it[1] for d in docs for r in d for it in r[-1] if it[1] is not None]
    assert any(b"\x01" in v for v in vals) and any(b"<" in v for v in itemvals)


@pytest.mark.parametrize("compile_", [False, True])
@pytest.mark.parametrize("seed", [4, 5])
def test_formats_on_the_evaluator(seed, compile_):
    ref, sim = RefMatcher(), HostSimMatcher(mode=1)
    for m in (ref, sim):
        define_formatted_rules(m, seed, 80, 12, nested=False)
        if compile_:
            m.compile()
    assert sim.info("eval_capable") == 1
    toks, offs = MC.random_documents(seed + 80, 6, 200, 12)
    docs = check(ref, sim, toks, offs, in_order=False)
    assert any(r[1] is not None for d in docs for r in d)
    assert any(it[1] is not None for d in docs for r in d for it in r[-1])


def test_format_string_errors():
    for cls in (RefMatcher, HostSimMatcher):
        for fmt in ("{nosuchvariable}", "{a", "a}", "{}"):
            m = cls()
            m.push_term(1)
            m.attach_variable("a")
            m.push_term(2)
            m.push_expression(MC.OP_SEQUENCE, 2, 3, 0)
            with pytest.raises(RuntimeError):
                m.define_pattern("P", fmt, True)


def test_resolve_value_needs_no_device():
    from struspattern_b200.capi import resolve_value
    try:
        got = resolve_value(b"EUR \x010,4,0,9\x02!", MONEY_TEXT)
    except RuntimeError as e:      # library not built on this machine
        pytest.skip(str(e))
    assert got == b"EUR 3'645!"
    assert resolve_value(b"\x011,4,1,9\x02", MONEY_TEXT) == b"\x011,4,1,9\x02"      # another segment: left to the caller


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["statemachine", "auto"])
def test_formats_through_the_c_abi(mode):
    from struspattern_b200 import Matcher
    from struspattern_b200.capi import resolve_value
    ref, gpu = RefMatcher(), Matcher()
    for m in (ref, gpu):
        define_money(m)
        m.compile()
    toks = money_tokens()

    class Run:      # Matcher with a context in the requested mode, the interface check() wants
        def __init__(self, m):
            self.m, self.ctx = m, m.create_context()
            self.ctx.counter("mode=" + mode)
            self.visible_results, self.result_name, self.item_name = m.visible_results, m.result_name, m.item_name

        def run(self, t, o):
            return self.ctx.run(t, o)
    docs = check(ref, Run(gpu), toks, [0, len(toks)], in_order=(mode == "statemachine"))
    resolved = {(r[0], resolve_value(r[1], MONEY_TEXT)) for r in docs[0]}
    assert ("MoneyAmount", b"EUR 3'645") in resolved and ("MoneyAmount", b"CHF 12") in resolved
    for nested, seed in ((True, 1), (False, 4)):
        ref, gpu = RefMatcher(), Matcher()
        for m in (ref, gpu):
            define_formatted_rules(m, seed, 80, 12, nested=nested)
            m.compile()
        assert gpu.table_info("eval_capable") == (0 if nested else 1)
        toks, offs = MC.random_documents(seed + 70, 40, 200, 12)
        docs = check(ref, Run(gpu), toks, offs, in_order=(mode == "statemachine" or nested))
        assert any(r[1] is not None for d in docs for r in d)


MONEY_RULES = r"""
# doc/webpage/introduction_struspattern.htm:148-157
AMOUNT^3        : /\b[0-9]{1,3}'[0-9]{3,3}'[0-9]{3,3}\b/;
AMOUNT^2        : /\b[0-9]{1,3}'[0-9]{3,3}\b/;
AMOUNT^1        : /\b[0-9]{1,3}\b/;
CURRENCY_CHF    : /\b[Ss]{0,1}[Ff][Rr][.]{0,1}\b/;
CURRENCY_EUR    : /\b[Ee][Uu][Rr][.]{0,1}\b/;
Currency        = CURRENCY_CHF ["CHF"];
Currency        = CURRENCY_EUR ["EUR"];
MoneyAmount     = sequence_imm( value=AMOUNT, currency=Currency) ["{currency} {value}"];
"""
MONEY_INPUT = b"pay 3'645 Eur or 12 sFr now, 1'000'000 EUR later"
MONEY_LISTING = ["Currency [2, 10]: 'EUR'", "MoneyAmount [1, 4]: 'EUR 3'645'", "Currency [4, 20]: 'CHF'", "MoneyAmount [3, 17]: 'CHF 12'",
                 "Currency [6, 39]: 'EUR'", "MoneyAmount [5, 29]: 'EUR 1'000'000'"]


def money_listing(lexer, matcher, run):
    from struspattern_b200 import rulefile
    from struspattern_b200.capi import resolve_value
    rulefile.load(MONEY_RULES, lexer, matcher)
    text = np.frombuffer(MONEY_INPUT, dtype=np.uint8)
    lo = lexer.scan(text, np.array([0, len(text)], dtype=np.uint64))
    return rulefile.format_results(text, run(lo), matcher.result_name, matcher.item_name, resolve=resolve_value)


def test_money_rule_file_oracle():
    """the article: an input "3'645 Eur" generates the value "EUR 3'645" """
    from oracle_util import OracleLexer
    mt = RefMatcher()
    try:
        got = money_listing(OracleLexer(), mt, lambda lo: mt.run(lo["lexems"], lo["offsets"], values=True))
    except RuntimeError as e:
        if "not built" in str(e):
            pytest.skip(str(e))
        raise
    assert sorted(got) == sorted(MONEY_LISTING)


@pytest.mark.gpu
def test_money_rule_file_gpu():
    from struspattern_b200 import Lexer, Matcher
    mt = Matcher()
    got = money_listing(Lexer(), mt, lambda lo: mt.visible_results(mt.run(lo["lexems"], lo["offsets"])))
    assert sorted(got) == sorted(MONEY_LISTING)


def random_format(rng):
    parts = []
    for _ in range(int(rng.integers(1, 6))):
        kind = int(rng.integers(0, 6))
        if kind == 0:
            parts.append("{%s}" % "abc"[int(rng.integers(0, 3))])
        elif kind == 1:
            parts.append("{ %s |%s}" % ("abc"[int(rng.integers(0, 3))], ("", ", ", "\\}", "|", " - ")[int(rng.integers(0, 5))]))
        elif kind == 2:
            parts.append(("\\{", "\\}", "\\\\", "|", " ", "x|y")[int(rng.integers(0, 6))])
        elif kind == 3:
            parts.append("{%s|}" % "abc"[int(rng.integers(0, 3))])
        else:
            parts.append("".join("lit-.:/"[int(rng.integers(0, 7))] for _ in range(int(rng.integers(1, 5)))))
    return "".join(parts)


@pytest.mark.parametrize("seed", [1, 2, 3, 4])
def test_random_format_strings_product_and_oracle_formatter_agree(seed):
    """the two formatters (csrc/result_format.hpp, oracle shim) are written independently: random format strings with
    escapes, separators (empty, with quoted braces, with '|') and blanks inside the braces give the same values."""
    rng = np.random.default_rng(seed)
    formats = [random_format(rng) for _ in range(40)]
    ref, sim = RefMatcher(), HostSimMatcher(mode=0)
    for m in (ref, sim):
        for i, v in enumerate("abc"):
            m.push_term(i + 1)
            m.attach_variable(v)
        m.push_expression(MC.OP_WITHIN, 3, 30, 0)
        m.define_pattern("KnownVars", "", False)
        for r, fmt in enumerate(formats):
            terms = [int(t) for t in np.random.default_rng(seed * 100 + r).integers(1, 9, size=3)]
            for i, t in enumerate(terms):
                m.push_term(t)
                m.attach_variable("abc"[(r + i) % 3] if i < 2 else "a")     # one variable twice: separators show
            m.push_expression(MC.OP_WITHIN, 3, 6, 0)
            m.define_pattern("R%d" % r, fmt, True)
    toks, offs = MC.random_documents(seed + 300, 4, 150, 8, delim_prob=0.0)
    docs = check(ref, sim, toks, offs, in_order=True)
    values = [r[1] for d in docs for r in d if r[1] is not None]
    assert len(values) > 20 and len(set(values)) > 10


@pytest.mark.parametrize("nested", [True, False])
def test_formats_with_the_exclusive_option(nested):
    """`exclusive` drops results covered by others (src/patternMatcher.cpp:192-246) before the values are made; the rule set
    runs on the state machine then."""
    ref, sim = RefMatcher(), HostSimMatcher(mode=1)
    for m in (ref, sim):
        m.define_option("exclusive", 0.0)
        m.define_option("maxResultSize", 30.0)
        define_formatted_rules(m, 7, 80, 12, nested=nested)
        m.compile()
    assert sim.info("eval_capable") == 0
    toks, offs = MC.random_documents(91, 6, 200, 12)
    docs = check(ref, sim, toks, offs, in_order=True)
    assert any(r[1] is not None for d in docs for r in d)
