"""Rule file front end: loads the `.rul` syntax of the strus pattern matcher into a lexer and a
matcher through their interface calls (define_lexem / define_symbol / push_term / push_expression / push_pattern /
attach_variable / define_pattern), so that the same file drives the product (struspattern_b200.Lexer / Matcher) and the
test oracles. The grammar lives in strusAnalyzer / strusUtilities, which are not part of the reference tree; it is
restated from the published examples (reference doc/webpage/introduction_struspattern.htm:82-162):

    NAME [^level] : /regex/ [~editdist] [ [resultindex] ] ( | /regex/ ... )* ;      lexem definitions (any delimiter char)
    [.]Name = expression [ ["format string"] ] ;                                    token patterns ('.' = private)
    expression := operator ( arg {, arg} [ | range [, cardinality]] )  |  term
    arg        := [variable =] expression
    term       := LEXEMNAME ["symbol"]  |  PatternName

Lexem ids are assigned in the order of first appearance (1, 2, ...), symbol ids behind them. Lines starting with '#'
are comments. An expression without "| range" gets range argc-1 if it is a sequence_imm (the article's examples omit it
there), else 0. A format string in square brackets behind the expression is the value of the pattern
(…htm:143-161; csrc/result_format.hpp)."""
import re

from .capi import BIND_CONTENT

OPERATORS = {"sequence": 0, "sequence_imm": 1, "sequence_struct": 2, "within": 3, "within_struct": 4, "any": 5, "and": 6}


class RuleSyntaxError(ValueError):
    pass


class _Scanner:
    TOKEN = re.compile(r"\s*(?:(?P<id>[A-Za-z_][A-Za-z0-9_]*)|(?P<num>[0-9]+)|(?P<str>\"(?:[^\"\\]|\\.)*\")|(?P<op>[\^:=|,;()\[\].~]))")

    def __init__(self, text):
        self.s = "\n".join("" if ln.lstrip().startswith("#") else ln for ln in text.splitlines())
        self.i = 0

    def skip(self):
        while self.i < len(self.s) and self.s[self.i].isspace():
            self.i += 1

    def eof(self):
        self.skip()
        return self.i >= len(self.s)

    def peek(self):
        self.skip()
        m = self.TOKEN.match(self.s, self.i)
        if not m:
            raise RuleSyntaxError("unexpected character %r at offset %d" % (self.s[self.i:self.i + 10], self.i))
        return m.lastgroup, m.group(m.lastgroup), m.end()

    def next(self):
        kind, val, end = self.peek()
        self.i = end
        return kind, val

    def expect(self, what):
        kind, val = self.next()
        if val != what:
            raise RuleSyntaxError("expected %r, found %r at offset %d" % (what, val, self.i))

    def accept(self, what):
        if self.eof():
            return False
        kind, val, end = self.peek()
        if val == what:
            self.i = end
            return True
        return False

    def regex(self):
        """a regular expression between two equal delimiter characters (/.../ or @...@ ...)"""
        self.skip()
        d = self.s[self.i]
        if d.isalnum() or d.isspace():
            raise RuleSyntaxError("regular expression delimiter expected at offset %d" % self.i)
        j = self.i + 1
        while j < len(self.s) and self.s[j] != d:
            j += 2 if (self.s[j] == "\\" and d == "/") else 1   # \/ inside /.../
        if j >= len(self.s):
            raise RuleSyntaxError("unterminated regular expression starting at offset %d" % self.i)
        rx = self.s[self.i + 1:j]
        self.i = j + 1
        return rx


class RuleProgram:
    """Result of load(): names of the lexems, symbols and what was defined."""

    def __init__(self):
        self.lexem_ids = {}        # name -> id
        self.lexem_names = {}      # id -> name
        self.symbol_ids = {}       # (lexem name, value) -> id
        self.patterns = []         # (name, visible, format)
        self.lexem_defs = []       # (id, expression, resultindex, level, editdist)

    def lexem_name(self, lexemid):
        return self.lexem_names.get(int(lexemid), "?")


def load(text, lexer, matcher, compile_=True):
    sc = _Scanner(text)
    prog = RuleProgram()
    statements = []
    # pass 1: split into statements, define the lexems (patterns may refer to lexems defined further down)
    while not sc.eof():
        private = sc.accept(".")
        kind, name = sc.next()
        if kind != "id":
            raise RuleSyntaxError("identifier expected, found %r" % name)
        level = 0
        if sc.accept("^"):
            level = int(sc.next()[1])
        if sc.accept(":"):
            if private:
                raise RuleSyntaxError("lexem %s cannot be private" % name)
            lid = prog.lexem_ids.setdefault(name, len(prog.lexem_ids) + 1)
            prog.lexem_names[lid] = name
            while True:
                rx = sc.regex()
                editdist, residx = 0, 0
                if sc.accept("~"):
                    editdist = int(sc.next()[1])
                if sc.accept("["):
                    residx = int(sc.next()[1])
                    sc.expect("]")
                prog.lexem_defs.append((lid, rx, residx, level, editdist))
                if not sc.accept("|"):
                    break
            sc.expect(";")
        else:
            sc.expect("=")
            start = sc.i
            depth = 0
            while True:     # the expression is parsed in pass 2: find its end
                kind, val = sc.next()
                if val == "(":
                    depth += 1
                elif val == ")":
                    depth -= 1
                elif val == ";" and depth == 0:
                    break
            statements.append((name, not private, sc.s[start:sc.i - 1]))
    for (lid, rx, residx, level, editdist) in prog.lexem_defs:
        lexer.define_lexem(lid, rx + (" ~%d" % editdist if editdist else ""), residx, level, BIND_CONTENT)
    # pass 2: patterns
    next_symbol = [len(prog.lexem_ids) + 1]

    def term(sc2):
        kind, name = sc2.next()
        if kind != "id":
            raise RuleSyntaxError("identifier expected in expression, found %r" % name)
        if name in OPERATORS and sc2.accept("("):
            argc = 0
            rng, card = None, 0
            while True:
                kind, val, end = sc2.peek()
                save = sc2.i
                variable = None
                if kind == "id":
                    sc2.i = end
                    if sc2.accept("="):
                        variable = val
                    else:
                        sc2.i = save
                term(sc2)
                if variable:
                    matcher.attach_variable(variable)
                argc += 1
                if sc2.accept(","):
                    continue
                if sc2.accept("|"):
                    rng = int(sc2.next()[1])
                    if sc2.accept(","):
                        card = int(sc2.next()[1])
                sc2.expect(")")
                break
            if rng is None:
                # no range in the source (the article's sequence_imm examples, …htm:140, 156): the arguments of sequence_imm
                # follow each other at once, so argc-1 positions span them when they are lexems; 0 (one position) otherwise
                rng = argc - 1 if name == "sequence_imm" else 0
            matcher.push_expression(OPERATORS[name], argc, rng, card)
        elif name in prog.lexem_ids:
            if not sc2.eof() and sc2.peek()[0] == "str":
                value = bytes(sc2.next()[1][1:-1], "utf-8").decode("unicode_escape")
                key = (name, value)
                if key not in prog.symbol_ids:
                    prog.symbol_ids[key] = next_symbol[0]
                    lexer.define_symbol(next_symbol[0], prog.lexem_ids[name], value)
                    next_symbol[0] += 1
                matcher.push_term(prog.symbol_ids[key])
            else:
                matcher.push_term(prog.lexem_ids[name])
        else:
            matcher.push_pattern(name)

    for (name, visible, expr) in statements:
        sc2 = _Scanner(expr)
        term(sc2)
        fmt = ""
        if sc2.accept("["):
            kind, val = sc2.next()
            if kind != "str":
                raise RuleSyntaxError("format string expected in pattern %s" % name)
            fmt = bytes(val[1:-1], "utf-8").decode("unicode_escape")
            sc2.expect("]")
        if not sc2.eof():
            raise RuleSyntaxError("unexpected %r behind the expression of pattern %s" % (sc2.s[sc2.i:sc2.i + 10], name))
        matcher.define_pattern(name, fmt, visible)
        prog.patterns.append((name, visible, fmt))
    for (name, value), sid in prog.symbol_ids.items():
        prog.lexem_names[sid] = "%s \"%s\"" % (name, value)
    if compile_:
        lexer.compile()
        matcher.compile()
    return prog


def format_tokens(prog, text, lexems):
    """the token listing of `strusPatternMatch -K` (reference doc/webpage/introduction_struspattern.htm:188-226)"""
    out = []
    for lx in lexems:
        out.append("%d: %s %s" % (int(lx["ordpos"]), prog.lexem_name(lx["id"]),
                                    bytes(text[int(lx["origpos"]):int(lx["origpos"]) + int(lx["origsize"])]).decode("utf-8", "replace")))
    return out


def format_results(text, out, result_name, item_name, base_offset=0, doc=0, resolve=None):
    """the result listing of strusPatternMatch (…htm:232-236): Name [ordpos, origpos]: var [ordpos, origpos, size] 'text' ...
    `out` may be the caller's view with value strings (Matcher.visible_results / RefMatcher.run(values=True)): a result
    with a format string prints its value (span references resolved against `text` by `resolve`, e.g.
    struspattern_b200.capi.resolve_value) instead of items, an item with a value prints the value instead of its span."""
    def value_text(v):
        if resolve is not None:
            v = resolve(v, text)
        return bytes(v).decode("utf-8", "replace")
    lines = []
    ro, io = out["result_offsets"], out["item_offsets"]
    rvals, ivals = out.get("result_values"), out.get("item_values")
    ib = int(io[doc])
    items = out["items"][ib:int(io[doc + 1])]
    for ri in range(int(ro[doc]), int(ro[doc + 1])):
        r = out["results"][ri]
        parts = []
        for k in range(int(r["item_begin"]), int(r["item_begin"]) + int(r["item_count"])):
            it = items[k]
            b, e = int(it["origpos"]), int(it["endpos"])
            shown = value_text(ivals[ib + k]) if ivals and ivals[ib + k] is not None else bytes(text[b:e]).decode("utf-8", "replace")
            parts.append("%s [%d, %d, %d] '%s'" % (item_name(it["variable"]), int(it["ordpos"]), b + base_offset, e - b, shown))
        val = " '%s'" % value_text(rvals[ri]) if rvals and rvals[ri] is not None else ""
        lines.append("%s [%d, %d]:%s %s" % (result_name(r["name"]), int(r["ordpos"]), int(r["origpos"]) + base_offset, val, " ".join(parts)))
    return [ln.rstrip() for ln in lines]
