#!/usr/bin/env python3
"""Extract the reference's golden vectors into JSON fixtures (run in the build container, where
/root/reference exists; the fixtures travel to the GPU box, the reference does not).

  charRegexMatch.json          <- tests/charRegexMatch/src/testCharRegexMatch.cpp:105-209 (g_tests)
  simpleTokenPatternMatch.json <- tests/simpleTokenPatternMatch/src/testSimpleTokenPatternMatch.cpp:79-97, 179-218
  worked_example.json          <- doc/webpage/introduction_struspattern.htm:82-102 (rules), :171-176 (input),
                                  :188-226 (tokens with ordinal positions), :232-236 (results with their items)
"""
import json
import os
import re
import sys

REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))


def c_string(lit):
    """Decode one C string literal body (without quotes) to bytes."""
    out = bytearray()
    i = 0
    while i < len(lit):
        c = lit[i]
        if c != "\\":
            out += c.encode()
            i += 1
            continue
        n = lit[i + 1]
        if n == "x":
            j = i + 2
            while j < len(lit) and lit[j] in "0123456789abcdefABCDEF":
                j += 1
            out.append(int(lit[i + 2:j], 16) & 0xFF)
            i = j
        else:
            out += {"n": b"\n", "t": b"\t", "\\": b"\\", '"': b'"', "0": b"\0"}.get(n, n.encode())
            i += 2
    return bytes(out)


def worked_example():
    """The published two-level example: rule file, input document, token listing and result listing, verbatim."""
    import html
    lines = open(os.path.join(REF, "doc", "webpage", "introduction_struspattern.htm"), encoding="utf-8").read().split("\n")
    unesc = lambda ls: html.unescape("\n".join(ls))
    rules = unesc(lines[81:102])                    # :82-102
    xml = unesc(lines[170:176]) + "\n"             # :171-176
    tokens = [ln.strip() for ln in lines[187:226] if re.match(r"\d+: ", ln.strip())]       # :188-226
    results = [html.unescape(ln.strip()) for ln in lines[231:236] if "[" in ln]             # :232-236
    begin = xml.index("<text>") + len("<text>")
    end = xml.index("</text>")
    return {"source": "doc/webpage/introduction_struspattern.htm:82-102, 171-176, 188-226, 232-236",
            "rules": rules, "xml": xml, "text": xml[begin:end], "text_offset": begin, "tokens": tokens, "results": results}


def char_regex_match():
    src = open(os.path.join(REF, "tests/charRegexMatch/src/testCharRegexMatch.cpp"), encoding="latin-1").read()
    body = src[src.index("g_tests[32] ="):src.index("int main(")]
    body = re.sub(r"\},\s*//[^\n]*", "},", body)  # trailing line comments after an entry
    cases = []
    # every test: { {patterns}, {symbols}, "src", {results} }
    for m in re.finditer(r'\{\s*\{((?:\s*\{[^{}]*\},?)+)\s*\},\s*\{((?:\s*\{[^{}]*\},?)+)\s*\},\s*("(?:[^"\\]|\\.)*"|0),\s*\{((?:\s*\{[^{}]*\},?\s*)+)\}\s*\}', body):
        pats, syms, text, res = m.groups()
        if text == "0":
            continue
        case = {"patterns": [], "symbols": [], "results": []}
        for pm in re.finditer(r'\{(\d+),\s*("(?:[^"\\]|\\.)*"),\s*(\d+),\s*(\d+),\s*(true|false)\}', pats):
            case["patterns"].append({"id": int(pm.group(1)), "expression_latin1": c_string(pm.group(2)[1:-1]).decode("latin-1"),
                                     "resultIndex": int(pm.group(3)), "level": int(pm.group(4)), "haspos": pm.group(5) == "true"})
        for sm in re.finditer(r'\{(\d+),\s*(\d+),\s*"([^"]*)"\}', syms):
            case["symbols"].append({"id": int(sm.group(1)), "patternid": int(sm.group(2)), "name": sm.group(3)})
        case["src_latin1"] = c_string(text[1:-1]).decode("latin-1")
        for rm in re.finditer(r'\{(\d+),\s*(\d+),\s*(\d+),\s*(\d+)\}', res):
            r = [int(x) for x in rm.groups()]
            if r[3] != 0:
                case["results"].append(r)
        cases.append(case)
    return {"source": "tests/charRegexMatch/src/testCharRegexMatch.cpp:105-209", "options": ["DOTALL"],
            "note": "strings are the raw bytes of the C literals, stored latin-1 decoded; haspos=false means BindPredecessor (:72-74)",
            "cases": cases}


def simple_token_pattern_match():
    src = open(os.path.join(REF, "tests/simpleTokenPatternMatch/src/testSimpleTokenPatternMatch.cpp")).read()
    body = src[src.index("testPatterns[32] ="):src.index("int main(")]
    pats = []
    for m in re.finditer(r'\{"([^"]+)",\s*\{\{Operation::Term,TOKEN\((\d+)\),(\d+)\},\s*\{Operation::Term,TOKEN\((\d+)\),(\d+)\},\s*'
                         r'\{Operation::Expression,0,0,PT::(\w+),(\d+),(\d+),(\d+)\}\},\s*\{([^}]*)\}', body):
        name, t1, v1, t2, v2, op, rng, card, argc, res = m.groups()
        pats.append({"name": name, "terms": [[int(t1), int(v1)], [int(t2), int(v2)]], "op": op, "range": int(rng),
                     "cardinality": int(card), "argc": int(argc), "results": [int(x) for x in res.split(",") if int(x) != 0]})
    return {"source": "tests/simpleTokenPatternMatch/src/testSimpleTokenPatternMatch.cpp:79-97,179-218",
            "document": "tokens i+1 at ordpos i+1 for i<100, DELIM (1<<24) at ordpos i+2 after every 10th, then TOKEN(1) at 101 and 102; "
                        "origpos = running index, origsize 1; variables are named A<n>; compile() is called",
            "patterns": pats}


if __name__ == "__main__":
    json.dump(worked_example(), open(os.path.join(HERE, "worked_example.json"), "w"), indent=1, ensure_ascii=False)
    json.dump(char_regex_match(), open(os.path.join(HERE, "charRegexMatch.json"), "w"), indent=1)
    json.dump(simple_token_pattern_match(), open(os.path.join(HERE, "simpleTokenPatternMatch.json"), "w"), indent=1)
    print("written")
