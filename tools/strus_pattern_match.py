#!/usr/bin/env python3
"""Command line front end in the shape of strusUtilities' strusPatternMatch (reference
doc/webpage/introduction_struspattern.htm:179-238; the original lives in strusUtilities, not in the reference tree):

    strus_pattern_match.py [-K] -p program.rul [--oracle] input [input ...]

Loads the rule file (struspattern_b200/rulefile.py), runs lexer + matcher on every input (an XML document: the content
of its <text> element, offsets relative to the file; or plain text) and prints the recognised entities, with -K also the
tokens. Default: the CUDA path (one batch over all inputs); --oracle: the CPU oracles (test infrastructure)."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("-K", action="store_true", help="print the tokens too")
    ap.add_argument("-p", required=True, help="rule file")
    ap.add_argument("--oracle", action="store_true")
    ap.add_argument("inputs", nargs="+")
    args = ap.parse_args()
    from struspattern_b200 import rulefile
    if args.oracle:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from oracle_util import OracleLexer as Lexer, RefMatcher as Matcher
    else:
        from struspattern_b200 import Lexer, Matcher
    lexer, matcher = Lexer(), Matcher()
    prog = rulefile.load(open(args.p, encoding="utf-8").read(), lexer, matcher)
    docs, bases = [], []
    for path in args.inputs:
        raw = open(path, "rb").read()
        b, e = raw.find(b"<text>"), raw.find(b"</text>")
        if b >= 0 and e > b:
            docs.append(raw[b + 6:e]); bases.append(b + 6)
        else:
            docs.append(raw); bases.append(0)
    text = np.frombuffer(b"".join(docs), dtype=np.uint8)
    offs = np.cumsum([0] + [len(d) for d in docs]).astype(np.uint64)
    lo = lexer.scan(text, offs)
    if args.oracle:
        mo, resolve = matcher.run(lo["lexems"], lo["offsets"], values=True), None
        from struspattern_b200.capi import resolve_value as resolve
    else:
        from struspattern_b200.capi import resolve_value as resolve
        mo = matcher.visible_results(matcher.run(lo["lexems"], lo["offsets"]))
    for d, path in enumerate(args.inputs):
        print("%s:" % path)
        doc = np.frombuffer(docs[d], dtype=np.uint8)
        if args.K:
            for line in rulefile.format_tokens(prog, doc, lo["lexems"][int(lo["offsets"][d]):int(lo["offsets"][d + 1])]):
                print(line)
        for line in rulefile.format_results(doc, mo, matcher.result_name, matcher.item_name, base_offset=bases[d], doc=d, resolve=resolve):
            print(line)
    print("OK done")


if __name__ == "__main__":
    main()
