#!/usr/bin/env python3
"""Latency of small batches through the C ABI (what an unmodified strus caller sees per document, SURVEY 8b: one
PatternLexerContext::match + putInput per lexem + fetchResults per document is a batch of one): 1, 16, 256 documents of the
bench corpus through sp_lexer_scan (host text) -> sp_matcher_run_lexer_output -> sp_matcher_fetch, C3 rule set.
Prints one JSON line: milliseconds per call and per document, for both level-2 formulations."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import bench
    import torch
    from struspattern_b200 import Lexer, Matcher
    vocab = bench.vocabulary()
    lx, mt = bench.define_workload(Lexer, Matcher, vocab)
    text, offs, keep = bench.make_corpus(16 << 20, 0, True, vocab)
    out = {"workload": "C3 rule set, documents of the bench corpus (4-64 KiB), host text in, host results out", "rows": []}
    for mode in ("auto", "statemachine"):
        lctx, mctx = lx.create_context(), mt.create_context()
        mctx.counter("mode=" + mode)
        for nd in (1, 16, 256):
            reps = {1: 50, 16: 20, 256: 5}[nd]
            times = []
            for r in range(reps + 2):
                b = (r * nd) % (len(offs) - 1 - nd)
                sub = np.ascontiguousarray(offs[b:b + nd + 1] - offs[b])
                t0 = time.perf_counter()
                lctx.launch(text[int(offs[b]):int(offs[b + nd])], sub, on_device=False)
                mctx.launch_lexer_output(lctx)
                res = mctx.fetch()
                torch.cuda.synchronize()
                if r >= 2:
                    times.append(time.perf_counter() - t0)
            out["rows"].append({"mode": mode, "documents": nd, "ms_per_call": 1000 * float(np.median(times)), "ms_per_document": 1000 * float(np.median(times)) / nd,
                                "bytes": int(offs[nd] - offs[0]), "results_last": int(len(res["results"]))})
    print(json.dumps(out))


if __name__ == "__main__":
    main()
