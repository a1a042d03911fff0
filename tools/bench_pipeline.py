#!/usr/bin/env python3
"""Full two-level pipeline (BASELINE config 3): ~200 lexem definitions + symbols + token patterns over synthetic UTF-8
text; the lexems stay in HBM between the two levels. Prints one JSON line: GB/s of text lexed+matched (device resident
and end to end from pinned host text to host results), matches/s, and the CPU pipeline (level-1 oracle + the
reference's own automaton, all host threads) on a bounded sample of the same corpus."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
SEED = 0x5354525553000003


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=256 << 20)
    ap.add_argument("--patterns", type=int, default=10000)
    ap.add_argument("--symbols", type=int, default=5000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--cpu-bytes", type=int, default=8 << 20)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--no-compile", action="store_true", help="skip PatternMatcherInstance::compile (no stop word / alternative key optimisation)")
    args = ap.parse_args()
    import torch
    from bench_util import corpus as CG
    from struspattern_b200 import Lexer, Matcher, RESULT_DT, ITEM_DT, STATS_DT
    from struspattern_b200 import workloads as W
    fixed = W.PHRASE_FIRST + W.PHRASE_SECOND + W.ABBREVIATIONS + ["the", "of", "and", "to", "in", "a", "is", "that", "for", "it"]
    vocab = CG.vocabulary(SEED, 50000, fixed)
    symbols = W.symbol_table(vocab, args.symbols)
    offs = CG.layout(SEED + 1, args.bytes)
    keep = torch.empty(args.bytes + 64, dtype=torch.uint8, pin_memory=True)
    text = keep.numpy()[:args.bytes]
    CG.fill(SEED + 1, vocab, text, offs)
    ndocs = len(offs) - 1
    lx = W.define_lexer(Lexer(), nphrases=140, symbols=symbols)
    mt = Matcher()
    W.define_c3_patterns(mt, 33, args.patterns, args.symbols)
    if not args.no_compile:
        mt.compile()
    lctx, mctx = lx.create_context(), mt.create_context()
    dev = torch.empty(args.bytes + 64, dtype=torch.uint8, device="cuda")
    dev[:args.bytes].copy_(keep[:args.bytes])
    for _ in range(2):
        lctx.launch(dev.data_ptr(), offs, on_device=True); lctx.count()
        mctx.launch_lexer_output(lctx); nres, nitems = mctx.count()
    torch.cuda.synchronize()
    mctx.counter("profile:on"); mctx.counter("profile:reset")
    tl = tm = 0.0
    t0 = time.perf_counter()
    for _ in range(args.steps):
        a = time.perf_counter()
        lctx.launch(dev.data_ptr(), offs, on_device=True); nlex = lctx.count()
        b = time.perf_counter()
        mctx.launch_lexer_output(lctx); nres, nitems = mctx.count()
        tm += time.perf_counter() - b
        tl += b - a
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    kern = {k: mctx.counter("time_us:" + k) / 1000.0 / args.steps for k in ("k_eval_count", "k_eval_write", "k_match_documents")}
    mctx.counter("profile:off")
    # end to end: pinned host text in, results + items + statistics out
    res = torch.empty((nres + 16) * RESULT_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(RESULT_DT)
    items = torch.empty((nitems + 16) * ITEM_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(ITEM_DT)
    roffs = torch.empty(ndocs + 1, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
    ioffs = torch.empty(ndocs + 1, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
    stats = torch.empty(ndocs * STATS_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(STATS_DT)
    status = torch.empty(ndocs, dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lctx.launch(text, offs, on_device=False)
        mctx.launch_lexer_output(lctx)
        mctx.fetch_into(res, roffs, items, ioffs, stats, status)
    e2e = (time.perf_counter() - t0) / args.steps
    line = {"workload": "C3 full pipeline: %d lexem definitions, %d symbols, %d token patterns over %d bytes in %d documents"
                        % (lx.table_info("definitions"), args.symbols, args.patterns, args.bytes, ndocs),
            "lexems": int(nlex), "results": int(nres), "items": int(nitems),
            "gbps_device_resident": args.bytes / dt / 1e9, "gbps_e2e": args.bytes / e2e / 1e9,
            "lexer_ms": 1000 * tl / args.steps, "matcher_ms": 1000 * tm / args.steps, "matches_per_s": nres / dt,
            "mtokens_per_s_matcher": nlex / (tm / args.steps) / 1e6,
            "installs": int(stats["nof_programs_installed"].sum()), "signals": int(stats["nof_signals_fired"].sum()), "altkey_installs": int(stats["nof_altkey_programs_installed"].sum()),
            "errors": int((status != 0).sum()), "arena_retries": mctx.counter("arena_retries"), "kernel_ms": kern,
            "eval_capable": mctx.counter("eval_capable"), "eval_flagged_docs": mctx.counter("eval_flagged"), "matcher_mode": mctx.counter("mode"),
            "h2d_bytes": int(args.bytes + 8 * (ndocs + 1)), "d2h_bytes": int(40 * nres + 32 * nitems + 60 * ndocs)}
    if not args.no_cpu:
        from oracle_util import OracleLexer, RefMatcher
        import matcher_cases as MC
        nd = max(1, int(np.searchsorted(offs, args.cpu_bytes, side="right")) - 1)
        nthreads = os.cpu_count() or 1
        olx = W.define_lexer(OracleLexer(), nphrases=140, symbols=symbols)
        ref = RefMatcher()
        W.define_c3_patterns(ref, 33, args.patterns, args.symbols)
        if not args.no_compile:
            ref.compile()
        t0 = time.perf_counter()
        ro = olx.scan(text[:int(offs[nd])], offs[:nd + 1], nthreads=nthreads)
        t1 = time.perf_counter()
        rm = ref.run(ro["lexems"], ro["offsets"], nthreads=nthreads, with_results=args.check)
        t2 = time.perf_counter()
        line["cpu_baseline"] = {"value": float(offs[nd]) / (t2 - t0) / 1e9, "unit": "GB/s", "cores": nthreads, "kind": "port+reference",
                                "sample": "first %d bytes: level-1 CPU restatement %.2f s + reference automaton %.2f s" % (int(offs[nd]), t1 - t0, t2 - t1),
                                "matcher_mtokens_per_s": len(ro["lexems"]) / (t2 - t1) / 1e6}
        if args.check:
            sub = lx.create_context().scan(text[:int(offs[nd])], offs[:nd + 1])
            assert (sub["lexems"] == ro["lexems"]).all()
            gm = mt.run(sub["lexems"], sub["offsets"])
            line["parity_results_checked"] = MC.compare_outputs(ref, rm, mt, gm, check_order=False)
    print(json.dumps(line))


if __name__ == "__main__":
    main()
