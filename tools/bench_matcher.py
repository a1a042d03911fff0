#!/usr/bin/env python3
"""Timing of the level-2 matcher alone (BASELINE config 1 shape: docs x tokens, Zipf features, random 2-term rules):
GPU through the C ABI (device resident tokens) next to the reference's own automaton on all host cores."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=1000)
    ap.add_argument("--tokens", type=int, default=1000)
    ap.add_argument("--rules", type=int, default=1000)
    ap.add_argument("--features", type=int, default=10000)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warps", type=int, default=0)
    ap.add_argument("--no-compile", action="store_true")
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--c4", action="store_true", help="BASELINE config 3 shape: --rules dictionary style sequence_imm patterns (2-5 terms, Zipf 0.8) "
                                                       "over a vocabulary of --features terms; --docs token streams of 1000-4000 tokens (Zipf 1.0)")
    args = ap.parse_args()
    import torch
    import matcher_cases as MC
    from oracle_util import RefMatcher
    from struspattern_b200 import Matcher
    from struspattern_b200 import workloads as W
    if args.c4:
        toks, offs = W.c4_token_documents(4712, args.docs, args.features)

        def define(m):
            W.define_c4_patterns(m, 4711, args.rules, args.features)
    else:
        toks, offs = MC.random_documents(4712, args.docs, args.tokens, args.features)

        def define(m):
            MC.define_c1_rules(m, 4711, args.rules, args.features)
    gpu = Matcher()
    t0 = time.perf_counter()
    define(gpu)
    t_define = time.perf_counter() - t0
    t0 = time.perf_counter()
    if not args.no_compile:
        gpu.compile()
    t_compile = time.perf_counter() - t0
    ctx = gpu.create_context()
    if args.warps:
        ctx.counter("warps_per_sm=%d" % args.warps)
    d_tok = torch.from_numpy(toks.view(np.uint8).copy()).cuda()
    d_off = torch.from_numpy(offs.view(np.int64).copy()).cuda()
    ctx.counter("profile:on")
    for _ in range(2):
        ctx.launch(d_tok.data_ptr(), d_off.data_ptr(), on_device=True, ndocs=args.docs)
        ctx.count()
    ctx.counter("profile:reset")
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ctx.launch(d_tok.data_ptr(), d_off.data_ptr(), on_device=True, ndocs=args.docs)
        nres, nitems = ctx.count()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / args.steps
    kms = ctx.counter("time_us:k_match_documents") / 1000.0 / max(1, ctx.counter("calls:k_match_documents"))
    out = ctx.fetch()
    line = {"workload": "C4 dictionary patterns" if args.c4 else "C1 random rules", "define_s": t_define, "compile_s": t_compile,
            "tokens": int(len(toks)), "docs": args.docs, "rules": args.rules, "results": int(nres), "items": int(nitems),
            "gpu_ms": dt * 1000.0, "kernel_ms": kms, "gpu_mtok_s": len(toks) / dt / 1e6,
            "installs": int(out["stats"]["nof_programs_installed"].sum()), "signals": int(out["stats"]["nof_signals_fired"].sum()),
            "arena_retries": ctx.counter("arena_retries"), "pool_retries": ctx.counter("pool_retries"), "warps_per_sm": ctx.counter("warps_per_sm"),
            "errors": int((out["status"] != 0).sum())}
    if not args.no_ref:
        ref = RefMatcher()
        define(ref)
        if not args.no_compile:
            ref.compile()
        nthreads = os.cpu_count() or 1
        t0 = time.perf_counter()
        ro = ref.run(toks, offs, nthreads=nthreads, with_results=args.check)
        rdt = time.perf_counter() - t0
        line.update({"cpu_ms": rdt * 1000.0, "cpu_threads": nthreads, "cpu_mtok_s": len(toks) / rdt / 1e6, "speedup": rdt / dt,
                     "cpu_results": int(ro["result_offsets"][-1])})
        if args.check:
            MC.compare_outputs(ref, ro, gpu, out, check_order=False)
            line["parity"] = "ok"
    print(json.dumps(line))


if __name__ == "__main__":
    main()
