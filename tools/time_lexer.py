#!/usr/bin/env python3
"""Per-kernel CUDA-event times of the lexer on the C2 corpus (device resident text), without the end-to-end leg of
bench.py: python tools/time_lexer.py [--bytes B] [--steps K]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

KERNELS = ["k_lex_count", "k_lex_verify", "k_lex_emit", "k_lex_finish", "k_lex_resolve", "k_lex_write", "k_scan", "k_isl_flags", "k_isl_resolve",
           "k_ordpos", "k_isl_slots", "k_slot_scan", "k_lex_fused", "k_fused_scan", "k_fused_write", "k_lex_resolve_chunks"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--contexts", type=int, default=1, help="cut the batch into that many sub-batches of whole documents, one host thread + context + stream each")
    args = ap.parse_args()
    import torch
    from struspattern_b200 import Lexer
    from struspattern_b200 import workloads as W
    lx = W.define_lexer(Lexer())
    text, offs, keep = bench.make_corpus(args.bytes, 0, False, bench.vocabulary())
    dev = torch.from_numpy(text).cuda()
    if args.contexts > 1:
        return concurrent(args, lx, dev, offs)
    ctx = lx.create_context()
    for _ in range(2):
        ctx.launch(dev.data_ptr(), offs, on_device=True)
        ctx.count()
    ctx.counter("profile:on"); ctx.counter("profile:reset")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        ctx.launch(dev.data_ptr(), offs, on_device=True)
        ctx.count()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.steps
    k = {n: ctx.counter("time_us:" + n) / 1000.0 / args.steps for n in KERNELS}
    print(json.dumps({"ms_per_step": ms, "gbps": args.bytes / ms / 1e6, "second_half": ctx.counter("second_half"),
                      "lexems": ctx.counter("lexems"), "kernels_ms": {n: v for n, v in k.items() if v}}))


def concurrent(args, lx, dev, offs):
    import threading
    import numpy as np
    import torch
    n = args.contexts
    ndocs = len(offs) - 1
    cuts = [int(np.searchsorted(offs, offs[-1] * k // n, side="left")) for k in range(n)] + [ndocs]
    subs = [(cuts[k], cuts[k + 1]) for k in range(n)]
    streams = [torch.cuda.Stream() for _ in range(n)]
    ctxs = [lx.create_context(s.cuda_stream) for s in streams]
    sub_offs = [np.ascontiguousarray(offs[b:e + 1] - offs[b]) for b, e in subs]
    counts = [0] * n

    def one_pass():
        def worker(k):
            torch.cuda.set_device(0)
            b, e = subs[k]
            ctxs[k].launch(dev.data_ptr() + int(offs[b]), sub_offs[k], on_device=True)
            counts[k] = ctxs[k].count()
        ts = [threading.Thread(target=worker, args=(k,)) for k in range(n)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()

    for _ in range(2):
        one_pass()
    torch.cuda.synchronize()
    import time
    t0 = time.perf_counter()
    for _ in range(args.steps):
        one_pass()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1000.0 / args.steps
    print(json.dumps({"contexts": n, "ms_per_step": ms, "gbps": args.bytes / ms / 1e6, "lexems": sum(counts)}))


if __name__ == "__main__":
    main()
