#!/usr/bin/env python3
"""Benchmark of the strusPattern hot path on B200 (BASELINE.json metric: GB/s of UTF-8 text lexed+matched).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--bytes B]

A step = one pass of the two-level hot path over one batch of synthetic text per GPU: workload C3 of SURVEY.md 8d
(186 lexem definitions + 5000 symbols, 10 000 token patterns: sequence_imm / within / *_struct / any nested in sequence)
over B bytes (default 256 MiB) of the multilingual synthetic corpus in documents of 4-64 KiB. Level 1 (sp_lexer_scan)
-> level 2 on the lexems still in HBM (sp_matcher_run_lexer_output) -> result and item counts. Weak scaling: every
rank owns its own shard generated from (seed, rank); there is no collective on the data path.
`value` is measured with the text resident in HBM (CUDA events), `e2e` through the C ABI with pinned host buffers: H2D
of the text and D2H of results, items, offsets, statistics and status inside the timed region.
After the timed region a sample of the timed output is compared with the CPU oracle (level-1 restatement + the
reference's own automaton): lexems element by element, results with their items as multisets per document, the cascade
counters exactly. A mismatch fails the run; the number of compared results is in the line (`parity_checked`).
--impl reference times the CPU path (level-1 oracle + oracle/_ref/libsp_ref_matcher.so, all host threads) on a bounded
sample of the same workload. Prints ONE JSON line on rank 0."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "GB/s UTF-8 text lexed+matched at 1/2/4/8 B200 vs same-box hyperscan CPU path"
SEED = 0x5354525553000003
NPATTERNS, NSYMBOLS, NPHRASES, PATTERN_SEED = 10000, 5000, 140, 33
# DRAM traffic of the dominant kernel (dram__bytes_read.sum + dram__bytes_write.sum of one ncu --set full launch), per token
NCU_TRAFFIC = {"k_eval_count": {"bytes_per_token": (490.784768e6 + 179.286e6) / 11344148, "source": "profiles/r02_eval_v11_ncu_raw.csv (k_eval_count_join<8>, 32 MiB of this corpus)"}}


def fixed_words():
    from struspattern_b200 import workloads as W
    return W.PHRASE_FIRST + W.PHRASE_SECOND + W.ABBREVIATIONS + ["the", "of", "and", "to", "in", "a", "is", "that", "for", "it"]


def vocabulary():
    from bench_util import corpus
    return corpus.vocabulary(SEED, 50000, fixed_words())


def make_corpus(nbytes, rank, pinned, vocab):
    """-> (host uint8 array, offsets, keepalive)"""
    from bench_util import corpus
    offs = corpus.layout(SEED + 1 + 1000 * rank, nbytes)
    keep = None
    if pinned:
        import torch
        keep = torch.empty(int(nbytes) + 64, dtype=torch.uint8, pin_memory=True)
        text = keep.numpy()[:nbytes]
    else:
        text = np.zeros(int(nbytes), dtype=np.uint8)
    corpus.fill(SEED + 1 + 1000 * rank, vocab, text, offs)
    return text, offs, keep


def define_workload(lexer_cls, matcher_cls, vocab):
    from struspattern_b200 import workloads as W
    lx = W.define_lexer(lexer_cls(), nphrases=NPHRASES, symbols=W.symbol_table(vocab, NSYMBOLS))
    mt = matcher_cls()
    W.define_c3_patterns(mt, PATTERN_SEED, NPATTERNS, NSYMBOLS)
    mt.compile()
    return lx, mt


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = "index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown," \
        "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
        self.f.flush()
        rows = [r.split(",") for r in open(self.f.name).read().strip().splitlines() if r.count(",") >= 8]
        os.unlink(self.f.name)
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if v.strip().lower().startswith("active"):
                    reasons.add(name)
        if sm:
            out = {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}
        return out


def bind_to_gpu_numa_node(local_rank):
    """Run this rank's host threads (and with them its pinned buffers: pages are placed where they are first touched) on the
    CPU socket its GPU hangs on; without it the copies of half of the ranks cross the socket link. -> what was done"""
    import torch
    try:
        p = torch.cuda.get_device_properties(local_rank)
        dev = "%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % dev).read())
        if node < 0:
            return {"gpu": dev, "node": node, "bound": False}
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            lo, _, hi = part.partition("-")
            cpus.update(range(int(lo), int(hi or lo) + 1))
        usable = cpus & os.sched_getaffinity(0)
        if usable:
            os.sched_setaffinity(0, usable)
        return {"gpu": dev, "node": node, "bound": bool(usable), "cpus": len(usable)}
    except (OSError, ValueError, AttributeError) as exc:
        return {"bound": False, "why": str(exc)[:80]}


def measured_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class CpuPipeline:
    """The CPU arm: level-1 oracle (CPU restatement; Hyperscan/libtre cannot be installed here) + the reference's own
    automaton compiled from its sources (oracle/_ref/libsp_ref_matcher.so), both on all host threads."""

    def __init__(self, vocab):
        from oracle_util import OracleLexer, RefMatcher
        self.lx, self.mt = define_workload(OracleLexer, RefMatcher, vocab)
        self.nthreads = os.cpu_count() or 1

    def run(self, text, offs, nd, with_results):
        t0 = time.perf_counter()
        ro = self.lx.scan(text[:int(offs[nd])], offs[:nd + 1], nthreads=self.nthreads)
        t1 = time.perf_counter()
        rm = self.mt.run(ro["lexems"], ro["offsets"], nthreads=self.nthreads, with_results=with_results)
        t2 = time.perf_counter()
        return ro, rm, t1 - t0, t2 - t1

    def sample_docs(self, text, offs, seconds):
        """number of leading documents that take about `seconds` (probe on ~1 MiB first)"""
        nd = max(1, min(int(np.searchsorted(offs, 1 << 20, side="right")) - 1, len(offs) - 1))
        _, _, tl, tm = self.run(text, offs, nd, False)
        want = float(offs[nd]) / max(tl + tm, 1e-9) * seconds
        return max(nd, min(int(np.searchsorted(offs, want, side="right")) - 1, len(offs) - 1))


def run_reference(args, rank, world):
    if rank != 0:
        return None
    vocab = vocabulary()
    cpu = CpuPipeline(vocab)
    text, offs, _ = make_corpus(min(args.bytes, 64 << 20), 0, False, vocab)
    per_step = max(2.0, 90.0 / max(1, args.steps + args.warmup))
    nd = cpu.sample_docs(text, offs, per_step)
    times, ntok = [], 0
    for i in range(args.warmup + args.steps):
        ro, rm, tl, tm = cpu.run(text, offs, nd, False)
        ntok = len(ro["lexems"])
        if i >= args.warmup:
            times.append((tl, tm))
    total = sum(a + b for a, b in times)
    sample = int(offs[nd])
    value = sample * len(times) / total / 1e9
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * total / len(times), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C3 lexer+matcher: %d lexem definitions + %d symbols, %d token patterns; CPU sample of the corpus" % (
            cpu.lx.table_info("definitions") if hasattr(cpu.lx, "table_info") else 186, NSYMBOLS, NPATTERNS), "sample_bytes": sample},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": cpu.nthreads, "kind": "port+reference",
                         "sample": "first %d bytes (%d documents, %d tokens) of the C3 corpus per step: level-1 CPU restatement %.2f s + the reference's automaton %.2f s "
                                   "(Hyperscan/libtre are not installable here)" % (sample, nd, ntok, times[-1][0], times[-1][1]),
                         "matcher_mtokens_per_s": ntok / times[-1][1] / 1e6},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


def parity_check(cpu, text, offs, nd, lexer_out, matcher_out, gpu_matcher):
    """GPU output of the first nd documents against the CPU oracle -> number of results compared (raises on a mismatch)."""
    import matcher_cases as MC
    ro, rm, tl, tm = cpu.run(text, offs, nd, True)
    assert (ro["offsets"] == lexer_out["offsets"]).all() and (ro["lexems"] == lexer_out["lexems"]).all(), "lexer output differs from the oracle"
    n = MC.compare_outputs_fast(cpu.mt, rm, gpu_matcher, matcher_out)
    return n, len(ro["lexems"]), tl, tm


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from struspattern_b200 import ITEM_DT, RESULT_DT, STATS_DT, Lexer, Matcher

    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else {"bound": False, "why": "single rank"}
    vocab = vocabulary()
    lx, mt = define_workload(Lexer, Matcher, vocab)
    text, offs, keep = make_corpus(args.bytes, rank, True, vocab)
    ndocs = len(offs) - 1
    dev_text = torch.empty(len(text) + 64, dtype=torch.uint8, device="cuda")
    dev_text[:len(text)].copy_(keep[:len(text)], non_blocking=False)
    lctx, mctx = lx.create_context(), mt.create_context()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        lctx.launch(dev_text.data_ptr(), offs, on_device=True)
        nlex = lctx.count()
        mctx.launch_lexer_output(lctx)
        nres, nitems = mctx.count()
        return nlex, nres, nitems

    # ---- device resident
    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        nlex, nres, nitems = step()
    mctx.counter("profile:on"); mctx.counter("profile:reset"); mctx.counter("launches:reset")
    lctx.counter("profile:on"); lctx.counter("profile:reset")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    launches, t_lex = 0, 0.0
    for _ in range(args.steps):
        nlex, nres, nitems = step()
        launches += lctx.counter("kernel_launches")
    e1.record()
    barrier()
    launches += mctx.counter("kernel_launches")
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    kern_ms = {k: mctx.counter("time_us:" + k) / 1000.0 / args.steps for k in ("k_eval_count", "k_eval_write", "k_match_documents")}
    lex_ms = sum(lctx.counter("time_us:" + k) for k in ("k_lex_count", "k_lex_verify", "k_lex_emit", "k_lex_finish", "k_lex_resolve_chunks", "k_fused_scan", "k_fused_write")) / 1000.0 / args.steps
    mctx.counter("profile:off"); lctx.counter("profile:off")
    nrules = mctx.counter("eval_rules") // max(1, mctx.counter("eval_runs"))
    info = {"eval_capable": mctx.counter("eval_capable"), "eval_flagged_docs": mctx.counter("eval_flagged") // max(1, mctx.counter("eval_runs")),
            "rule_instances_per_step": nrules, "rules_by_full_fold_per_step": mctx.counter("eval_queued") // max(1, mctx.counter("eval_runs"))}
    stats = mctx.fetch()["stats"] if args.bytes <= (64 << 20) else None

    # ---- parity gate on the timed output: the first documents of the batch against the CPU oracle
    parity = None
    cpu_line = None
    parity_mo = None
    if rank == 0 and not args.no_parity:
        cpu = CpuPipeline(vocab)
        seconds = args.cpu_seconds if (world == 1 and not args.no_cpu) else 2.0
        nd = cpu.sample_docs(text, offs, seconds)
        # the first nd documents of the output of the last timed step, as it lies in HBM
        lo, mo = lctx.fetch_docs(0, nd), mctx.fetch_docs(0, nd)
        n, ntok, tl, tm = parity_check(cpu, text, offs, nd, lo, mo, mt)
        parity = {"results": n, "documents": nd, "tokens": ntok}
        parity_mo = mo
        if world == 1 and not args.no_cpu:
            cpu_line = {"value": float(offs[nd]) / (tl + tm) / 1e9, "unit": "GB/s", "cores": cpu.nthreads, "kind": "port+reference",
                        "sample": "first %d bytes (%d documents) of the same corpus: level-1 CPU restatement %.2f s + the reference's automaton %.2f s" % (int(offs[nd]), nd, tl, tm),
                        "matcher_mtokens_per_s": ntok / tm / 1e6}

    # ---- end to end through the C ABI with pinned host buffers: sub-batches of whole documents go through (lexer context,
    # matcher context, stream) pairs on several host threads, so that the copies of one sub-batch overlap the kernels of another
    del lctx, mctx
    e2e_steps = max(1, min(args.steps, 2))
    cuts = [0]
    while cuts[-1] < ndocs:
        nxt = int(np.searchsorted(offs, offs[cuts[-1]] + np.uint64(args.e2e_sub_bytes), side="right")) - 1
        cuts.append(min(ndocs, max(nxt, cuts[-1] + 1)))
    subs = [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)]
    sub_offs = [np.ascontiguousarray(offs[b:e + 1] - offs[b]) for b, e in subs]
    res_begin, item_begin = [0] * (len(subs) + 1), [0] * (len(subs) + 1)
    streams = [torch.cuda.Stream() for _ in range(args.e2e_threads)]
    pairs = [(lx.create_context(s.cuda_stream), mt.create_context(s.cuda_stream)) for s in streams]
    from struspattern_b200.capi import ITEM_PACKED_DT, RESULT_PACKED_DT, unpack_results
    out_res = torch.empty((nres + 4096) * RESULT_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(RESULT_DT)
    out_items = torch.empty((nitems + 4096) * ITEM_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(ITEM_DT)
    pk_res = pk_items = None
    if args.e2e_wire != "records" and world == 1:     # (one rank only: every rank would pin another 3.8 GB of host memory)
        # second measurement: packed records cross the link (16 / 12 bytes instead of 40 / 32) and stay packed in pinned memory
        pk_res = torch.empty((nres + 4096) * RESULT_PACKED_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(RESULT_PACKED_DT)
        pk_items = torch.empty((nitems + 4096) * ITEM_PACKED_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(ITEM_PACKED_DT)
    wire = ["records"]
    out_roffs = torch.empty(ndocs + len(subs) + 1, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
    out_ioffs = torch.empty(ndocs + len(subs) + 1, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
    out_stats = torch.empty(ndocs * STATS_DT.itemsize, dtype=torch.uint8, pin_memory=True).numpy().view(STATS_DT)
    out_status = torch.empty(ndocs, dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    errors = []

    trace_e2e = bool(os.environ.get("SP_BENCH_TRACE"))
    trace_t0 = [0.0]

    def e2e_pass(measure_only):
        nxt = [0]
        trace_t0[0] = time.perf_counter()
        lock = threading.Lock()

        def worker(w):
            try:
                torch.cuda.set_device(local_rank)
                lc, mc = pairs[w]
                while True:
                    with lock:
                        k = nxt[0]
                        nxt[0] += 1
                    if k >= len(subs):
                        return
                    b, e = subs[k]
                    t_a = time.perf_counter()
                    lc.launch(text[int(offs[b]):int(offs[e])], sub_offs[k], on_device=False)
                    t_b = time.perf_counter()
                    mc.launch_lexer_output(lc)
                    if measure_only:
                        res_begin[k + 1], item_begin[k + 1] = mc.count()
                    else:
                        rb, ib = res_begin[k], item_begin[k]
                        mc.count()
                        t_c = time.perf_counter()
                        if wire[0] == "packed":
                            re_, ie_ = res_begin[k + 1], item_begin[k + 1]
                            mc.fetch_packed_into(pk_res[rb:re_ + 1], out_roffs[b + k:e + k + 1], pk_items[ib:ie_ + 1], out_ioffs[b + k:e + k + 1], out_stats[b:e], out_status[b:e])
                        else:
                            mc.fetch_into(out_res[rb:res_begin[k + 1] + 1], out_roffs[b + k:e + k + 1], out_items[ib:item_begin[k + 1] + 1],
                                          out_ioffs[b + k:e + k + 1], out_stats[b:e], out_status[b:e])
                        if trace_e2e:
                            t_d = time.perf_counter()
                            sys.stderr.write("[e2e] thread %d sub %d: start %.1f lexer %.1f matcher %.1f fetch %.1f ms\n" % (w, k, (t_a - trace_t0[0]) * 1e3, (t_b - t_a) * 1e3, (t_c - t_b) * 1e3, (t_d - t_c) * 1e3))
            except Exception as exc:  # surfaced after the join
                errors.append(exc)

        threads = [threading.Thread(target=worker, args=(w,)) for w in range(len(pairs))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]

    e2e_pass(True)
    for k in range(len(subs)):
        res_begin[k + 1] += res_begin[k]
        item_begin[k + 1] += item_begin[k]
    assert res_begin[-1] == nres and item_begin[-1] == nitems, "sub-batches give other counts than the whole batch"
    e2e_pass(False)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_pass(False)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if int(out_status.max(initial=0)) != 0:
        raise RuntimeError("document errors on the synthetic corpus")
    if parity_mo is not None:
        # what the end to end path delivered to the host (unpacked, if packed) is the checked output of the device-resident step
        nr, ni = len(parity_mo["results"]), len(parity_mo["items"])
        if not ((out_res[:nr] == parity_mo["results"]).all() and (out_items[:ni] == parity_mo["items"]).all()):
            raise RuntimeError("end to end output differs from the output of the device-resident step")
        parity["e2e_records_equal"] = nr
    d2h = int(RESULT_DT.itemsize * nres + ITEM_DT.itemsize * nitems + 16 * (ndocs + len(subs)) + (STATS_DT.itemsize + 4) * ndocs)
    packed_line = None
    if pk_res is not None:
        wire[0] = "packed"
        e2e_pass(False)
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_pass(False)
        torch.cuda.synchronize()
        packed_s = time.perf_counter() - t0
        wire[0] = "records"
        # untimed: the packed records unpack into the records of the first measurement
        k_chk = min(len(subs), 2)
        for k in range(k_chk):
            b, e = subs[k]
            rb, re_, ib, ie_ = res_begin[k], res_begin[k + 1], item_begin[k], item_begin[k + 1]
            chk_res, chk_items = np.zeros(re_ - rb + 1, dtype=RESULT_DT), np.zeros(ie_ - ib + 1, dtype=ITEM_DT)
            unpack_results(pk_res[rb:re_ + 1], out_roffs[b + k:e + k + 1], pk_items[ib:ie_ + 1], out_ioffs[b + k:e + k + 1], chk_res, chk_items)
            if not ((chk_res[:re_ - rb] == out_res[rb:re_]).all() and (chk_items[:ie_ - ib] == out_items[ib:ie_]).all()):
                raise RuntimeError("packed records do not unpack into the records of the full wire format")
        packed_times = torch.tensor([packed_s * 1000.0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(packed_times, op=dist.ReduceOp.MAX)
        packed_line = {"value": float(args.bytes) * world * e2e_steps / (float(packed_times[0]) / 1000.0) / 1e9, "unit": "GB/s",
                       "d2h_bytes_per_step": int(RESULT_PACKED_DT.itemsize * nres + ITEM_PACKED_DT.itemsize * nitems + 16 * (ndocs + len(subs)) + (STATS_DT.itemsize + 4) * ndocs),
                       "what": "same pipeline with sp_matcher_fetch_packed: 16 B per result, 12 B per item, left packed in pinned memory (sp_unpack_results restores the records: "
                               "checked after the timed region on %d sub-batches; unpacking all of them on the host costs more than the link saves)" % k_chk}

    times = torch.tensor([ms, e2e_s * 1000.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(times[0]), float(times[1])
    if rank != 0:
        return None
    total_bytes = float(args.bytes) * world
    value = total_bytes * args.steps / (ms_max / 1000.0) / 1e9
    e2e_value = total_bytes * e2e_steps / (e2e_ms_max / 1000.0) / 1e9
    peak, peak_kind = measured_peak()
    step_s = ms_max / 1000.0 / args.steps
    # fused pipeline (lexems never leave the device): text + result records + item records + document offsets
    b_step = args.bytes + RESULT_DT.itemsize * nres + ITEM_DT.itemsize * nitems + 8 * ndocs
    dominant = max(kern_ms, key=lambda k: kern_ms[k])
    # algorithmic bytes of the dominant kernel per launch: count pass = token record (16 B) + event slot and
    # work count (8 B) read, result/item counts (8 B) written per token, one bit per rule instance
    b_dom = {"k_eval_count": 32.0 * nlex + nrules / 8.0, "k_eval_write": 16.0 * nlex + RESULT_DT.itemsize * nres + ITEM_DT.itemsize * nitems + nrules / 8.0,
             "k_match_documents": 16.0 * nlex + RESULT_DT.itemsize * nres + ITEM_DT.itemsize * nitems}[dominant]
    achieved = b_dom / (kern_ms[dominant] / 1000.0) / 1e9 if kern_ms[dominant] else 0.0
    traffic = NCU_TRAFFIC.get(dominant)
    line = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u32", "data": "synthetic",
        "config": {"workload": "C3 lexer+matcher: %d lexem definitions + %d symbols, %d token patterns over %d bytes of synthetic multilingual UTF-8 text per GPU in %d documents "
                               "(4-64 KiB); one batch of the 4 GiB corpus of BASELINE config 3" % (lx.table_info("definitions"), NSYMBOLS, NPATTERNS, args.bytes, ndocs),
                   "l2": "input (%.0f MiB text, %.0f MiB lexems) larger than L2, no flush needed" % (args.bytes / 2**20, 16.0 * nlex / 2**20),
                   "lexems_per_step": int(nlex), "results_per_step": int(nres), "items_per_step": int(nitems), "matcher": info},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": int(args.bytes + 8 * (ndocs + len(subs))), "d2h_bytes_per_step": d2h, "steps": e2e_steps,
                "pipeline": "%d host threads x (lexer context + matcher context + stream), sub-batches of %d MiB" % (args.e2e_threads, args.e2e_sub_bytes >> 20),
                "wire": "sp_result / sp_item records (40 / 32 B) copied to pinned memory", "packed_wire": packed_line,
                "numa_rank0": numa, "host_cpus": len(os.sched_getaffinity(0))},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": (traffic["bytes_per_token"] * nlex) if traffic else None, "traffic_source": traffic["source"] if traffic else None,
                     "algorithmic_bytes": b_dom, "peak_kind": peak_kind,
                     "step": {"algorithmic_bytes": b_step, "achieved": b_step / step_s / 1e9, "frac": b_step / step_s / 1e9 / peak,
                              "what": "SURVEY 8d fused form: text + result records + item records + document offsets over the whole step"},
                     "kernel_ms_per_step": dict(kern_ms, lexer_kernels=lex_ms)},
        "matches_per_s": nres * world * args.steps / (ms_max / 1000.0),
        "mtokens_per_s": nlex * world * args.steps / (ms_max / 1000.0) / 1e6,
        "lexer_only": {"value": args.bytes / (lex_ms / 1000.0) / 1e9 if lex_ms else None, "unit": "GB/s", "what": "level-1 kernels of the same step (C3 lexem set), per GPU"},
    }
    if parity:
        line["parity_checked"] = parity["results"]
        line["parity"] = parity
    if stats is not None:
        line["config"]["installs_per_step"] = int(stats["nof_programs_installed"].sum())
    if cpu_line:
        line["cpu_baseline"] = cpu_line
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=256 << 20)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--e2e-threads", type=int, default=4)
    ap.add_argument("--e2e-sub-bytes", type=int, default=8 << 20)
    ap.add_argument("--e2e-wire", default="both", choices=["both", "records"],
                    help="both: a second end-to-end measurement with the packed wire format (reported as e2e.packed_wire)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        line = run_reference(args, rank, world)
    else:
        if world > 1:
            import torch
            import torch.distributed as dist
            torch.cuda.set_device(local_rank)
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        line = run_ours(args, rank, world, local_rank)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
            dist.destroy_process_group()
    if line is not None:
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
