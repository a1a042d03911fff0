#!/usr/bin/env python3
"""Benchmark of the strusPattern hot path on B200 (BASELINE.json metric: GB/s of UTF-8 text lexed).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--bytes B]

A step = one pass of the lexer over one batch of synthetic text (workload C2 of SURVEY.md 8d: ~100 lexem
regexes over 1 GiB of UTF-8 text in documents of 4-64 KiB), per GPU (weak scaling: every rank owns its own
shard generated from (seed, rank)). `value` is measured with the text resident in HBM, `e2e` through the
C ABI with pinned host buffers (H2D of the text and D2H of the lexems inside the timed region).
--impl reference times the reference's CPU path of level 1; Hyperscan is not available here, so this is the
CPU restatement (oracle/lexer_oracle.cpp, kind "port") on all host cores over a bounded sample.
Prints ONE JSON line on rank 0."""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GB/s UTF-8 text lexed+matched at 1/2/4/8 B200 vs same-box hyperscan CPU path"
SEED = 0x5354525553000002
FIXED_WORDS = None


def fixed_words():
    from struspattern_b200 import workloads as W
    return W.PHRASE_FIRST + W.PHRASE_SECOND + W.ABBREVIATIONS + ["the", "of", "and", "to", "in", "a", "is", "that", "for", "it"]


def make_corpus(nbytes, rank, pinned):
    """-> (host uint8 array, offsets, keepalive)"""
    from bench_util import corpus
    vocab = corpus.vocabulary(SEED, 50000, fixed_words())
    offs = corpus.layout(SEED + 1000 * rank, nbytes)
    keep = None
    if pinned:
        import torch
        keep = torch.empty(int(nbytes) + 64, dtype=torch.uint8, pin_memory=True)
        text = keep.numpy()[:nbytes]
    else:
        text = np.zeros(int(nbytes), dtype=np.uint8)
    corpus.fill(SEED + 1000 * rank, vocab, text, offs)
    return text, offs, keep


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


def measured_peak():
    try:
        d = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(d["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def cpu_lexer_rate(text, offs, seconds_target, nthreads):
    """Oracle (CPU restatement of level 1) on a bounded prefix of the corpus -> (GB/s, sample bytes, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle_util import OracleLexer
    from struspattern_b200 import workloads as W
    o = W.define_lexer(OracleLexer())
    # probe on ~1 MiB, then size the sample for the target duration
    nd = int(np.searchsorted(offs, 1 << 20, side="right"))
    nd = max(1, min(nd, len(offs) - 1))
    t = time.perf_counter()
    o.scan(text[:int(offs[nd])], offs[:nd + 1], nthreads=nthreads)
    dt = time.perf_counter() - t
    rate = float(offs[nd]) / max(dt, 1e-9)
    want = rate * seconds_target
    nd2 = int(np.searchsorted(offs, want, side="right"))
    nd2 = max(nd, min(nd2, len(offs) - 1))
    t = time.perf_counter()
    o.scan(text[:int(offs[nd2])], offs[:nd2 + 1], nthreads=nthreads)
    dt = time.perf_counter() - t
    return float(offs[nd2]) / dt / 1e9, int(offs[nd2]), dt


KERNELS = ["k_lex_count", "k_lex_verify", "k_lex_emit", "k_lex_finish", "k_lex_resolve", "k_lex_write", "k_scan",
           "k_isl_flags", "k_isl_resolve", "k_ordpos", "k_isl_slots", "k_slot_scan", "k_lex_fused", "k_fused_scan", "k_fused_write", "k_lex_resolve_chunks"]


def algorithmic_bytes(kernel, c):
    """Bytes one launch of `kernel` has to move by the algorithm (DESIGN.md 'kernels'), c = batch counters."""
    n_text, nev, nlex, ndocs, nchunks = c["text_bytes"], c["raw_events"], c["lexems"], c["ndocs"], c["chunks"]
    return {
        "k_lex_count": n_text + 12 * nchunks,                 # text read once, start/end state + count per chunk
        "k_lex_verify": 9 * nchunks,
        "k_lex_emit": n_text + 16 * nev + 16 * nchunks,       # text read once, one 16 B event record per raw event
        "k_lex_finish": 36 * nev,                             # event record read + written, document word written (text re-reads hit L2)
        "k_lex_resolve": 16 * nev + 16 * nlex + 16 * ndocs,   # events read, kept match events written
        "k_lex_write": 32 * nlex + 24 * ndocs,                # kept match events read, 16 B lexem records written
        "k_scan": 12 * (nchunks + ndocs),
        # island pipeline (per step totals of the kernel group; divided by its launches by the caller)
        "k_isl_flags": 20 * nev + nev + 4 * c.get("islands", 0),       # events + document ids read, flags + island starts written
        "k_isl_resolve": 2 * 20 * nev + 20 * c.get("elements", 0),    # events read by both passes, dense elements written
        "k_ordpos": 20 * c.get("elements", 0) + 16 * nlex,           # dense elements read once, lexem records written
        # slot pipeline (default): events + document ids read once, one 16 B slot per event written; slots read once, lexems written
        "k_isl_slots": 20 * nev + 16 * nev * c.get("ar_mult", 1),
        "k_slot_scan": 16 * nev * c.get("ar_mult", 1) + 16 * nlex + 16 * ndocs,
        # fused segment pipeline (default): text read once, one 16 B record per kept element (>= lexems) written, per chunk
        # start state + event offset read and 30 B of counters written; then kept elements read once, 16 B lexem records written
        "k_lex_fused": n_text + 16 * nlex + 42 * nchunks,
        "k_fused_scan": 40 * nchunks + 28 * ndocs,
        "k_fused_write": 32 * nlex + 20 * nchunks,
        # chunk resolution over completed events (default): event + document id + segment flag read, kept elements written
        "k_lex_resolve_chunks": 20 * nev + 16 * nlex + 42 * nchunks,
    }[kernel]


# DRAM traffic (dram__bytes_read.sum + dram__bytes_write.sum) of one launch from the ncu --set full capture summarised in
# profiles/r01_lexer_v10_ncu.txt (256 MiB of the same corpus), as bytes per unit the launch processes; scaled to this run's units
NCU_TRAFFIC = {
    "k_lex_count": (0.323008e9 + 0.010925e9, "text_bytes", 268435456),
    "k_lex_emit": (0.539853e9 + 1.577043e9, "raw_events", 89434349),
    "k_lex_finish": (1.699961e9 + 1.731728e9, "raw_events", 89434349),
    "k_lex_resolve_chunks": (2.123800e9 + 1.211899e9, "raw_events", 89434349),
    "k_fused_write": (1.182111e9 + 1.070465e9, "lexems", 69551577),
}


def ncu_traffic(kernel, c):
    if kernel not in NCU_TRAFFIC:
        return None
    nbytes, unit, units = NCU_TRAFFIC[kernel]
    return nbytes * c[unit] / units


def run_reference(args, rank, world):
    if rank != 0:
        return None
    nthreads = os.cpu_count() or 1
    text, offs, _ = make_corpus(min(args.bytes, 256 << 20), 0, pinned=False)
    per_step = max(2.0, 60.0 / max(1, args.steps + args.warmup))
    rates = []
    sample = 0
    for i in range(args.warmup + args.steps):
        r, sample, dt = cpu_lexer_rate(text, offs, per_step, nthreads)
        if i >= args.warmup:
            rates.append((r, dt))
    total_t = sum(dt for _, dt in rates)
    value = sample * len(rates) / total_t / 1e9
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1000.0 * total_t / len(rates), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C2 lexer only: %d lexem definitions, CPU sample of the 1 GiB corpus" % args.ndefs,
                   "sample_bytes": sample},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": nthreads, "kind": "port",
                         "sample": "first %d bytes of the C2 corpus per step; level-1 CPU restatement (Hyperscan/libtre not installable here)" % sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }


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


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from struspattern_b200 import LEXEM_DT, Lexer
    from struspattern_b200 import workloads as W

    torch.cuda.set_device(local_rank)
    numa = bind_to_gpu_numa_node(local_rank) if world > 1 else {"bound": False, "why": "single rank"}
    lx = W.define_lexer(Lexer())
    text, offs, keep = make_corpus(args.bytes, rank, pinned=True)
    ndocs = len(offs) - 1
    dev_text = torch.empty(len(text) + 64, dtype=torch.uint8, device="cuda")
    dev_text[:len(text)].copy_(keep[:len(text)], non_blocking=False)
    ctx = lx.create_context()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device resident
    sampler = ClockSampler(local_rank)
    sampler.start()
    for _ in range(args.warmup):
        ctx.launch(dev_text.data_ptr(), offs, on_device=True)
        ctx.count()
    counters = {k: ctx.counter(k) for k in ("raw_events", "lexems", "chunks", "fixup_chunks", "fixup_rounds", "text_bytes", "kernel_launches", "islands", "elements", "fused", "fused_fallbacks", "second_half")}
    counters["ndocs"] = ndocs
    counters["ar_mult"] = 2 if lx.table_info("symbols") else 1
    ctx.counter("profile:on"); ctx.counter("profile:reset")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    launches = 0
    for _ in range(args.steps):
        ctx.launch(dev_text.data_ptr(), offs, on_device=True)
        ctx.count()
        launches += ctx.counter("kernel_launches")
    e1.record()
    barrier()
    clocks = sampler.stop()
    ms = e0.elapsed_time(e1)
    ktimes = {k: (ctx.counter("time_us:" + k), ctx.counter("calls:" + k)) for k in KERNELS}
    ctx.counter("profile:off")

    # ---- end to end through the C ABI with pinned host buffers: the batch is cut into sub-batches of whole documents
    # that several host threads (one context + CUDA stream each) push through scan + fetch, so that the H2D copy of one
    # sub-batch, the kernels of another and the D2H copy of a third overlap (PCIe is full duplex)
    import threading
    nlex = counters["lexems"]
    del ctx
    out_keep = torch.empty((nlex + 1024) * 16, dtype=torch.uint8, pin_memory=True)
    out_lex = out_keep.numpy().view(LEXEM_DT)
    out_offs = torch.empty(ndocs + 1, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64)
    out_status = torch.empty(ndocs, dtype=torch.int32, pin_memory=True).numpy().view(np.uint32)
    e2e_steps = max(1, min(args.steps, 3))
    cuts = [0]
    while cuts[-1] < ndocs:
        nxt = int(np.searchsorted(offs, offs[cuts[-1]] + np.uint64(args.e2e_sub_bytes), side="right")) - 1
        cuts.append(min(ndocs, max(nxt, cuts[-1] + 1)))
    subs = [(cuts[i], cuts[i + 1]) for i in range(len(cuts) - 1)]
    sub_offs = [np.ascontiguousarray(offs[b:e + 1] - offs[b]) for b, e in subs]
    sub_lex_begin = [0] * (len(subs) + 1)     # filled by the untimed first pass (sizes are deterministic)
    streams = [torch.cuda.Stream() for _ in range(args.e2e_threads)]
    ctxs = [lx.create_context(s.cuda_stream) for s in streams]
    # several ranks on one host: waiting host threads sleep instead of spinning (ranks x threads is about the CPU count)
    blocking = world * args.e2e_threads * 2 > len(os.sched_getaffinity(0))
    if blocking:
        for c in ctxs:
            c.counter("sync:blocking")
    scratch_offs = [np.zeros(max(e - b for b, e in subs) + 1, dtype=np.uint64) for _ in ctxs]
    errors = []

    def e2e_pass(measure_only):
        nxt = [0]
        lock = threading.Lock()

        def worker(w):
            try:
                torch.cuda.set_device(local_rank)
                while True:
                    with lock:
                        k = nxt[0]
                        nxt[0] += 1
                    if k >= len(subs):
                        return
                    b, e = subs[k]
                    ctxs[w].launch(text[int(offs[b]):int(offs[e])], sub_offs[k], on_device=False)
                    if measure_only:
                        sub_lex_begin[k + 1] = ctxs[w].count()
                    else:
                        lb = sub_lex_begin[k]
                        ctxs[w].fetch_into(out_lex[lb:sub_lex_begin[k + 1] + 1], scratch_offs[w], out_status[b:e])
                        out_offs[b:e + 1] = scratch_offs[w][:e - b + 1] + np.uint64(lb)
            except Exception as exc:  # surfaced after the join
                errors.append(exc)

        threads = [threading.Thread(target=worker, args=(w,)) for w in range(len(ctxs))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]

    e2e_pass(True)
    for k in range(len(subs)):
        sub_lex_begin[k + 1] += sub_lex_begin[k]
    assert sub_lex_begin[-1] == nlex, "sub-batches give another lexem count than the whole batch"
    e2e_pass(False)
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_pass(False)
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if int(out_status.max(initial=0)) != 0:
        raise RuntimeError("lexer reported document errors on the synthetic corpus")
    if args.verify_e2e:
        whole = lx.create_context()
        ref = whole.scan(dev_text.data_ptr(), offs, on_device=True)
        assert (ref["offsets"] == out_offs).all() and (ref["lexems"] == out_lex[:nlex]).all(), "pipelined end-to-end output differs from the single batch"
        del whole

    # what the link allows: plain pinned copies of the same sizes, both directions at once (the end-to-end leg is PCIe bound)
    pc0, pc1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s_up, s_down = torch.cuda.Stream(), torch.cuda.Stream()
    dev_out = torch.empty(16 * nlex, dtype=torch.uint8, device="cuda")
    link_ms = None
    barrier()   # all ranks copy at the same time: the bound is what the host gives every rank then
    for _ in range(2):
        torch.cuda.synchronize()
        pc0.record()
        s_up.wait_event(pc0); s_down.wait_event(pc0)
        with torch.cuda.stream(s_up):
            dev_text[:len(text)].copy_(keep[:len(text)], non_blocking=True)
        with torch.cuda.stream(s_down):
            out_keep[:16 * nlex].copy_(dev_out, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s_up); torch.cuda.current_stream().wait_stream(s_down)
        pc1.record()
        torch.cuda.synchronize()
        link_ms = pc0.elapsed_time(pc1) if link_ms is None else min(link_ms, pc0.elapsed_time(pc1))
    del dev_out

    times = torch.tensor([ms, e2e_s * 1000.0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_max, e2e_ms_max = float(times[0]), float(times[1])
    if rank != 0:
        return None
    total_bytes = float(args.bytes) * world
    value = total_bytes * args.steps / (ms_max / 1000.0) / 1e9
    e2e_value = total_bytes * e2e_steps / (e2e_ms_max / 1000.0) / 1e9
    dominant = max(KERNELS, key=lambda k: ktimes[k][0])
    dom_us, dom_calls = ktimes[dominant]
    peak, peak_kind = measured_peak()
    # timed groups may cover several launches per step: algorithmic bytes of the group per step / its time per step
    achieved = algorithmic_bytes(dominant, counters) / (dom_us / args.steps * 1e-6) / 1e9 if dom_us else 0.0
    line = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_max / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic",
        "config": {"workload": "C2 lexer only: %d lexem definitions over %d bytes of synthetic UTF-8 text per GPU in %d documents (4-64 KiB)"
                               % (lx.table_info("definitions"), args.bytes, ndocs),
                   "l2": "input (%.0f MiB) larger than L2, no flush needed" % (args.bytes / 2**20),
                   "lexems_per_step": counters["lexems"], "raw_events_per_step": counters["raw_events"],
                   "fwd_states": lx.table_info("fwd_states"), "fwd_table_bytes": lx.table_info("fwd_table_bytes"),
                   "fixup_chunks": counters["fixup_chunks"], "chunks": counters["chunks"],
                   "second_half": {0: "per document loop", 1: "dense island pipeline", 2: "island/slot pipeline", 3: "fused segment kernel",
                                   4: "chunk resolution over completed events"}.get(counters["second_half"], "?")},
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": int(args.bytes + 8 * (ndocs + 1)),
                "d2h_bytes_per_step": int(16 * nlex + 8 * (ndocs + len(subs)) + 4 * ndocs), "steps": e2e_steps,
                "pipeline": "%d host threads x (context + stream), sub-batches of %d MiB" % (args.e2e_threads, args.e2e_sub_bytes >> 20),
                "numa_rank0": numa, "host_cpus": len(os.sched_getaffinity(0)), "blocking_waits": bool(blocking),
                "link_bound": {"value": float(args.bytes) / (link_ms / 1000.0) / 1e9, "unit": "GB/s",
                               "what": "text bytes / time of plain pinned copies of the same H2D and D2H bytes issued together (rank 0)"}},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "kernel": dominant, "achieved": achieved, "peak": peak, "unit": "GB/s",
                     "frac": achieved / peak, "traffic": ncu_traffic(dominant, counters), "algorithmic_bytes": algorithmic_bytes(dominant, counters),
                     "traffic_source": "profiles/r01_lexer_v10_ncu.txt (ncu --set full, 256 MiB of this corpus, scaled by units)", "peak_kind": peak_kind,
                     "kernel_ms_per_step": {k: ktimes[k][0] / 1000.0 / args.steps for k in KERNELS}},
        "matches_per_s": counters["lexems"] * world * args.steps / (ms_max / 1000.0),
    }
    if world == 1 and not args.no_cpu:
        r, sample, dt = cpu_lexer_rate(text, offs, args.cpu_seconds, os.cpu_count() or 1)
        line["cpu_baseline"] = {"value": r, "unit": "GB/s", "cores": os.cpu_count() or 1, "kind": "port",
                                "sample": "first %d bytes of the same corpus, %.1f s; level-1 CPU restatement (Hyperscan not installable here)" % (sample, dt)}
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--bytes", type=int, default=1 << 30)
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--e2e-threads", type=int, default=4)
    ap.add_argument("--e2e-sub-bytes", type=int, default=64 << 20)
    ap.add_argument("--verify-e2e", action="store_true")
    args = ap.parse_args()
    args.ndefs = 86
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
