"""N>1 path on the CPU: two gloo ranks shard a batch by documents, run their shard (here through the ORACLES, the
only engines a CPU box has) and gather; the gathered batch must equal the single-rank run. This covers the host-side
logic bench.py / a multi-GPU caller uses: shard bounds, offset rebasing, gather in document order."""
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import os, sys, pickle
import numpy as np
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch.distributed as dist
from struspattern_b200 import sharding
from lexer_cases import define_set, make_batch
from oracle_util import OracleLexer, RefMatcher
import matcher_cases as MC
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
text, offs = make_batch(seed=3, ndocs=40, mean_bytes=800)
lx = define_set(OracleLexer(), "words_numbers", with_symbols=False)
bounds = sharding.shard_bounds(offs, world)
mytext, myoffs = sharding.take_shard(text, offs, bounds[rank])
lex = lx.scan(mytext, myoffs)
lex = {k: lex[k] for k in ("lexems", "offsets", "status")}
m = RefMatcher()
for k, (op, terms, rng) in enumerate([(MC.OP_SEQUENCE, (1, 2), 2), (MC.OP_WITHIN, (2, 1), 4), (MC.OP_ANY, (2, 4), 1)]):
    for i, t in enumerate(terms):
        m.push_term(t); m.attach_variable("v%d" % i)
    m.push_expression(op, len(terms), rng, 0)
    m.define_pattern("p%d" % k, "", True)
m.compile()
res = m.run(lex["lexems"], lex["offsets"])
res["results"]["name"] = [0 if n == 0 else int(m.name(n)[1:]) + 1 for n in res["results"]["name"]]   # rank independent handles
res["items"]["variable"] = [0 if n == 0 else int(m.name(n)[1:]) + 1 for n in res["items"]["variable"]]
glex = sharding.gather_batches(lex, {"lexems": "offsets"}, group=dist.group.WORLD)
gres = sharding.gather_batches(res, {"results": "result_offsets", "items": "item_offsets"}, group=dist.group.WORLD)
if rank == 0:
    pickle.dump({"bounds": bounds, "lex": glex, "res": gres}, open(OUT, "wb"))
dist.barrier()
dist.destroy_process_group()
'''


def run_world(world, out):
    code = "ROOT=%r\nOUT=%r\n" % (ROOT, out) + WORKER
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world + os.getpid() % 400), "-c", code]
    # torch.distributed.run has no -c: write the worker to a file
    script = out + ".py"
    open(script, "w").write(code)
    cmd = cmd[:-2] + [script]
    res = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-3000:]


def test_shard_bounds_partition_every_document_once():
    from struspattern_b200 import sharding
    rng = np.random.default_rng(1)
    for world in (1, 2, 3, 8):
        sizes = rng.integers(0, 5000, size=57)
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.uint64)
        b = sharding.shard_bounds(offs, world)
        assert b[0][0] == 0 and b[-1][1] == 57 and all(b[i][1] == b[i + 1][0] for i in range(world - 1))
        loads = [int(offs[e] - offs[s]) for s, e in b]
        assert max(loads) - min(loads) <= 2 * 5000 or world == 1
    assert sharding.shard_bounds(np.zeros(1, dtype=np.uint64), 4) == [(0, 0)] * 4


def test_two_gloo_ranks_equal_one(tmp_path):
    import pickle
    one, two = str(tmp_path / "w1.pkl"), str(tmp_path / "w2.pkl")
    run_world(1, one)
    run_world(2, two)
    a, b = pickle.load(open(one, "rb")), pickle.load(open(two, "rb"))
    assert len(b["bounds"]) == 2 and b["bounds"][0][1] > 0
    for part in ("lex", "res"):
        for k in a[part]:
            assert np.array_equal(a[part][k], b[part][k]), (part, k)
    assert len(a["lex"]["lexems"]) > 1000 and len(a["res"]["results"]) > 10
