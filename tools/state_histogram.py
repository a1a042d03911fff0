"""Which rows of the forward table does the C2 corpus read? (decides what a shared-memory copy of hot rows would hold)
TEST/ANALYSIS TOOL: uses tests/hostsim (the product's host compiler + a plain CPU walk)."""
import ctypes, os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import bench
from hostsim_util import HostSimLexer as HostSim
from struspattern_b200 import workloads as W

nbytes = int(sys.argv[1]) if len(sys.argv) > 1 else (8 << 20)
text, offs, _ = bench.make_corpus(nbytes, 0, pinned=False)
hs = HostSim(chunk=512, warmup=64, mode=3)
W.define_lexer(hs)
hs.compile()
n = hs.info("fwd_states"); stride = hs.info("fwd_stride")
hist = np.zeros(n, dtype=np.uint64)
lib = HostSim.lib()
lib.hsim_state_histogram.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32, ctypes.c_void_p]
offs = np.ascontiguousarray(offs, dtype=np.uint64)
rc = lib.hsim_state_histogram(hs.h, text.ctypes.data, offs.ctypes.data, len(offs) - 1, hist.ctypes.data)
assert rc == 0
tot = hist.sum()
print("states", n, "stride", stride, "row bytes", stride * 2, "visited states", int((hist > 0).sum()))
srt = np.sort(hist)[::-1]
for H in (32, 64, 128, 160, 192, 256, 384, 512, 1024):
    print("H=%4d (%6.1f KB): best rows cover %.4f, lowest ids cover %.4f" % (H, H * stride * 2 / 1024, srt[:H].sum() / tot, hist[:H].sum() / tot))

# by group: [plain][plain,cut][acc,cut][acc]
accbase, cutlo, cutn = hs.info("acc_base"), hs.info("cut_lo"), hs.info("cut_states")
groups = [("plain", 0, cutlo), ("plain,cut", cutlo, accbase), ("acc,cut", accbase, cutlo + cutn), ("acc", cutlo + cutn, n)]
for name, lo, hi in groups:
    h = np.sort(hist[lo:hi])[::-1]
    print("%-10s states %6d share of reads %.4f; top 8/16/32/64 rows of the group cover %s of all reads" % (
        name, hi - lo, h.sum() / tot, " ".join("%.4f" % (h[:k].sum() / tot) for k in (8, 16, 32, 64))))
# one budget of H rows spent greedily over all groups = the best-rows line above; how many rows per group is that?
order = np.argsort(hist)[::-1]
for H in (32, 64, 128):
    top = order[:H]
    print("H=%d: rows per group" % H, [int(((top >= lo) & (top < hi)).sum()) for _, lo, hi in groups])
