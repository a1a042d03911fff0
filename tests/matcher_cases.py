"""Shared rule sets and seeded token documents for the matcher parity tests (reference oracle, CPU host
simulation and GPU). Every generator drives a matcher object through the common interface
(push_term / attach_variable / push_expression / push_pattern / define_pattern / compile / run) so that the very
same call sequence reaches the reference and the product."""
import numpy as np

from oracle_util import LEXEM_DT, canonical_results

OP_SEQUENCE, OP_SEQUENCE_IMM, OP_SEQUENCE_STRUCT, OP_WITHIN, OP_WITHIN_STRUCT, OP_ANY, OP_AND = range(7)
OP_NAMES = {OP_SEQUENCE: "sequence", OP_SEQUENCE_IMM: "sequence_imm", OP_SEQUENCE_STRUCT: "sequence_struct",
            OP_WITHIN: "within", OP_WITHIN_STRUCT: "within_struct", OP_ANY: "any", OP_AND: "and"}
DELIM = 1 << 24  # termId(SentenceDelim, 0), tests/utils/testUtils.cpp:74-77


class Zipf:
    """ZipfDistribution of tests/utils/testUtils.cpp:39-72 (rank 1..size, weight 1/k^S, S=0 means exponent 1)."""

    def __init__(self, size, S=0.0):
        k = np.arange(1, size + 1, dtype=np.float64)
        self.cum = np.cumsum(1.0 / (k ** S if S > 1e-12 else k))

    def draw(self, rng, n=None):
        u = rng.random(n) * self.cum[-1]
        return np.searchsorted(self.cum, u, side="right") + 1


def random_documents(seed, ndocs, ntokens, nfeatures, delim_prob=1.0 / 13, tokens_per_pos=1.0):
    """createRandomDocument (tests/utils/testUtils.cpp:79-96): Zipf tokens at ordpos 1..n, a sentence delimiter at
    the SAME ordpos with probability 1/13; origpos = running index, origsize 1
    (tests/randomTokenPatternMatch/src/testRandomTokenPatternMatch.cpp:145)."""
    rng = np.random.default_rng(seed)
    feat = Zipf(nfeatures)
    toks, offs = [], [0]
    for _ in range(ndocs):
        n = int(ntokens if np.isscalar(ntokens) else rng.integers(ntokens[0], ntokens[1] + 1))
        ids = feat.draw(rng, n)
        delim = rng.random(n) < delim_prob
        pos = 1
        idx = 0
        for i in range(n):
            toks.append((int(ids[i]), pos, idx, 1)); idx += 1
            if delim[i]:
                toks.append((DELIM, pos, idx, 1)); idx += 1
            if tokens_per_pos >= 1.0 or rng.random() < tokens_per_pos:
                pos += 1
        offs.append(len(toks))
    arr = np.zeros(len(toks), dtype=LEXEM_DT)
    if toks:
        a = np.array(toks, dtype=np.uint32)
        arr["id"], arr["ordpos"], arr["origpos"], arr["origsize"] = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    return arr, np.array(offs, dtype=np.uint64)


def define_term_rule(m, op, terms, rng_, card, variables=True):
    """createTermOpRule (testRandomTokenPatternMatch.cpp:48-74)."""
    argc = len(terms)
    if op in (OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT):
        m.push_term(DELIM)
        argc += 1
    for i, t in enumerate(terms):
        m.push_term(int(t))
        if variables:
            m.attach_variable("A%u" % i)
    m.push_expression(op, argc, int(rng_), int(card))


def define_c1_rules(m, seed, nrules, nfeatures, ops=(OP_SEQUENCE, OP_WITHIN, OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT, OP_ANY),
                    nterms=(2, 2), card_prob=0.0, feature_skew=0.8):
    """createRules (testRandomTokenPatternMatch.cpp:90-123): terms Zipf(F, 0.8), range Zipf(10, 1.7)+1,
    operator Zipf(5) over `ops`; every rule is a visible pattern named <op>_0_1..."""
    rng = np.random.default_rng(seed)
    feat, rangedist, seldist = Zipf(nfeatures, feature_skew), Zipf(10, 1.7), Zipf(len(ops))
    for _ in range(nrules):
        rng_ = int(rangedist.draw(rng)) + 1
        n = int(rng.integers(nterms[0], nterms[1] + 1))
        terms = feat.draw(rng, n)
        op = ops[int(seldist.draw(rng)) - 1]
        card = 0
        if card_prob and rng.random() < card_prob and n > 1:
            card = int(rng.integers(1, n + 1))
        define_term_rule(m, op, terms, rng_, card)
        m.define_pattern(OP_NAMES[op] + "".join("_%u" % i for i in range(n)), "", True)


def define_tree_rules(m, seed, nrules, nfeatures, depth=2, ops=(OP_SEQUENCE, OP_SEQUENCE_IMM, OP_WITHIN, OP_ANY, OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT),
                      ref_prob=0.15, card_prob=0.2, max_range=12):
    """Random expression trees: nested expressions, references to earlier (visible and private) patterns,
    cardinalities, variables on leaves and on sub expressions."""
    rng = np.random.default_rng(seed)
    feat = Zipf(nfeatures, 0.8)
    names = []
    varcnt = [0]

    def node(level):
        if level > 0 and rng.random() < 0.45:
            op = ops[int(rng.integers(0, len(ops)))]
            argc = int(rng.integers(2, 4))
            if op in (OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT):
                m.push_term(DELIM)
            for _ in range(argc):
                node(level - 1)
            total = argc + (1 if op in (OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT) else 0)
            card = int(rng.integers(1, argc + 1)) if rng.random() < card_prob else 0
            m.push_expression(op, total, int(rng.integers(1, max_range + 1)) * (level + 1), card)
        elif names and rng.random() < ref_prob:
            m.push_pattern(names[int(rng.integers(0, len(names)))])
        else:
            m.push_term(int(feat.draw(rng)))
        if rng.random() < 0.7:
            varcnt[0] += 1
            m.attach_variable("V%u" % (varcnt[0] % 7))

    for r in range(nrules):
        op = ops[int(rng.integers(0, len(ops)))]
        argc = int(rng.integers(2, 4))
        if op in (OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT):
            m.push_term(DELIM)
        for _ in range(argc):
            node(depth - 1)
        total = argc + (1 if op in (OP_SEQUENCE_STRUCT, OP_WITHIN_STRUCT) else 0)
        card = int(rng.integers(1, argc + 1)) if rng.random() < card_prob else 0
        m.push_expression(op, total, int(rng.integers(1, max_range + 1)) * depth, card)
        name = "P%u" % r
        visible = rng.random() < 0.8
        m.define_pattern(name, "", bool(visible))
        names.append(name)


def compare_outputs(ref_m, ref_out, got_m, got_out, check_stats=True, check_order=False):
    """Parity contract: per document the multiset of complete result records incl. ordered items,
    the document status, and the exact event cascade counters."""
    rs, gs = ref_out["status"], got_out["status"]
    assert [int(x != 0) for x in rs] == [int(x != 0) for x in gs], "document status differs"
    ok = [d for d in range(len(rs)) if rs[d] == 0]
    rc = canonical_results(ref_out, ref_m.result_name, ref_m.item_name, sort=not check_order)
    gc = canonical_results(got_out, got_m.result_name, got_m.item_name, sort=not check_order)
    for d in ok:
        assert rc[d] == gc[d], "results of document %d differ:\nref %r\ngot %r" % (d, rc[d][:6], gc[d][:6])
    if check_stats:
        for f in ("nof_programs_installed", "nof_altkey_programs_installed", "nof_signals_fired", "nof_open_patterns", "nof_events"):
            for d in ok:
                assert int(ref_out["stats"][f][d]) == int(got_out["stats"][f][d]), "%s of document %d: ref %d got %d" % (
                    f, d, int(ref_out["stats"][f][d]), int(got_out["stats"][f][d]))
    return sum(len(rc[d]) for d in ok)


def _mix(h, v):
    """64 bit mixing step (splitmix64 finaliser) of numpy uint64 arrays; wraps around on purpose"""
    with np.errstate(over="ignore"):
        h = (h ^ v.astype(np.uint64)) * np.uint64(0x9E3779B97F4A7C15)
        h ^= h >> np.uint64(29)
        h = h * np.uint64(0xBF58476D1CE4E5B9)
        h ^= h >> np.uint64(32)
    return h


def _name_hashes(name_of, handles):
    import zlib
    top = int(handles.max(initial=0)) + 1
    table = np.zeros(top, dtype=np.uint64)
    for h in np.unique(handles):
        table[int(h)] = zlib.crc32(name_of(int(h)).encode()) + 1
    return table[handles]


def multiset_hashes(m, out):
    """Per document: (number of results, order independent 64 bit hash of the complete result records with their items
    in order). Vectorised form of canonical_results for samples of millions of results."""
    res, items = out["results"], out["items"]
    ro, io = out["result_offsets"].astype(np.int64), out["item_offsets"].astype(np.int64)
    ndocs = len(ro) - 1
    ih = _name_hashes(m.item_name, items["variable"]) if len(items) else np.zeros(0, dtype=np.uint64)
    for f in ("ordpos", "ordend", "origseg", "origpos", "endseg", "endpos"):
        ih = _mix(ih, items[f])
    doc_of_res = np.repeat(np.arange(ndocs), np.diff(ro))
    first_item = io[doc_of_res] + res["item_begin"].astype(np.int64)        # absolute index of the first item of a result
    cnt = res["item_count"].astype(np.int64)
    rh = _name_hashes(m.result_name, res["name"]) if len(res) else np.zeros(0, dtype=np.uint64)
    for f in ("ordpos", "ordend", "origseg", "origpos", "endseg", "endpos", "item_count"):
        rh = _mix(rh, res[f])
    if len(items) and len(res):
        owner = np.repeat(np.arange(len(res)), cnt)                         # result of every item, in result order
        idx = np.repeat(first_item, cnt) + (np.arange(int(cnt.sum())) - np.repeat(np.cumsum(cnt) - cnt, cnt))
        pos = (idx - np.repeat(first_item, cnt)).astype(np.uint64)
        with np.errstate(over="ignore"):
            contrib = _mix(ih[idx], pos + np.uint64(1))                     # position dependent: the item order is part of the record
        isum = np.zeros(len(res), dtype=np.uint64)
        np.add.at(isum, owner, contrib)
        rh = _mix(rh, isum)
    dh = np.zeros(ndocs, dtype=np.uint64)
    with np.errstate(over="ignore"):
        np.add.at(dh, doc_of_res, rh)
    return np.diff(ro), dh


def compare_outputs_fast(ref_m, ref_out, got_m, got_out, exact_docs=32):
    """compare_outputs for large samples: the first `exact_docs` documents record by record, all documents by result count,
    multiset hash and the exact cascade counters. Returns the number of results compared."""
    rs, gs = ref_out["status"], got_out["status"]
    assert ((rs != 0) == (gs != 0)).all(), "document status differs"
    ok = rs == 0
    nd = min(exact_docs, len(rs))
    sub = lambda o: {"results": o["results"][:int(o["result_offsets"][nd])], "result_offsets": o["result_offsets"][:nd + 1],
                     "items": o["items"][:int(o["item_offsets"][nd])], "item_offsets": o["item_offsets"][:nd + 1],
                     "stats": o["stats"][:nd], "status": o["status"][:nd]}
    compare_outputs(ref_m, sub(ref_out), got_m, sub(got_out), check_order=False)
    rc, rh = multiset_hashes(ref_m, ref_out)
    gc, gh = multiset_hashes(got_m, got_out)
    bad = np.nonzero(ok & ((rc != gc) | (rh != gh)))[0]
    assert len(bad) == 0, "results of document %d differ (%d vs %d results)" % (int(bad[0]), int(rc[bad[0]]), int(gc[bad[0]]))
    for f in ("nof_programs_installed", "nof_altkey_programs_installed", "nof_signals_fired", "nof_open_patterns", "nof_events"):
        diff = np.nonzero(ok & (ref_out["stats"][f] != got_out["stats"][f]))[0]
        assert len(diff) == 0, "%s of document %d: ref %d got %d" % (f, int(diff[0]), int(ref_out["stats"][f][diff[0]]), int(got_out["stats"][f][diff[0]]))
    return int(rc[ok].sum())
