"""Lexem sets of the benchmark configurations (BASELINE.json configs). The expressions follow
the published strus rule example (reference doc/webpage/introduction_struspattern.htm:82-97, 149-151):
words / capitalised / lower case words, numbers, decimals, amounts, dates, e-mail, URL, sentence delimiter,
abbreviations, punctuation and fixed multi word phrases, with the levels of that example."""
from .capi import BIND_CONTENT, BIND_PREDECESSOR

ABBREVIATIONS = ["mr", "mrs", "ms", "dr", "prof", "st", "jr", "sr", "inc", "ltd", "co", "corp", "vs", "etc", "eg", "ie",
                 "no", "vol", "pp", "ca", "fig", "dept", "univ", "assn", "bros", "gen", "col", "lt", "sgt", "capt"]
PHRASE_FIRST = ["new", "san", "los", "united", "high", "prime", "world", "south", "north", "east", "west", "real",
                "white", "black", "red", "open", "free", "last", "next", "first"]
PHRASE_SECOND = ["york", "diego", "angeles", "states", "school", "minister", "war", "africa", "korea", "coast"]
TLDS = ["com", "org", "net", "edu", "gov", "de", "ch", "uk", "fr", "io"]

# ids of the main lexem classes (tests and the matcher workloads refer to them)
WORD, CAPWORD, LOWORD, NUMBER, DECIMAL, AMOUNT, DATE, EMAIL, URL, SENT, ABBREV, PUNCT, QUOTE, PHRASE = range(1, 15)


def phrase_list(n):
    out = []
    for i in range(n):
        out.append((PHRASE_FIRST[i % len(PHRASE_FIRST)], PHRASE_SECOND[(i // len(PHRASE_FIRST) + i) % len(PHRASE_SECOND)]))
    return out


def lexer_definitions(nphrases=40, nabbrev=30):
    """-> list of (id, expression, resultIndex, level, posbind); about 100 definitions with the defaults (C2),
    200 with nphrases=140 (C3)."""
    defs = [
        (WORD, r"\b\w+\b", 0, 1, BIND_CONTENT),
        (CAPWORD, r"\b\p{Lu}\p{Ll}*\b", 0, 1, BIND_CONTENT),
        (LOWORD, r"\b\p{Ll}+\b", 0, 1, BIND_CONTENT),
        (NUMBER, r"\b[0-9]+\b", 0, 2, BIND_CONTENT),
        (DECIMAL, r"\b[0-9]+[.,][0-9]+\b", 0, 3, BIND_CONTENT),
        (AMOUNT, r"[$€£][0-9]+\b", 0, 1, BIND_CONTENT),
        (AMOUNT, r"[$€£][0-9]+[.,][0-9]{2}\b", 0, 2, BIND_CONTENT),
        (AMOUNT, r"\b[0-9]+[.,][0-9]{2}\s(?:USD|EUR|CHF|GBP)\b", 0, 3, BIND_CONTENT),
        (DATE, r"\b[0-9]{4}-[0-9]{2}-[0-9]{2}\b", 0, 4, BIND_CONTENT),
        (DATE, r"\b[0-9]{1,2}\.[0-9]{1,2}\.[0-9]{4}\b", 0, 4, BIND_CONTENT),
        (EMAIL, r"\b[a-z0-9._-]+@[a-z0-9-]+(?:\.[a-z0-9-]+)*\.(?:%s)\b" % "|".join(TLDS), 0, 7, BIND_CONTENT),
        (URL, r"\bhttps?://[a-z0-9-]+(?:\.[a-z0-9-]+)*\.(?:%s)(?:/[a-z0-9._/-]*)?\b" % "|".join(TLDS), 0, 6, BIND_CONTENT),
        (URL, r"\bwww\.[a-z0-9-]+(?:\.[a-z0-9-]+)*\.(?:%s)\b" % "|".join(TLDS), 0, 5, BIND_CONTENT),
        (SENT, r"[.]", 0, 2, BIND_PREDECESSOR),
        (PUNCT, r"[,;:!?]", 0, 1, BIND_PREDECESSOR),
        (QUOTE, r"[\"'()\[\]]", 0, 1, BIND_PREDECESSOR),
    ]
    for i, a in enumerate(ABBREVIATIONS[:nabbrev]):
        cap = a[0].upper() + a[1:]
        defs.append((ABBREV, r"\b(?:%s|%s)\." % (a, cap), 0, 3 if i % 2 else 4, BIND_CONTENT))
    for i, (a, b) in enumerate(phrase_list(nphrases)):
        defs.append((PHRASE + i, r"\b%s\s%s\b" % (a, b), 0, 2, BIND_CONTENT))
    return defs


def define_lexer(lx, nphrases=40, nabbrev=30, symbols=()):
    for d in lexer_definitions(nphrases, nabbrev):
        lx.define_lexem(*d)
    for (sid, lid, name) in symbols:
        lx.define_symbol(sid, lid, name)
    lx.compile()
    return lx


# ---------------------------------------------------------------------------------------------------------
# Token pattern sets of the full pipeline configurations (C3 and C4). The generators only call the
# common matcher interface (push_term / attach_variable / push_expression / define_pattern), so the same call
# sequence reaches the product and the reference oracle.
OP_SEQUENCE, OP_SEQUENCE_IMM, OP_SEQUENCE_STRUCT, OP_WITHIN, OP_WITHIN_STRUCT, OP_ANY, OP_AND = range(7)
SYMBOL_BASE = 1000


def _zipf_ranks(rng, n, size, s):
    import numpy as np
    k = np.arange(1, size + 1, dtype=np.float64)
    cum = np.cumsum(1.0 / k ** s)
    return np.searchsorted(cum, rng.random(n) * cum[-1], side="right")


def symbol_table(vocab, nsymbols, lexemid=WORD):
    """(symbolid, lexemid, name) for the most frequent vocabulary words; symbol id = SYMBOL_BASE + rank."""
    out = []
    for rank, w in enumerate(vocab[:nsymbols]):
        out.append((SYMBOL_BASE + rank, lexemid, w))
    return out


def define_c3_patterns(m, seed, npatterns, nsymbols, lexem_share=0.1):
    """C3: 50 % sequence_imm of 2-4 terms (range = argc), 30 % within of 2-3 terms (range 3-10), 10 % *_struct with the
    sentence delimiter, 10 % any of 2-5 terms nested as first argument of a sequence; cardinality 0 except 5 % of
    within/any with cardinality 2; variables on all leaf terms; terms are symbols of vocabulary words (Zipf 0.8) or,
    with probability lexem_share, plain lexem ids. The terms of one pattern are distinct: a pattern that names a term
    twice (within(the, the, x)) binds its variables in the order of the reference's trigger arrays, i.e. depending on
    the history of the document; such rule sets are valid but run on the state machine, not on the window evaluator."""
    import numpy as np
    rng = np.random.default_rng(seed)
    plain = [NUMBER, DECIMAL, DATE, CAPWORD, ABBREV, AMOUNT]

    used = set()

    def term(var):
        while True:
            if rng.random() < lexem_share:
                t = int(plain[int(rng.integers(0, len(plain)))])
            else:
                t = SYMBOL_BASE + int(_zipf_ranks(rng, 1, nsymbols, 0.8)[0])
            if t not in used:
                break
        used.add(t)
        m.push_term(t)
        m.attach_variable("v%d" % var)

    for r in range(npatterns):
        used.clear()
        u = rng.random()
        if u < 0.5:
            n = int(rng.integers(2, 5))
            for i in range(n):
                term(i)
            m.push_expression(OP_SEQUENCE_IMM, n, n, 0)
            name = "seqimm%d" % n
        elif u < 0.8:
            n = int(rng.integers(2, 4))
            for i in range(n):
                term(i)
            m.push_expression(OP_WITHIN, n, int(rng.integers(3, 11)), 2 if rng.random() < 0.05 else 0)
            name = "within%d" % n
        elif u < 0.9:
            n = int(rng.integers(2, 4))
            m.push_term(SENT)
            for i in range(n):
                term(i)
            op = OP_SEQUENCE_STRUCT if rng.random() < 0.5 else OP_WITHIN_STRUCT
            m.push_expression(op, n + 1, int(rng.integers(3, 11)), 0)
            name = "struct%d" % n
        else:
            n = int(rng.integers(2, 6))
            for i in range(n):
                term(i)
            m.push_expression(OP_ANY, n, 1, 2 if rng.random() < 0.05 else 0)
            m.attach_variable("alt")
            term(9)
            m.push_expression(OP_SEQUENCE, 2, int(rng.integers(2, 8)), 0)
            name = "anyseq%d" % n
        m.define_pattern("%s_%d" % (name, r % 97), "", True)


def define_c4_patterns(m, seed, npatterns, vocab_size):
    """C4 (dictionary style): sequence_imm patterns of 2-5 terms drawn Zipf(0.8) from the vocabulary, so that the most
    frequent terms are key events of very many patterns (high trigger fan-out, exercises the alt-key optimisation)."""
    import numpy as np
    rng = np.random.default_rng(seed)
    for r in range(npatterns):
        n = int(rng.integers(2, 6))
        ranks = _zipf_ranks(rng, n, vocab_size, 0.8)
        for i in range(n):
            m.push_term(SYMBOL_BASE + int(ranks[i]))
            m.attach_variable("t%d" % i)
        m.push_expression(OP_SEQUENCE_IMM, n, n, 0)
        m.define_pattern("dict%d" % (r % 1009), "", True)


def c4_token_documents(seed, ndocs, vocab_size, min_tokens=1000, max_tokens=4000):
    """C4 input: token records (one token per ordinal position, ids Zipf(1) over the vocabulary)."""
    import numpy as np
    from .capi import LEXEM_DT
    rng = np.random.default_rng(seed)
    sizes = rng.integers(min_tokens, max_tokens + 1, size=ndocs)
    offs = np.zeros(ndocs + 1, dtype=np.uint64)
    offs[1:] = np.cumsum(sizes)
    total = int(offs[-1])
    toks = np.zeros(total, dtype=LEXEM_DT)
    toks["id"] = SYMBOL_BASE + _zipf_ranks(rng, total, vocab_size, 1.0)
    pos = np.arange(total, dtype=np.uint64) - np.repeat(offs[:-1], sizes)
    toks["ordpos"] = pos + 1
    toks["origpos"] = pos * 6
    toks["origsize"] = 5
    return toks, offs
