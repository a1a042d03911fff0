"""Shared lexem sets and seeded text generators for the lexer parity tests (CPU host simulation and GPU)."""
import numpy as np

BIND_CONTENT, BIND_SUCCESSOR, BIND_PREDECESSOR, BIND_UNIQUE = 0, 1, 2, 3

# (id, expression, resultIndex, level, posbind)
LEXEM_SETS = {
    "words_numbers": [
        (1, r"\b\w+\b", 0, 1, BIND_CONTENT),
        (2, r"\b[0-9]+\b", 0, 2, BIND_CONTENT),
        (3, r"\b[0-9]+[.,][0-9]+\b", 0, 3, BIND_CONTENT),
        (4, r"[.]", 0, 2, BIND_PREDECESSOR),
        (5, r"[,;:!?]", 0, 1, BIND_PREDECESSOR),
    ],
    "unicode_case": [
        (1, r"\b\w+\b", 0, 1, BIND_CONTENT),
        (2, r"\p{Lu}\p{Ll}*", 0, 1, BIND_CONTENT),
        (3, r"\p{Ll}+", 0, 1, BIND_CONTENT),
        (4, r"[^\s\w]", 0, 0, BIND_PREDECESSOR),
        (5, r"\p{Lo}+", 0, 2, BIND_CONTENT),
    ],
    "no_boundaries": [
        (1, r"[a-z]+", 0, 0, BIND_CONTENT),
        (2, r"[a-z]+ [a-z]+", 0, 1, BIND_CONTENT),
        (3, r"(?:ab|ba)+", 0, 2, BIND_CONTENT),
        (4, r"a{2,3}", 0, 3, BIND_UNIQUE),
        (5, r"#", 0, 0, BIND_SUCCESSOR),
    ],
    "phrases_dates": [
        (1, r"\b\w+\b", 0, 1, BIND_CONTENT),
        (2, r"\bnew\syork\b", 0, 2, BIND_CONTENT),
        (3, r"\b[0-9]{4}-[0-9]{2}-[0-9]{2}\b", 0, 4, BIND_CONTENT),
        (4, r"\b[0-9]{1,2}\.[0-9]{1,2}\.[0-9]{4}\b", 0, 4, BIND_CONTENT),
        (5, r"\b[a-z0-9._-]+@[a-z0-9.-]+\.[a-z]{2,4}\b", 0, 7, BIND_CONTENT),
        (6, r"\bhttps?://[a-z0-9./_-]+\b", 0, 5, BIND_CONTENT),
        (7, r"\b([0-9]+)th\b", 1, 3, BIND_CONTENT),
        (8, r"[.]", 0, 2, BIND_PREDECESSOR),
        (9, r"\b(?:mr|dr|prof)\.", 0, 3, BIND_CONTENT),
    ],
    "lines": [
        (1, r"^[a-z]+", 0, 1, BIND_CONTENT),
        (2, r"[a-z]+$", 0, 1, BIND_CONTENT),
        (3, r"\Bb+\B", 0, 2, BIND_CONTENT),
        (4, r"\A.", 0, 3, BIND_CONTENT),
        (5, r".\z", 0, 3, BIND_CONTENT),
        (6, r"[0-9]+", 0, 0, BIND_CONTENT),
    ],
}
OPTIONS = {"lines": ["MULTILINE"], "phrases_dates": ["CASELESS"]}
SYMBOLS = {
    "words_numbers": [(101, 1, "the"), (102, 1, "new"), (103, 1, "york"), (104, 2, "42")],
    "phrases_dates": [(201, 1, "a"), (202, 1, "b"), (203, 2, "new york")],
}

_ASCII_WORDS = ["the", "new", "york", "a", "b", "ab", "ba", "abba", "aaa", "aa", "quick", "Brown", "fox", "Mr", "dr",
                "prof", "x9", "hello_world", "NEW", "York", "th", "zebra", "b", "bb", "abb"]
_UNI_WORDS = ["größe", "Ärger", "naïve", "Ωmega", "αβγ", "Привет", "мир", "日本語", "中文", "été", "Łódź", "😀"]
_NUMBERS = ["42", "3.14", "1,5", "2024-01-31", "7.11.1999", "12th", "100th", "0", "99999"]
_SPECIAL = ["me@example.com", "a.b-c@d.org", "http://x.y/z_1", "https://strus.net/a-b", ".", ",", ";", "!", "?", "#", "mr.",
            "Dr.", "prof."]


def make_text(rng, nbytes, unicode_share=0.2, newline_every=40):
    """Random but token-structured valid UTF-8 text of about `nbytes` bytes."""
    out = []
    size = 0
    col = 0
    while size < nbytes:
        r = rng.random()
        if r < 0.55:
            w = _ASCII_WORDS[rng.integers(len(_ASCII_WORDS))]
        elif r < 0.55 + unicode_share:
            w = _UNI_WORDS[rng.integers(len(_UNI_WORDS))]
        elif r < 0.85:
            w = _NUMBERS[rng.integers(len(_NUMBERS))]
        else:
            w = _SPECIAL[rng.integers(len(_SPECIAL))]
        sep = " " if rng.random() < 0.9 else ("" if rng.random() < 0.5 else "  ")
        b = (w + sep).encode()
        col += len(b)
        if col > newline_every:
            b += b"\n"
            col = 0
        out.append(b)
        size += len(b)
    return b"".join(out)


def make_batch(seed, ndocs, mean_bytes, unicode_share=0.2, with_empty=True):
    """-> (uint8 array, uint64 offsets); document lengths vary widely and a few are empty."""
    rng = np.random.default_rng(seed)
    docs = []
    for d in range(ndocs):
        if with_empty and rng.random() < 0.1:
            docs.append(b"")
            continue
        n = int(rng.integers(1, 2 * mean_bytes))
        docs.append(make_text(rng, n, unicode_share))
    offs = np.zeros(ndocs + 1, dtype=np.uint64)
    offs[1:] = np.cumsum([len(d) for d in docs])
    return np.frombuffer(b"".join(docs), dtype=np.uint8).copy(), offs


def define_set(lx, name, with_symbols=True):
    for opt in OPTIONS.get(name, []):
        lx.define_option(opt)
    for (id_, expr, ri, level, bind) in LEXEM_SETS[name]:
        lx.define_lexem(id_, expr, ri, level, bind)
    if with_symbols:
        for (sid, lid, sname) in SYMBOLS.get(name, []):
            lx.define_symbol(sid, lid, sname)
    lx.compile()
    return lx
