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
