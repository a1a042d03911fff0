"""Oracle and product share include/sp_unicode_tables.h (general categories, simple case pairs), so a wrong range would be
invisible to every parity test. The table is regenerated here from python's unicodedata with the generator that made it
(tools/gen_unicode_tables.py) and compared byte for byte; a few properties the workloads rely on are checked directly."""
import os
import subprocess
import sys
import unicodedata

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_table_is_what_the_generator_produces(tmp_path):
    out = tmp_path / "sp_unicode_tables.h"
    subprocess.run([sys.executable, os.path.join(ROOT, "tools", "gen_unicode_tables.py"), str(out)], check=True)
    have = open(os.path.join(ROOT, "include", "sp_unicode_tables.h")).read()
    want = open(out).read()
    if ("unicodedata %s " % unicodedata.unidata_version) not in have.splitlines()[0]:
        import pytest
        pytest.skip("the table was generated from another Unicode version than this interpreter has")
    assert have == want


def test_categories_of_the_corpus_blocks():
    import re
    text = open(os.path.join(ROOT, "include", "sp_unicode_tables.h")).read()

    def ranges(cat):
        m = re.search(r"sp_cat_%s\[\] = \{(.*?)\};" % cat, text)
        return [(int(a, 16), int(b, 16)) for a, b in re.findall(r"\{0x([0-9A-F]+),0x([0-9A-F]+)\}", m.group(1))]

    def inside(cp, rs):
        return any(lo <= cp <= hi for lo, hi in rs)

    lu, ll, lo_ = ranges("Lu"), ranges("Ll"), ranges("Lo")
    for cp in list(range(0x41, 0x5B)) + [0xC0, 0xD6, 0x391, 0x3A9, 0x410, 0x42F]:      # A-Z, Latin-1, Greek, Cyrillic capitals
        assert inside(cp, lu) and unicodedata.category(chr(cp)) == "Lu"
    for cp in list(range(0x61, 0x7B)) + [0xDF, 0xE0, 0xFF, 0x3B1, 0x3C9, 0x430, 0x44F]:
        assert inside(cp, ll) and unicodedata.category(chr(cp)) == "Ll"
    for cp in (0x4E00, 0x9FA5):                                                          # CJK unified ideographs
        assert inside(cp, lo_)
    for cat in ("Lu", "Ll", "Lo", "Nd"):
        rs = ranges(cat)
        assert all(a[1] < b[0] for a, b in zip(rs, rs[1:])), "ranges of %s are not sorted and disjoint" % cat
