"""The strus C++ interface classes (struspattern_b200/csrc/strus_iface.cpp) driven by a C++ program that reads like
the reference's own tests (tests/cpp/test_strus_iface.cpp): compile + link check on the CPU, run on the GPU."""
import json
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
BUILD = os.path.join(ROOT, "tests", "hostsim", "_build")
EXE = os.path.join(BUILD, "test_strus_iface")
OPS = {"OpSequence": 0, "OpSequenceImm": 1, "OpSequenceStruct": 2, "OpWithin": 3, "OpWithinStruct": 4, "OpAny": 5, "OpAnd": 6}


def build_exe():
    from struspattern_b200 import build, capi
    build.build()
    src = os.path.join(ROOT, "tests", "cpp", "test_strus_iface.cpp")
    if os.path.exists(EXE) and os.path.getmtime(EXE) >= max(os.path.getmtime(src), os.path.getmtime(capi.LIB_PATH)):
        return
    os.makedirs(BUILD, exist_ok=True)
    libdir = os.path.dirname(capi.LIB_PATH)
    subprocess.run(["g++", "-O1", "-std=c++17", "-w", "-I", os.path.join(ROOT, "include"), "-o", EXE, src,
                    "-L", libdir, "-lstrus_pattern_b200", "-ldl", "-Wl,-rpath," + libdir, "-Wl,-rpath-link," + libdir,
                    "-L/usr/local/cuda/lib64", "-Wl,-rpath-link,/usr/local/cuda/lib64"], check=True)


def write_casefile(path):
    from test_oracle_matcher import simple_document
    lines = []
    case = json.load(open(os.path.join(GOLDEN, "charRegexMatch.json")))["cases"][0]
    lines += ["LEXER", "O DOTALL"]
    for p in case["patterns"]:
        lines.append("L %d %d %d %d %s" % (p["id"], p["resultIndex"], p["level"], 0 if p["haspos"] else 2, p["expression_latin1"].encode("latin-1").hex()))
    for s in case["symbols"]:
        lines.append("S %d %d %s" % (s["id"], s["patternid"], s["name"].encode().hex()))
    lines.append("T " + case["src_latin1"].encode("latin-1").hex())
    lines += ["E %d %d %d %d" % tuple(r) for r in case["results"]]
    lines.append("END")
    kat = json.load(open(os.path.join(GOLDEN, "simpleTokenPatternMatch.json")))
    lines.append("MATCHER")
    for p in kat["patterns"]:
        for termid, var in p["terms"]:
            lines.append("t %d %s" % (termid, ("A%u" % var) if var else "-"))
        lines.append("x %d %d %d %d" % (OPS[p["op"]], p["argc"], p["range"], p["cardinality"]))
        lines.append("p %s 1" % p["name"])
        lines += ["e %s %d" % (p["name"], pos) for pos in p["results"]]
    for t in simple_document():
        lines.append("d %d %d %d %d" % (t["id"], t["ordpos"], t["origpos"], t["origsize"]))
    lines.append("END")
    lines.append("FORMATS")
    open(path, "w").write("\n".join(lines) + "\n")


def test_interface_program_compiles_and_links():
    build_exe()
    assert os.path.exists(EXE)


@pytest.mark.gpu
def test_interface_program_runs_like_the_reference_tests(tmp_path):
    from struspattern_b200 import capi
    build_exe()
    casefile = str(tmp_path / "cases.txt")
    write_casefile(casefile)
    res = subprocess.run([EXE, casefile, capi.LIB_PATH], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=300)
    assert res.returncode == 0, res.stdout
    assert "lexer case: 36 lexems OK" in res.stdout and "matcher case: 5 results OK" in res.stdout and "module entry point OK" in res.stdout
    assert "format case: 6 results OK" in res.stdout, res.stdout
