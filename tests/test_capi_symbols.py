"""The C-ABI library loads on a CPU-only box and exports every entry point include/strus_pattern_b200.h declares
(no compute calls here: there is no CPU path behind the ABI)."""
import ctypes
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    src = open(os.path.join(ROOT, "include", "strus_pattern_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sp_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from struspattern_b200 import build, capi
    build.build()
    L = ctypes.CDLL(capi.LIB_PATH)
    names = declared_symbols()
    assert len(names) > 40
    missing = [n for n in names if not hasattr(L, n)]
    assert not missing, "declared in the header but not exported: %s" % missing


def test_definition_calls_work_without_a_device():
    from struspattern_b200 import Matcher, PatternError
    m = Matcher()
    m.push_term(1); m.attach_variable("a")
    m.push_term(2)
    m.push_expression(0, 2, 3, 0)
    m.define_pattern("seq", "", True)
    m.compile()
    assert m.table_info("programs") == 1 and m.table_info("trigger_defs") == 2
    assert m.result_name(1) == "seq" and m.item_name(1) == "a"
    try:
        m.push_expression(0, 5, 1, 0)
        assert False, "expected an error"
    except PatternError as e:
        assert e.code == -1


def test_reference_artefact_names_load_and_export_the_strus_symbols():
    """libstrus_pattern.so.0.17 and modstrus_analyzer_pattern.so (reference src/CMakeLists.txt:48-57) are build targets: they
    load under those names and export the two factory functions (src/libstrus_pattern.cpp:21-47) and the module entry point
    (src/modstrus_analyzer_pattern.cpp:59-61)."""
    from struspattern_b200 import build
    build.build()
    for name in build.ALIASES:
        path = os.path.join(build.PKG, name)
        assert os.path.exists(path), name
        L = ctypes.CDLL(path)
        for sym in ("entryPoint", "_ZN5strus22createPatternLexer_stdEPNS_20ErrorBufferInterfaceE",
                    "_ZN5strus24createPatternMatcher_stdEPNS_20ErrorBufferInterfaceE"):
            assert hasattr(L, sym), "%s lacks %s" % (name, sym)
