"""B200-native strusPattern hot path: regex lexer (level 1) + token pattern matcher (level 2) behind a C ABI.
See DESIGN.md; capi.py is the ctypes mirror of include/strus_pattern_b200.h."""
from .capi import (BIND_CONTENT, BIND_PREDECESSOR, BIND_SUCCESSOR, BIND_UNIQUE, ITEM_DT, LEXEM_DT, RESULT_DT,
                   STATS_DT, Lexer, LexerContext, Matcher, MatcherContext, PatternError, lib)

__all__ = ["Lexer", "LexerContext", "Matcher", "MatcherContext", "PatternError", "lib", "LEXEM_DT", "RESULT_DT", "ITEM_DT", "STATS_DT",
           "BIND_CONTENT", "BIND_SUCCESSOR", "BIND_PREDECESSOR", "BIND_UNIQUE"]
