/* Exported factory functions of libstrus_pattern (same signatures as the reference's
 * include/strus/lib/pattern.hpp:27-32; implemented in struspattern_b200/csrc/strus_iface.cpp). */
#ifndef _STRUS_PATTERN_LIB_HPP_INCLUDED
#define _STRUS_PATTERN_LIB_HPP_INCLUDED
#include <cstdio>
namespace strus {
class PatternLexerInterface;
class PatternMatcherInterface;
class ErrorBufferInterface;
PatternLexerInterface* createPatternLexer_std( ErrorBufferInterface* errorhnd);
PatternMatcherInterface* createPatternMatcher_std( ErrorBufferInterface* errorhnd);
}
#endif
