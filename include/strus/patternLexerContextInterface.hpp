/* Interface shim (strusAnalyzer PatternLexerContextInterface), reference implementation:
 * src/patternLexer.cpp:681-959. */
#ifndef _STRUS_PATTERN_LEXER_CONTEXT_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_LEXER_CONTEXT_INTERFACE_HPP_INCLUDED
#include "strus/analyzer/patternLexem.hpp"
#include <vector>
#include <cstddef>
namespace strus {
class PatternLexerContextInterface
{
public:
	virtual ~PatternLexerContextInterface(){}
	virtual std::vector<analyzer::PatternLexem> match( const char* src, std::size_t srclen)=0;
	virtual void reset()=0;
};
}
#endif
