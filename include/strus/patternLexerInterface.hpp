/* Interface shim (strusAnalyzer PatternLexerInterface), reference: src/patternLexer.hpp:21-38. */
#ifndef _STRUS_PATTERN_LEXER_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_LEXER_INTERFACE_HPP_INCLUDED
#include "strus/structView.hpp"
#include <vector>
#include <string>
namespace strus {
class PatternLexerInstanceInterface;
class PatternLexerInterface
{
public:
	virtual ~PatternLexerInterface(){}
	virtual std::vector<std::string> getCompileOptionNames() const=0;
	virtual PatternLexerInstanceInterface* createInstance() const=0;
	virtual const char* name() const=0;
	virtual StructView view() const=0;
};
}
#endif
