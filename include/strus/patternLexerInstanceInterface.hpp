/* Interface shim (strusAnalyzer PatternLexerInstanceInterface), reference implementation:
 * src/patternLexer.cpp:961-1151. */
#ifndef _STRUS_PATTERN_LEXER_INSTANCE_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_LEXER_INSTANCE_INTERFACE_HPP_INCLUDED
#include "strus/analyzer/positionBind.hpp"
#include "strus/structView.hpp"
#include <string>
namespace strus {
class PatternLexerContextInterface;
class PatternLexerInstanceInterface
{
public:
	virtual ~PatternLexerInstanceInterface(){}
	virtual void defineOption( const std::string& name, double value)=0;
	virtual void defineLexemName( unsigned int id, const std::string& name)=0;
	virtual void defineLexem( unsigned int id, const std::string& expression, unsigned int resultIndex, unsigned int level, analyzer::PositionBind posbind)=0;
	virtual void defineSymbol( unsigned int id, unsigned int lexemid, const std::string& name)=0;
	virtual unsigned int getSymbol( unsigned int lexemid, const std::string& name) const=0;
	virtual const char* getLexemName( unsigned int id) const=0;
	virtual bool compile()=0;
	virtual PatternLexerContextInterface* createContext() const=0;
	virtual const char* name() const=0;
	virtual StructView view() const=0;
};
}
#endif
