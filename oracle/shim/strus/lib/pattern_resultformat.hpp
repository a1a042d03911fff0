/* Oracle build shim (strusAnalyzer lib/pattern_resultformat.hpp). Only what
 * src/patternMatcher.cpp:45-65, 79, 174-181, 253-262, 562-566 touches. Format strings are not
 * exercised by the oracle (formatstring == "" on every parity case), so creating one throws. */
#ifndef _STRUS_PATTERN_RESULTFORMAT_HPP_INCLUDED
#define _STRUS_PATTERN_RESULTFORMAT_HPP_INCLUDED
#include "strus/analyzer/patternMatcherResultItem.hpp"
#include <string>
#include <stdexcept>
namespace strus {
class ErrorBufferInterface;
struct PatternResultFormat {};
class PatternResultFormatVariableMap
{
public:
	virtual ~PatternResultFormatVariableMap(){}
	virtual const char* getVariable( const std::string& name) const=0;
};
class PatternResultFormatTable
{
public:
	PatternResultFormatTable( const PatternResultFormatVariableMap*, ErrorBufferInterface*){}
	const PatternResultFormat* createResultFormat( const char*)
		{throw std::runtime_error( "result format strings are not part of the oracle build");}
};
class PatternResultFormatContext
{
public:
	explicit PatternResultFormatContext( ErrorBufferInterface*){}
	const char* map( const PatternResultFormat*, const analyzer::PatternMatcherResultItem*, std::size_t)
		{return 0;}
};
}
#endif
