/* Interface shim (strusAnalyzer PatternMatcherContextInterface), reference implementation:
 * src/patternMatcher.cpp:107-341. */
#ifndef _STRUS_PATTERN_MATCHER_CONTEXT_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_MATCHER_CONTEXT_INTERFACE_HPP_INCLUDED
#include "strus/analyzer/patternLexem.hpp"
#include "strus/analyzer/patternMatcherResult.hpp"
#include "strus/analyzer/patternMatcherStatistics.hpp"
#include <vector>
namespace strus {
class PatternMatcherContextInterface
{
public:
	virtual ~PatternMatcherContextInterface(){}
	virtual void putInput( const analyzer::PatternLexem& token)=0;
	virtual std::vector<analyzer::PatternMatcherResult> fetchResults()=0;
	virtual analyzer::PatternMatcherStatistics getStatistics() const=0;
	virtual void reset()=0;
};
}
#endif
