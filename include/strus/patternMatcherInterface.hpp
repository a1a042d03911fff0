/* Interface shim (strusAnalyzer PatternMatcherInterface), reference: src/patternMatcher.hpp:23-39. */
#ifndef _STRUS_PATTERN_MATCHER_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_MATCHER_INTERFACE_HPP_INCLUDED
#include "strus/structView.hpp"
#include <vector>
#include <string>
namespace strus {
class PatternMatcherInstanceInterface;
class PatternMatcherInterface
{
public:
	virtual ~PatternMatcherInterface(){}
	virtual std::vector<std::string> getCompileOptionNames() const=0;
	virtual PatternMatcherInstanceInterface* createInstance() const=0;
	virtual const char* name() const=0;
	virtual StructView view() const=0;
};
}
#endif
