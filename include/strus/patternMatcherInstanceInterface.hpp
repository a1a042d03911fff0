/* Interface shim (strusAnalyzer PatternMatcherInstanceInterface), reference implementation:
 * src/patternMatcher.cpp:345-704. JoinOperation enumerators by NAME as used at
 * src/patternMatcher.cpp:403-439. */
#ifndef _STRUS_PATTERN_MATCHER_INSTANCE_INTERFACE_HPP_INCLUDED
#define _STRUS_PATTERN_MATCHER_INSTANCE_INTERFACE_HPP_INCLUDED
#include "strus/structView.hpp"
#include <string>
#include <cstddef>
namespace strus {
class PatternMatcherContextInterface;
class PatternMatcherInstanceInterface
{
public:
	enum JoinOperation
	{
		OpSequence,		///< arguments in ascending position order
		OpSequenceImm,		///< arguments immediately following each other
		OpSequenceStruct,	///< sequence not crossing the delimiter given as first argument
		OpWithin,		///< arguments in any order within a proximity range
		OpWithinStruct,		///< within, not crossing the delimiter given as first argument
		OpAny,			///< any of the arguments
		OpAnd			///< all arguments at the same ordinal position
	};
	virtual ~PatternMatcherInstanceInterface(){}
	virtual void defineOption( const std::string& name, double value)=0;
	virtual void defineTermFrequency( unsigned int termid, double df)=0;
	virtual void pushTerm( unsigned int termid)=0;
	virtual void pushExpression( JoinOperation operation, std::size_t argc, unsigned int range, unsigned int cardinality)=0;
	virtual void pushPattern( const std::string& name)=0;
	virtual void attachVariable( const std::string& name)=0;
	virtual void definePattern( const std::string& name, const std::string& formatstring, bool visible)=0;
	virtual bool compile()=0;
	virtual PatternMatcherContextInterface* createContext() const=0;
	virtual const char* name() const=0;
	virtual StructView view() const=0;
};
}
#endif
