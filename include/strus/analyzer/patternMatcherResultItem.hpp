/* Interface shim (strusAnalyzer analyzer::PatternMatcherResultItem). */
#ifndef _STRUS_ANALYZER_PATTERN_MATCHER_RESULT_ITEM_HPP_INCLUDED
#define _STRUS_ANALYZER_PATTERN_MATCHER_RESULT_ITEM_HPP_INCLUDED
#include "strus/analyzer/position.hpp"
namespace strus {
namespace analyzer {
class PatternMatcherResultItem
{
public:
	PatternMatcherResultItem( const char* name_, const char* value_, int ordpos_, int ordend_, const Position& origpos_, const Position& origend_)
		:m_name(name_),m_value(value_),m_ordpos(ordpos_),m_ordend(ordend_),m_origpos(origpos_),m_origend(origend_){}
	PatternMatcherResultItem( const PatternMatcherResultItem& o)
		:m_name(o.m_name),m_value(o.m_value),m_ordpos(o.m_ordpos),m_ordend(o.m_ordend),m_origpos(o.m_origpos),m_origend(o.m_origend){}
	const char* name() const {return m_name;}
	const char* value() const {return m_value;}
	int ordpos() const {return m_ordpos;}
	int ordend() const {return m_ordend;}
	const Position& origpos() const {return m_origpos;}
	const Position& origend() const {return m_origend;}
private:
	const char* m_name;
	const char* m_value;
	int m_ordpos;
	int m_ordend;
	Position m_origpos;
	Position m_origend;
};
}}
#endif
