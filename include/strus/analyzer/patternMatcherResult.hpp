/* Interface shim (strusAnalyzer analyzer::PatternMatcherResult). */
#ifndef _STRUS_ANALYZER_PATTERN_MATCHER_RESULT_HPP_INCLUDED
#define _STRUS_ANALYZER_PATTERN_MATCHER_RESULT_HPP_INCLUDED
#include "strus/analyzer/patternMatcherResultItem.hpp"
#include <vector>
namespace strus {
namespace analyzer {
class PatternMatcherResult
{
public:
	typedef PatternMatcherResultItem Item;
	PatternMatcherResult( const char* name_, const char* value_, int ordpos_, int ordend_, const Position& origpos_, const Position& origend_, const std::vector<Item>& itemlist_=std::vector<Item>())
		:m_name(name_),m_value(value_),m_ordpos(ordpos_),m_ordend(ordend_),m_origpos(origpos_),m_origend(origend_),m_items(itemlist_){}
	PatternMatcherResult( const PatternMatcherResult& o)
		:m_name(o.m_name),m_value(o.m_value),m_ordpos(o.m_ordpos),m_ordend(o.m_ordend),m_origpos(o.m_origpos),m_origend(o.m_origend),m_items(o.m_items){}
	const char* name() const {return m_name;}
	const char* value() const {return m_value;}
	int ordpos() const {return m_ordpos;}
	int ordend() const {return m_ordend;}
	const Position& origpos() const {return m_origpos;}
	const Position& origend() const {return m_origend;}
	const std::vector<Item>& items() const {return m_items;}
private:
	const char* m_name;
	const char* m_value;
	int m_ordpos;
	int m_ordend;
	Position m_origpos;
	Position m_origend;
	std::vector<Item> m_items;
};
}}
#endif
