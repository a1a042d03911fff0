/* Interface shim (strusAnalyzer analyzer::PatternMatcherStatistics). */
#ifndef _STRUS_ANALYZER_PATTERN_MATCHER_STATISTICS_HPP_INCLUDED
#define _STRUS_ANALYZER_PATTERN_MATCHER_STATISTICS_HPP_INCLUDED
#include <vector>
namespace strus {
namespace analyzer {
class PatternMatcherStatistics
{
public:
	class Item
	{
	public:
		Item( const char* name_, double value_) :m_name(name_),m_value(value_){}
		Item( const Item& o) :m_name(o.m_name),m_value(o.m_value){}
		const char* name() const {return m_name;}
		double value() const {return m_value;}
	private:
		const char* m_name;
		double m_value;
	};
	PatternMatcherStatistics(){}
	PatternMatcherStatistics( const PatternMatcherStatistics& o) :m_items(o.m_items){}
	void define( const char* name, double value) {m_items.push_back( Item( name, value));}
	const std::vector<Item>& items() const {return m_items;}
private:
	std::vector<Item> m_items;
};
}}
#endif
