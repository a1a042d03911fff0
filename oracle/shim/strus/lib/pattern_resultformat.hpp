/* Oracle build shim (strusAnalyzer lib/pattern_resultformat.hpp) for what src/patternMatcher.cpp:45-65, 79, 174-181,
 * 253-262, 562-566 touches. TEST INFRASTRUCTURE: only the oracle build (oracle/Makefile) compiles this.
 *
 * The real formatter is strusAnalyzer 0.17 code that is not part of the reference tree: parity of value strings is
 * UNPINNED against an executed strusAnalyzer. What the reference itself fixes - which items reach the formatter, which
 * results and items get a value, which items disappear from the result - is the reference's own code compiled around
 * this shim. The formatter is restated from the published behaviour (doc/webpage/introduction_struspattern.htm:143-161:
 * text is copied, {variable} stands for the items of that name, an item with a value contributes the value, any other
 * item the piece of source it spans) and written independently of the product's csrc/result_format.hpp:
 *   format  := ( text | '{' name [ '|' separator ] '}' )*      '\' quotes the next character, default separator " "
 *   a span  := "\x01" seg "," pos "," endseg "," endpos "\x02"  (the matcher never sees the source text)            */
#ifndef _STRUS_PATTERN_RESULTFORMAT_HPP_INCLUDED
#define _STRUS_PATTERN_RESULTFORMAT_HPP_INCLUDED
#include "strus/analyzer/patternMatcherResultItem.hpp"
#include <string>
#include <list>
#include <vector>
#include <cstring>
#include <cstdio>
#include <stdexcept>
namespace strus {
class ErrorBufferInterface;

class PatternResultFormatVariableMap
{
public:
	virtual ~PatternResultFormatVariableMap(){}
	virtual const char* getVariable( const std::string& name) const=0;
};

struct PatternResultFormat
{
	struct Part
	{
		const char* variable;	/* NULL: `text` is copied; else the key of the variable, `text` its separator */
		std::string text;
	};
	std::vector<Part> parts;
};

class PatternResultFormatTable
{
public:
	PatternResultFormatTable( const PatternResultFormatVariableMap* variableMap_, ErrorBufferInterface*)
		:m_variableMap(variableMap_){}

	const PatternResultFormat* createResultFormat( const char* src)
	{
		PatternResultFormat fmt;
		PatternResultFormat::Part cur; cur.variable = 0;
		char const* si = src;
		while (*si)
		{
			if (*si == '\\' && si[1]) {cur.text.push_back( si[1]); si += 2;}
			else if (*si == '{')
			{
				if (!cur.text.empty()) {fmt.parts.push_back( cur); cur.text.clear();}
				std::string field[ 2]; int fi = 0;
				for (++si; *si && *si != '}'; ++si)
				{
					if (*si == '\\' && si[1]) {field[ fi].push_back( *++si);}
					else if (*si == '{') throw std::runtime_error( "nested '{' in result format string");
					else if (*si == '|' && fi == 0) fi = 1;
					else field[ fi].push_back( *si);
				}
				if (*si != '}') throw std::runtime_error( "missing '}' in result format string");
				++si;
				std::string::size_type b = field[0].find_first_not_of( ' '), e = field[0].find_last_not_of( ' ');
				if (b == std::string::npos) throw std::runtime_error( "empty variable name in result format string");
				std::string name( field[0], b, e - b + 1);
				PatternResultFormat::Part var;
				var.variable = m_variableMap->getVariable( name);
				if (!var.variable) throw std::runtime_error( "unknown variable '" + name + "' in result format string");
				var.text = fi ? field[1] : std::string( " ");
				fmt.parts.push_back( var);
			}
			else if (*si == '}') throw std::runtime_error( "unbalanced '}' in result format string");
			else cur.text.push_back( *si++);
		}
		if (!cur.text.empty()) fmt.parts.push_back( cur);
		m_formats.push_back( fmt);
		return &m_formats.back();
	}
private:
	const PatternResultFormatVariableMap* m_variableMap;
	std::list<PatternResultFormat> m_formats;
};

class PatternResultFormatContext
{
public:
	explicit PatternResultFormatContext( ErrorBufferInterface*){}

	/* the returned string lives as long as the context */
	const char* map( const PatternResultFormat* fmt, const analyzer::PatternMatcherResultItem* items, std::size_t nofItems)
	{
		std::string out;
		std::vector<PatternResultFormat::Part>::const_iterator pi = fmt->parts.begin(), pe = fmt->parts.end();
		for (; pi != pe; ++pi)
		{
			if (!pi->variable) {out.append( pi->text); continue;}
			bool first = true;
			for (std::size_t ii=0; ii<nofItems; ++ii)
			{
				if (!items[ ii].name() || 0!=std::strcmp( items[ ii].name(), pi->variable)) continue;
				if (!first) out.append( pi->text);
				first = false;
				if (items[ ii].value()) out.append( items[ ii].value());
				else
				{
					char ref[ 80];	/* (no iostreams here: number output through the locale facets of a statically linked libstdc++ crashed in threads) */
					std::snprintf( ref, sizeof(ref), "\x01%d,%d,%d,%d\x02", (int)items[ ii].origpos().seg(), (int)items[ ii].origpos().ofs(),
							(int)items[ ii].origend().seg(), (int)items[ ii].origend().ofs());
					out.append( ref);
				}
			}
		}
		m_values.push_back( out);
		return m_values.back().c_str();
	}
private:
	std::list<std::string> m_values;
};
}
#endif
