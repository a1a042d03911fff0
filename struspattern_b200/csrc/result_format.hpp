// Result format strings: the value of a result or of a result item defined by a format string
// in definePattern (src/patternMatcher.cpp:545, :560-566).
//
// The formatter itself (PatternResultFormatTable / PatternResultFormatContext, strusAnalyzer 0.17
// lib/pattern_resultformat) is NOT part of the reference tree (un-vendored dependency; parity unpinned against an
// executed strusAnalyzer). It is restated here from the reference's call sites and the published behaviour:
//   - a format is created once per definePattern call; variables it names must be known to the variable map at that
//     time (VariableMap::getVariable, src/patternMatcher.cpp:54-59), an unknown variable is an error;
//   - the value of a result is map( format, items gathered for the result) and the result gets NO items then
//     (:253-262); the value of an item that refers to a sub expression with a format is map( format, items of that sub
//     expression), and those items do not join the list of the parent (:172-181, :185-188);
//   - "{currency} {value}" on "3'645 Eur" with Currency = CURRENCY_EUR ["EUR"] gives "EUR 3'645"
//     (doc/webpage/introduction_struspattern.htm:143-161): text outside braces is copied, {name} stands for the items of
//     that variable in list order, an item with a value of its own contributes that value, any other item the piece of
//     the source it spans.
// Syntax restated: '{' name [ '|' separator ] '}' (separator between several items of one variable, default one
// space), '\' takes the next character literally. The matcher never sees the source text, so a span is written into the
// value as a reference  \x01 seg ',' pos ',' endseg ',' endpos \x02  (decimal) that the caller resolves against its
// segments (sp_resolve_value in include/strus_pattern_b200.h, struspattern_b200.capi.resolve_value).
#ifndef SP_RESULT_FORMAT_HPP_INCLUDED
#define SP_RESULT_FORMAT_HPP_INCLUDED
#include "strus_pattern_b200.h"
#include <string>
#include <vector>
#include <stdexcept>
#include <cstdio>

namespace sp {

struct ResultFormatElement
{
	uint32_t variable;		// 0 = literal text
	std::string text;		// literal, or separator of a variable
};
typedef std::vector<ResultFormatElement> ResultFormat;

// VariableLookup: uint32_t operator()( const std::string& name) -> handle, 0 = unknown
template <class VariableLookup>
inline ResultFormat parseResultFormat( const std::string& src, const VariableLookup& lookup)
{
	ResultFormat rt;
	std::string literal;
	for (std::size_t i=0; i<src.size(); ++i)
	{
		const char ch = src[ i];
		if (ch == '\\' && i + 1 < src.size()) {literal.push_back( src[ ++i]); continue;}
		if (ch == '}') throw std::runtime_error( "unbalanced '}' in result format string");
		if (ch != '{') {literal.push_back( ch); continue;}
		std::string name, separator( " ");
		bool insep = false, closed = false;
		for (++i; i<src.size(); ++i)
		{
			char c = src[ i];
			if (c == '\\' && i + 1 < src.size()) c = src[ ++i];
			else if (c == '}') {closed = true; break;}
			else if (c == '{') throw std::runtime_error( "nested '{' in result format string");
			else if (c == '|' && !insep) {insep = true; separator.clear(); continue;}
			(insep ? separator : name).push_back( c);
		}
		if (!closed) throw std::runtime_error( "missing '}' in result format string");
		while (!name.empty() && name[ name.size()-1] == ' ') name.erase( name.size()-1);
		while (!name.empty() && name[ 0] == ' ') name.erase( 0, 1);
		if (name.empty()) throw std::runtime_error( "empty variable name in result format string");
		const uint32_t handle = lookup( name);
		if (!handle) throw std::runtime_error( "unknown variable '" + name + "' in result format string");
		if (!literal.empty()) {ResultFormatElement e; e.variable = 0; e.text.swap( literal); rt.push_back( e);}
		ResultFormatElement e; e.variable = handle; e.text = separator;
		rt.push_back( e);
	}
	if (!literal.empty()) {ResultFormatElement e; e.variable = 0; e.text = literal; rt.push_back( e);}
	return rt;
}

inline void appendSpanReference( std::string& out, const sp_item& it)
{
	char buf[ 64];
	std::snprintf( buf, sizeof(buf), "\x01%u,%u,%u,%u\x02", (unsigned int)it.origseg, (unsigned int)it.origpos, (unsigned int)it.endseg, (unsigned int)it.endpos);
	out.append( buf);
}

// map( format `handle`, the item records rec[0..n) ): nested records (SP_ITEM_NESTED) belong to the item in front of them.
inline void mapResultFormat( const std::vector<ResultFormat>& formats, uint32_t handle, const sp_item* rec, uint32_t n, std::string& out, int depth=0)
{
	if (handle == 0 || handle > formats.size()) throw std::runtime_error( "unknown result format handle");
	if (depth > 64) throw std::runtime_error( "result formats nested too deep");
	for (const ResultFormatElement& e: formats[ handle-1])
	{
		if (!e.variable) {out.append( e.text); continue;}
		uint32_t cnt = 0;
		for (uint32_t i=0; i<n; i += 1 + SP_ITEM_NESTED( rec[ i]))
		{
			if (i + SP_ITEM_NESTED( rec[ i]) >= n) throw std::runtime_error( "corrupt item list: nested records beyond its end");
			if (rec[ i].variable != e.variable) continue;
			if (cnt++) out.append( e.text);
			if (SP_ITEM_FORMAT( rec[ i])) mapResultFormat( formats, SP_ITEM_FORMAT( rec[ i]), rec + i + 1, SP_ITEM_NESTED( rec[ i]), out, depth + 1);
			else appendSpanReference( out, rec[ i]);
		}
	}
}

// Replaces the span references of segment 0 by the bytes of `source`; references to other segments stay as they are.
inline void resolveValue( const char* value, std::size_t len, const uint8_t* source, std::size_t sourcelen, std::string& out)
{
	for (std::size_t i=0; i<len; ++i)
	{
		if (value[ i] != '\x01') {out.push_back( value[ i]); continue;}
		std::size_t e = i + 1;
		unsigned long num[ 4] = {0,0,0,0}; int k = 0;
		bool ok = true;
		for (; e<len && value[ e] != '\x02'; ++e)
		{
			if (value[ e] == ',') {if (++k > 3) {ok = false; break;}}
			else if (value[ e] >= '0' && value[ e] <= '9') num[ k] = num[ k] * 10 + (unsigned long)(value[ e] - '0');
			else {ok = false; break;}
		}
		if (!ok || e >= len || k != 3) throw std::runtime_error( "corrupt span reference in value string");
		if (num[ 0] == 0 && num[ 2] == 0)
		{
			if (num[ 1] > num[ 3] || num[ 3] > sourcelen) throw std::runtime_error( "span reference beyond the end of the source");
			out.append( reinterpret_cast<const char*>( source) + num[ 1], num[ 3] - num[ 1]);
		}
		else out.append( value + i, e + 1 - i);
		i = e;
	}
}

} // namespace sp
#endif
