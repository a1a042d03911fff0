/* Interface shim (strusAnalyzer analyzer::PatternLexem): output of the lexer, input of the matcher. */
#ifndef _STRUS_ANALYZER_PATTERN_LEXEM_HPP_INCLUDED
#define _STRUS_ANALYZER_PATTERN_LEXEM_HPP_INCLUDED
#include "strus/analyzer/position.hpp"
namespace strus {
namespace analyzer {
class PatternLexem
{
public:
	PatternLexem( unsigned int id_, int ordpos_, const Position& origpos_, int origsize_)
		:m_id(id_),m_ordpos(ordpos_),m_origpos(origpos_),m_origsize(origsize_){}
	PatternLexem( const PatternLexem& o)
		:m_id(o.m_id),m_ordpos(o.m_ordpos),m_origpos(o.m_origpos),m_origsize(o.m_origsize){}
	unsigned int id() const {return m_id;}
	int ordpos() const {return m_ordpos;}
	const Position& origpos() const {return m_origpos;}
	int origsize() const {return m_origsize;}
	void setOrigPosition( const Position& p) {m_origpos = p;}
private:
	unsigned int m_id;
	int m_ordpos;
	Position m_origpos;
	int m_origsize;
};
}}
#endif
