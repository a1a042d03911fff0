/* Interface shim (strusAnalyzer analyzer::PositionBind). Enumerator order as implied by
 * the reference (src/patternLexer.cpp:256, :902-916). */
#ifndef _STRUS_ANALYZER_POSITION_BIND_HPP_INCLUDED
#define _STRUS_ANALYZER_POSITION_BIND_HPP_INCLUDED
namespace strus {
namespace analyzer {
enum PositionBind
{
	BindContent,		///< element gets an own ordinal position
	BindSuccessor,		///< element gets the ordinal position of the succeeding content element
	BindPredecessor,	///< element gets the ordinal position of the preceding content element
	BindUnique		///< like content, but dropped if it directly follows another unique element
};
}}
#endif
