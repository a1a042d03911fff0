/* Interface shim (strusAnalyzer analyzer::Position): segment + byte offset in the segment. */
#ifndef _STRUS_ANALYZER_POSITION_HPP_INCLUDED
#define _STRUS_ANALYZER_POSITION_HPP_INCLUDED
namespace strus {
namespace analyzer {
class Position
{
public:
	Position() :m_seg(0),m_ofs(0){}
	Position( int seg_, int ofs_) :m_seg(seg_),m_ofs(ofs_){}
	Position( const Position& o) :m_seg(o.m_seg),m_ofs(o.m_ofs){}
	int seg() const {return m_seg;}
	int ofs() const {return m_ofs;}
	void setSeg( int seg_) {m_seg = seg_;}
	void setOfs( int ofs_) {m_ofs = ofs_;}
private:
	int m_seg;
	int m_ofs;
};
}}
#endif
