// Regex front end of the B200 lexer compiler: expression string -> syntax tree over code point sets.
//
// Replaces the parsing half of hs_compile_ext_multi (reference call site src/patternLexer.cpp:1083-1086);
// the dialect is the PCRE subset Hyperscan accepts as used by strus lexem definitions:
// literals, classes, \d \w \s \h \v (+ negations), \p{..} general categories, '.', alternation, groups,
// inline flags (?ism), greedy/lazy quantifiers (laziness is irrelevant under all-matches semantics),
// assertions \b \B ^ $ \A \z. Rejected like Hyperscan does: back references, look-around, atomic groups,
// possessive quantifiers.
#ifndef SP_REGEX_HPP_INCLUDED
#define SP_REGEX_HPP_INCLUDED
#include <stdint.h>
#include <string>
#include <vector>
#include <memory>
#include <stdexcept>
#include <algorithm>
#include <cstring>
#include <cstdio>
#include "sp_unicode_tables.h"

namespace sp {

enum RegexFlags {RX_CASELESS=1, RX_DOTALL=2, RX_MULTILINE=4, RX_ALLOWEMPTY=8, RX_UCP=16};
static const uint32_t kMaxCodePoint = 0x10FFFF;

struct SyntaxError :public std::runtime_error
{
	explicit SyntaxError( const std::string& msg) :std::runtime_error(msg){}
};
struct UnsupportedError :public std::runtime_error
{
	explicit UnsupportedError( const std::string& msg) :std::runtime_error(msg){}
};

// ---- set of code points as sorted disjoint intervals
class CodeSet
{
public:
	struct Iv {uint32_t lo, hi;};

	CodeSet(){}
	static CodeSet single( uint32_t c) {CodeSet s; s.insert( c, c); return s;}
	static CodeSet all() {CodeSet s; s.insert( 0, kMaxCodePoint); return s;}

	void insert( uint32_t lo, uint32_t hi) {if (lo <= hi) {Iv v = {lo,hi}; m_iv.push_back( v); m_clean = false;}}
	void insert( const CodeSet& o) {m_iv.insert( m_iv.end(), o.m_iv.begin(), o.m_iv.end()); m_clean = false;}
	const std::vector<Iv>& intervals() const {tidy(); return m_iv;}
	bool empty() const {return m_iv.empty();}

	bool has( uint32_t c) const
	{
		tidy();
		std::size_t a = 0, b = m_iv.size();
		while (a < b)
		{
			std::size_t m = (a+b) >> 1;
			if (c < m_iv[m].lo) b = m; else if (c > m_iv[m].hi) a = m+1; else return true;
		}
		return false;
	}
	CodeSet complement() const
	{
		tidy();
		CodeSet rt;
		uint32_t from = 0;
		for (std::size_t i=0; i<m_iv.size(); ++i)
		{
			if (m_iv[i].lo > from) rt.insert( from, m_iv[i].lo-1);
			from = m_iv[i].hi + 1;
		}
		if (from <= kMaxCodePoint) rt.insert( from, kMaxCodePoint);
		return rt;
	}
	bool operator<( const CodeSet& o) const
	{
		tidy(); o.tidy();
		if (m_iv.size() != o.m_iv.size()) return m_iv.size() < o.m_iv.size();
		for (std::size_t i=0; i<m_iv.size(); ++i)
		{
			if (m_iv[i].lo != o.m_iv[i].lo) return m_iv[i].lo < o.m_iv[i].lo;
			if (m_iv[i].hi != o.m_iv[i].hi) return m_iv[i].hi < o.m_iv[i].hi;
		}
		return false;
	}
private:
	void tidy() const
	{
		if (m_clean) return;
		std::sort( m_iv.begin(), m_iv.end(), []( const Iv& a, const Iv& b) {return a.lo != b.lo ? a.lo < b.lo : a.hi < b.hi;});
		std::vector<Iv> out;
		for (std::size_t i=0; i<m_iv.size(); ++i)
		{
			if (!out.empty() && m_iv[i].lo <= out.back().hi + 1) out.back().hi = std::max( out.back().hi, m_iv[i].hi);
			else out.push_back( m_iv[i]);
		}
		m_iv.swap( out);
		m_clean = true;
	}
	mutable std::vector<Iv> m_iv;
	mutable bool m_clean = true;
};

// ---- unicode helpers on top of the generated tables (include/sp_unicode_tables.h)
inline void insertGeneralCategory( CodeSet& set, const std::string& name)
{
	bool any = false;
	for (uint32_t ci=0; ci<sp_num_categories; ++ci)
	{
		const sp_cp_category& cat = sp_categories[ci];
		bool match = (name.size() == 1) ? (cat.name[0] == name[0]) : (name == cat.name);
		if (!match) continue;
		any = true;
		for (uint32_t ri=0; ri<cat.count; ++ri) set.insert( cat.ranges[ri].lo, cat.ranges[ri].hi);
	}
	if (!any) throw SyntaxError( "unknown unicode property '" + name + "'");
}
inline CodeSet foldCase( const CodeSet& in)
{
	CodeSet cur = in;
	for (int pass=0; pass<2; ++pass)	// pairs are one-to-one; two passes close chains like K, k, KELVIN SIGN
	{
		CodeSet add;
		for (uint32_t pi=0; pi<sp_num_case_pairs; ++pi)
		{
			if (cur.has( sp_case_pairs[pi].cp)) add.insert( sp_case_pairs[pi].other, sp_case_pairs[pi].other);
		}
		cur.insert( add);
	}
	return cur;
}
inline CodeSet classWord( unsigned int flags)
{
	CodeSet s; s.insert('0','9'); s.insert('A','Z'); s.insert('a','z'); s.insert('_','_');
	if (flags & RX_UCP) {insertGeneralCategory( s, "L"); insertGeneralCategory( s, "N");}
	return s;
}
inline CodeSet classDigit( unsigned int flags)
{
	CodeSet s; s.insert('0','9');
	if (flags & RX_UCP) insertGeneralCategory( s, "Nd");
	return s;
}
inline CodeSet classSpace( unsigned int flags)
{
	CodeSet s; s.insert( 0x09, 0x0D); s.insert( 0x20, 0x20);
	if (flags & RX_UCP) {insertGeneralCategory( s, "Z"); s.insert( 0x85, 0x85);}
	return s;
}

// ---- syntax tree
struct Rx;
typedef std::shared_ptr<Rx> RxRef;
enum RxKind {RxEmpty, RxSet, RxConcat, RxChoice, RxLoop, RxGroup, RxAssert};
enum AssertKind {AsWordBoundary, AsNotWordBoundary, AsTextBegin, AsTextEnd, AsTextEndOptNewline, AsLineBegin, AsLineEnd};

struct Rx
{
	RxKind kind;
	CodeSet set;			// RxSet
	std::vector<RxRef> sub;		// RxConcat, RxChoice (n), RxLoop, RxGroup (1)
	int lo, hi;			// RxLoop: bounds, hi < 0 = unbounded
	int capture;			// RxGroup: 1-based capture index or 0
	AssertKind assertion;		// RxAssert

	explicit Rx( RxKind k) :kind(k),lo(0),hi(0),capture(0),assertion(AsWordBoundary){}
	static RxRef make( RxKind k) {return RxRef( new Rx( k));}
	static RxRef makeSet( const CodeSet& s) {RxRef r = make( RxSet); r->set = s; return r;}
	static RxRef makeAssert( AssertKind a) {RxRef r = make( RxAssert); r->assertion = a; return r;}
	static RxRef concatOf( const std::vector<RxRef>& v, std::size_t b, std::size_t e)
	{
		if (b >= e) return make( RxEmpty);
		if (e - b == 1) return v[b];
		RxRef r = make( RxConcat); r->sub.assign( v.begin()+b, v.begin()+e); return r;
	}
};

// ---- recursive descent parser over the UTF-8 expression string
class RegexParser
{
public:
	RegexParser( const std::string& expr, unsigned int flags) :m_s(expr),m_i(0),m_flags(flags),m_captures(0){}

	RxRef parse()
	{
		RxRef r = choice();
		if (m_i < m_s.size()) error( "unmatched ')'");
		return r;
	}
	int captureCount() const {return m_captures;}

private:
	[[noreturn]] void error( const std::string& what) const
	{
		char pos[ 32]; snprintf( pos, sizeof(pos), " at offset %u", (unsigned int)m_i);
		throw SyntaxError( what + pos);
	}
	bool more() const {return m_i < m_s.size();}
	unsigned char cur() const {return (unsigned char)m_s[ m_i];}
	bool lookingAt( char c) const {return more() && m_s[ m_i] == c;}

	uint32_t takeCodePoint()
	{
		// expression strings are UTF-8; a malformed byte stands for itself
		unsigned char c = cur();
		int extra = c >= 0xF0 ? 3 : c >= 0xE0 ? 2 : c >= 0xC0 ? 1 : 0;
		if (!extra || m_i + extra >= m_s.size()) {++m_i; return c;}
		uint32_t cp = c & (0x3F >> extra);
		for (int k=1; k<=extra; ++k)
		{
			unsigned char cc = (unsigned char)m_s[ m_i+k];
			if ((cc & 0xC0) != 0x80) {++m_i; return c;}
			cp = (cp << 6) | (cc & 0x3F);
		}
		m_i += extra + 1;
		return cp;
	}
	RxRef literalOf( uint32_t cp) const
	{
		CodeSet s = CodeSet::single( cp);
		return Rx::makeSet( (m_flags & RX_CASELESS) ? foldCase( s) : s);
	}

	RxRef choice()
	{
		std::vector<RxRef> alts( 1, sequence());
		while (lookingAt( '|')) {++m_i; alts.push_back( sequence());}
		if (alts.size() == 1) return alts[0];
		RxRef r = Rx::make( RxChoice); r->sub = alts; return r;
	}
	RxRef sequence()
	{
		std::vector<RxRef> items;
		while (more() && cur() != '|' && cur() != ')') items.push_back( quantified( atom()));
		return Rx::concatOf( items, 0, items.size());
	}
	bool bounds( int& lo, int& hi)
	{
		// m_i is at '{'; accepts {n} {n,} {n,m}; anything else leaves '{' a literal
		std::size_t p = m_i + 1;
		auto num = [&]( int& out) -> bool
		{
			std::size_t b = p; long v = 0;
			for (; p < m_s.size() && m_s[p] >= '0' && m_s[p] <= '9'; ++p) {v = v*10 + (m_s[p]-'0'); if (v > 32767) error( "repeat count too large");}
			out = (int)v; return p != b;
		};
		if (!num( lo)) return false;
		if (p < m_s.size() && m_s[p] == '}') {hi = lo; m_i = p+1; return true;}
		if (p >= m_s.size() || m_s[p] != ',') return false;
		++p;
		if (p < m_s.size() && m_s[p] == '}') {hi = -1; m_i = p+1; return true;}
		if (!num( hi) || p >= m_s.size() || m_s[p] != '}') return false;
		if (hi < lo) error( "repeat bounds out of order");
		m_i = p+1;
		return true;
	}
	RxRef quantified( RxRef operand)
	{
		while (more())
		{
			int lo, hi;
			switch (cur())
			{
				case '*': lo = 0; hi = -1; ++m_i; break;
				case '+': lo = 1; hi = -1; ++m_i; break;
				case '?': lo = 0; hi = 1; ++m_i; break;
				case '{': if (!bounds( lo, hi)) return operand; break;
				default: return operand;
			}
			if (operand->kind == RxAssert) error( "quantifier applied to an assertion");
			if (lookingAt( '?')) ++m_i;
			else if (lookingAt( '+')) error( "possessive quantifiers are not supported");
			RxRef loop = Rx::make( RxLoop);
			loop->lo = lo; loop->hi = hi; loop->sub.push_back( operand);
			operand = loop;
		}
		return operand;
	}
	RxRef atom()
	{
		switch (cur())
		{
			case '(': return group();
			case '[': return bracket();
			case '\\': return escape();
			case '.':
			{
				++m_i;
				if (m_flags & RX_DOTALL) return Rx::makeSet( CodeSet::all());
				return Rx::makeSet( CodeSet::single( '\n').complement());
			}
			case '^': ++m_i; return Rx::makeAssert( (m_flags & RX_MULTILINE) ? AsLineBegin : AsTextBegin);
			case '$': ++m_i; return Rx::makeAssert( (m_flags & RX_MULTILINE) ? AsLineEnd : AsTextEndOptNewline);
			case '*': case '+': case '?': error( "quantifier without operand");
			default: return literalOf( takeCodePoint());
		}
	}
	RxRef group()
	{
		++m_i;
		unsigned int outer = m_flags;
		int cap = 0;
		if (lookingAt( '?'))
		{
			++m_i;
			if (!more()) error( "incomplete group");
			char c = m_s[ m_i];
			if (c == ':') ++m_i;
			else if (c == '#')
			{
				while (more() && cur() != ')') ++m_i;
				if (!more()) error( "unterminated comment");
				++m_i;
				return Rx::make( RxEmpty);
			}
			else if (std::strchr( "=!<>P'|", c)) error( "look-around, atomic, named and branch-reset groups are not supported");
			else
			{
				bool enable = true;
				for (;; ++m_i)
				{
					if (!more()) error( "incomplete inline option");
					char f = m_s[ m_i];
					if (f == ')' || f == ':') break;
					unsigned int bit = f == 'i' ? RX_CASELESS : f == 's' ? RX_DOTALL : f == 'm' ? RX_MULTILINE : 0;
					if (f == '-') {enable = false; continue;}
					if (!bit) error( "unsupported inline option");
					if (enable) m_flags |= bit; else m_flags &= ~bit;
				}
				if (m_s[ m_i] == ')') {++m_i; return Rx::make( RxEmpty);}	// (?i) stays on up to the end of the enclosing group
				++m_i;
			}
		}
		else cap = ++m_captures;
		RxRef body = choice();
		if (!lookingAt( ')')) error( "missing ')'");
		++m_i;
		m_flags = outer;
		RxRef g = Rx::make( RxGroup);
		g->capture = cap; g->sub.push_back( body);
		return g;
	}
	static int hexDigit( unsigned char c)
	{
		return (c >= '0' && c <= '9') ? c-'0' : (c >= 'a' && c <= 'f') ? c-'a'+10 : (c >= 'A' && c <= 'F') ? c-'A'+10 : -1;
	}
	uint32_t hexNumber()
	{
		uint32_t v = 0;
		if (lookingAt( '{'))
		{
			++m_i;
			int n = 0;
			for (; more() && cur() != '}'; ++m_i, ++n)
			{
				int d = hexDigit( cur());
				if (d < 0) error( "bad hex escape");
				v = v*16 + d;
				if (v > kMaxCodePoint) error( "hex escape out of range");
			}
			if (!more() || !n) error( "bad hex escape");
			++m_i;
			return v;
		}
		for (int n=0; n<2 && more() && hexDigit( cur()) >= 0; ++n, ++m_i) v = v*16 + hexDigit( cur());
		return v;
	}
	bool shorthandClass( uint32_t e, CodeSet& out) const
	{
		switch (e)
		{
			case 'd': out = classDigit( m_flags); return true;
			case 'D': out = classDigit( m_flags).complement(); return true;
			case 'w': out = classWord( m_flags); return true;
			case 'W': out = classWord( m_flags).complement(); return true;
			case 's': out = classSpace( m_flags); return true;
			case 'S': out = classSpace( m_flags).complement(); return true;
			case 'h': case 'H':
			{
				static const uint32_t hs[][2] = {{0x09,0x09},{0x20,0x20},{0xA0,0xA0},{0x1680,0x1680},{0x180E,0x180E},{0x2000,0x200A},{0x202F,0x202F},{0x205F,0x205F},{0x3000,0x3000}};
				CodeSet s; for (auto& r: hs) s.insert( r[0], r[1]);
				out = (e == 'h') ? s : s.complement(); return true;
			}
			case 'v': case 'V':
			{
				CodeSet s; s.insert( 0x0A, 0x0D); s.insert( 0x85, 0x85); s.insert( 0x2028, 0x2029);
				out = (e == 'v') ? s : s.complement(); return true;
			}
			default: return false;
		}
	}
	static bool controlEscape( uint32_t e, uint32_t& cp)
	{
		switch (e)
		{
			case 't': cp = 9; return true;
			case 'n': cp = 10; return true;
			case 'r': cp = 13; return true;
			case 'f': cp = 12; return true;
			case 'a': cp = 7; return true;
			case 'e': cp = 27; return true;
			case '0': cp = 0; return true;
			default: return false;
		}
	}
	CodeSet property( bool negate)
	{
		std::string name;
		if (!more()) error( "incomplete \\p");
		if (lookingAt( '{'))
		{
			++m_i;
			if (lookingAt( '^')) {negate = !negate; ++m_i;}
			while (more() && cur() != '}') name.push_back( m_s[ m_i++]);
			if (!more()) error( "unterminated \\p{");
			++m_i;
		}
		else name.push_back( m_s[ m_i++]);
		CodeSet s;
		if (name == "Any") s = CodeSet::all();
		else if (name == "L&") {insertGeneralCategory( s, "Lu"); insertGeneralCategory( s, "Ll"); insertGeneralCategory( s, "Lt");}
		else if (name.size() == 1 || name.size() == 2) insertGeneralCategory( s, name);
		else error( "unsupported unicode property (only general categories are available)");
		return negate ? s.complement() : s;
	}
	RxRef escape()
	{
		++m_i;
		if (!more()) error( "trailing backslash");
		uint32_t e = takeCodePoint();
		CodeSet cls; uint32_t cp;
		if (shorthandClass( e, cls)) return Rx::makeSet( cls);
		if (controlEscape( e, cp)) return literalOf( cp);
		switch (e)
		{
			case 'b': return Rx::makeAssert( AsWordBoundary);
			case 'B': return Rx::makeAssert( AsNotWordBoundary);
			case 'A': return Rx::makeAssert( AsTextBegin);
			case 'z': return Rx::makeAssert( AsTextEnd);
			case 'Z': return Rx::makeAssert( AsTextEndOptNewline);
			case 'p': return Rx::makeSet( property( false));
			case 'P': return Rx::makeSet( property( true));
			case 'x': return literalOf( hexNumber());
			case 'Q':
			{
				std::vector<RxRef> lits;
				while (more())
				{
					if (cur() == '\\' && m_i+1 < m_s.size() && m_s[ m_i+1] == 'E') {m_i += 2; break;}
					lits.push_back( literalOf( takeCodePoint()));
				}
				return Rx::concatOf( lits, 0, lits.size());
			}
			default: break;
		}
		if (e >= '1' && e <= '9') error( "back references are not supported");
		if ((e|0x20) >= 'a' && (e|0x20) <= 'z' && e < 128) error( "unsupported escape sequence");
		return literalOf( e);
	}
	void posixNamed( const std::string& name, CodeSet& s)
	{
		struct Def {const char* name; uint32_t r[4][2]; int n;};
		static const Def defs[] = {
			{"alpha",{{'a','z'},{'A','Z'}},2}, {"digit",{{'0','9'}},1}, {"alnum",{{'a','z'},{'A','Z'},{'0','9'}},3},
			{"upper",{{'A','Z'}},1}, {"lower",{{'a','z'}},1}, {"space",{{9,13},{32,32}},2}, {"blank",{{9,9},{32,32}},2},
			{"punct",{{0x21,0x2F},{0x3A,0x40},{0x5B,0x60},{0x7B,0x7E}},4}, {"xdigit",{{'0','9'},{'a','f'},{'A','F'}},3},
			{"word",{{'a','z'},{'A','Z'},{'0','9'},{'_','_'}},4}, {"cntrl",{{0,0x1F},{0x7F,0x7F}},2}, {"print",{{0x20,0x7E}},1},
			{"graph",{{0x21,0x7E}},1}, {"ascii",{{0,0x7F}},1}};
		for (const Def& d: defs)
		{
			if (name != d.name) continue;
			for (int k=0; k<d.n; ++k) s.insert( d.r[k][0], d.r[k][1]);
			return;
		}
		error( "unknown posix class");
	}
	// one range end point or single member inside [...]; returns false if a whole set was added to `acc`
	bool bracketMember( CodeSet& acc, uint32_t& cp, bool asRangeEnd)
	{
		if (cur() == '\\')
		{
			++m_i;
			if (!more()) error( "trailing backslash");
			uint32_t e = takeCodePoint();
			CodeSet cls;
			if (!asRangeEnd)
			{
				if (shorthandClass( e, cls)) {acc.insert( cls); return false;}
				if (e == 'p' || e == 'P') {acc.insert( property( e == 'P')); return false;}
			}
			if (e == 'b') {cp = 8; return true;}
			if (e == 'x') {cp = hexNumber(); return true;}
			if (controlEscape( e, cp)) return true;
			if (e < 128 && (e|0x20) >= 'a' && (e|0x20) <= 'z') error( "unsupported escape sequence in character class");
			cp = e;
			return true;
		}
		cp = takeCodePoint();
		return true;
	}
	RxRef bracket()
	{
		++m_i;
		bool negate = false;
		if (lookingAt( '^')) {negate = true; ++m_i;}
		CodeSet acc;
		for (bool first=true;; first=false)
		{
			if (!more()) error( "unterminated character class");
			if (cur() == ']' && !first) {++m_i; break;}
			if (cur() == '[' && m_i+1 < m_s.size() && m_s[ m_i+1] == ':')
			{
				std::size_t close = m_s.find( ":]", m_i+2);
				if (close == std::string::npos) error( "unterminated posix class");
				std::string name = m_s.substr( m_i+2, close-m_i-2);
				m_i = close + 2;
				CodeSet named;
				bool neg = !name.empty() && name[0] == '^';
				posixNamed( neg ? name.substr(1) : name, named);
				acc.insert( neg ? named.complement() : named);
				continue;
			}
			uint32_t lo, hi;
			if (!bracketMember( acc, lo, false)) continue;
			hi = lo;
			if (m_i+1 < m_s.size() && cur() == '-' && m_s[ m_i+1] != ']')
			{
				++m_i;
				bracketMember( acc, hi, true);
				if (hi < lo) error( "range out of order in character class");
			}
			acc.insert( lo, hi);
		}
		if (m_flags & RX_CASELESS) acc = foldCase( acc);
		return Rx::makeSet( negate ? acc.complement() : acc);
	}

	std::string m_s;
	std::size_t m_i;
	unsigned int m_flags;
	int m_captures;
};

} // namespace sp
#endif
