// ORACLE (test infrastructure, not product code) -- CPU restatement of strusPattern level 1.
//
// Restates, for checking the CUDA path, what the reference's lexer computes
// (all file:line below are under /root/reference):
//   * definition handling            src/patternLexer.cpp:245-262 (definePattern), :605-626 (" ~N" suffix),
//                                    :264-317 (symbols), :333-412 (complete: engine id = 1-based definition index)
//   * the engine contract            hs_compile_ext_multi / hs_scan with HS_FLAG_UTF8|HS_FLAG_SOM_LEFTMOST,
//                                    HS_MODE_BLOCK (src/patternLexer.cpp:403, :1083-1086, :879)
//   * match_event_handler            src/patternLexer.cpp:717-826 (supersede logic incl. the double-insert quirk)
//   * ordinal position assignment    src/patternLexer.cpp:893-945
//   * capture group selection        src/patternLexer.cpp:414-426, 488-507 (libtre re-match, POSIX leftmost-longest)
//
// The regex arithmetic of the reference lives in Intel Hyperscan (v5.1.1 per INSTALL.Ubuntu.md:15-20) and
// libtre; neither is in /root/reference nor installed, so their published semantics are restated:
//   - every end offset `to` at which some non-empty substring [from,to) matches pattern p is reported once
//     as (p, from, to) with `from` the LEFTMOST start for that end; events are delivered by ascending `to`;
//     the canonical order among patterns at one `to` is ascending definition index (consistent with the
//     golden vector tests/charRegexMatch/src/testCharRegexMatch.cpp:155-156).
//   - HS_FLAG_UTF8: '.', classes and literals consume whole code points; \w \d \s \b are ASCII unless UCP.
// The matcher here is deliberately the simplest correct one: code-point level Thompson NFA simulated from
// every start position in ascending order (first start to reach an end = leftmost start).
//
// PARITY PIN: golden case 1 of tests/charRegexMatch/src/testCharRegexMatch.cpp:107-160 (tests/test_oracle_lexer.py).
// Cases 2 and 3 (edit distance "~1") need Hyperscan's approximate matching + libtre regaexec and are a
// "next" row: definitions with edit distance / option BYTECHAR report "unsupported".
// Everything else about Hyperscan's behaviour is unpinned by the reference's tests.
#include "strus_pattern_b200.h"
#include "sp_unicode_tables.h"
#include <vector>
#include <string>
#include <map>
#include <unordered_map>
#include <algorithm>
#include <stdexcept>
#include <thread>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <memory>

namespace {

enum {FLAG_CASELESS=1, FLAG_DOTALL=2, FLAG_MULTILINE=4, FLAG_ALLOWEMPTY=8, FLAG_UCP=16};
enum {MaxPatternId=(1<<30)-1};
static const uint32_t CP_MAX = 0x10FFFF;
static const uint32_t NONE = 0xFFFFFFFFu;

// ---------------------------------------------------------------- code point sets
struct CpSet
{
	std::vector<std::pair<uint32_t,uint32_t> > r;	// sorted, disjoint, non adjacent after normalize()

	void add( uint32_t lo, uint32_t hi) {if (lo <= hi) r.push_back( std::make_pair( lo, hi));}
	void add( uint32_t c) {add( c, c);}
	void add( const CpSet& o) {r.insert( r.end(), o.r.begin(), o.r.end());}
	void normalize()
	{
		std::sort( r.begin(), r.end());
		std::vector<std::pair<uint32_t,uint32_t> > n;
		for (std::size_t i=0; i<r.size(); ++i)
		{
			if (!n.empty() && r[i].first <= n.back().second + 1) {if (r[i].second > n.back().second) n.back().second = r[i].second;}
			else n.push_back( r[i]);
		}
		r.swap( n);
	}
	void negate()
	{
		normalize();
		std::vector<std::pair<uint32_t,uint32_t> > n;
		uint32_t next = 0;
		for (std::size_t i=0; i<r.size(); ++i)
		{
			if (r[i].first > next) n.push_back( std::make_pair( next, r[i].first-1));
			next = r[i].second + 1;
		}
		if (next <= CP_MAX) n.push_back( std::make_pair( next, CP_MAX));
		r.swap( n);
	}
	bool contains( uint32_t c) const
	{
		std::size_t lo = 0, hi = r.size();
		while (lo < hi)
		{
			std::size_t mid = (lo + hi) / 2;
			if (c < r[mid].first) hi = mid; else if (c > r[mid].second) lo = mid+1; else return true;
		}
		return false;
	}
};

static void addCategory( CpSet& set, const char* name)
{
	// name: one letter (all sub categories) or two letters
	bool found = false;
	for (uint32_t ci=0; ci<sp_num_categories; ++ci)
	{
		const sp_cp_category& cat = sp_categories[ci];
		bool take = name[1] ? (0==std::strcmp( cat.name, name)) : (cat.name[0] == name[0]);
		if (!take) continue;
		found = true;
		for (uint32_t ri=0; ri<cat.count; ++ri) set.add( cat.ranges[ri].lo, cat.ranges[ri].hi);
	}
	if (!found) throw std::runtime_error( std::string("unknown unicode property '") + name + "'");
}

static void addCaseVariants( CpSet& set)
{
	// simple one-to-one case pairs, closed transitively by two rounds
	for (int round=0; round<2; ++round)
	{
		set.normalize();
		CpSet extra;
		for (uint32_t pi=0; pi<sp_num_case_pairs; ++pi)
		{
			if (set.contains( sp_case_pairs[pi].cp)) extra.add( sp_case_pairs[pi].other);
		}
		set.add( extra);
	}
	set.normalize();
}

static CpSet wordSet( unsigned int flags)
{
	CpSet s;
	s.add( '0','9'); s.add( 'A','Z'); s.add( 'a','z'); s.add( '_');
	if (flags & FLAG_UCP) {addCategory( s, "L"); addCategory( s, "N");}
	s.normalize();
	return s;
}
static CpSet digitSet( unsigned int flags)
{
	CpSet s;
	s.add( '0','9');
	if (flags & FLAG_UCP) addCategory( s, "Nd");
	s.normalize();
	return s;
}
static CpSet spaceSet( unsigned int flags)
{
	CpSet s;
	s.add( 0x09, 0x0D); s.add( 0x20);
	if (flags & FLAG_UCP) {addCategory( s, "Z"); s.add( 0x85);}
	s.normalize();
	return s;
}

// ---------------------------------------------------------------- regex AST
enum NodeType {N_EMPTY, N_SET, N_CAT, N_ALT, N_REPEAT, N_GROUP, N_ASSERT};
enum AssertKind {A_WORDB, A_NWORDB, A_BEGIN_TEXT, A_END_TEXT, A_END_TEXT_OPT_LF, A_BEGIN_LINE, A_END_LINE};

struct Node
{
	NodeType type;
	CpSet set;
	std::vector<std::shared_ptr<Node> > kids;
	int rmin, rmax;		// repeat bounds, rmax < 0 = unbounded
	int capidx;		// N_GROUP: capture index (0 = non capturing)
	AssertKind akind;

	explicit Node( NodeType t) :type(t),rmin(0),rmax(0),capidx(0),akind(A_WORDB){}
};
typedef std::shared_ptr<Node> NodeRef;

static uint32_t decodeUtf8( const unsigned char* s, std::size_t n, std::size_t& len)
{
	// strict decoder; malformed bytes decode as one code point U+FFFD per byte
	unsigned char c = s[0];
	if (c < 0x80) {len = 1; return c;}
	int need = (c >= 0xF0 && c <= 0xF4) ? 3 : (c >= 0xE0 && c < 0xF0) ? 2 : (c >= 0xC2 && c < 0xE0) ? 1 : -1;
	if (need < 0 || (std::size_t)need >= n) {len = 1; return 0xFFFD;}
	uint32_t cp = (need == 1) ? (c & 0x1F) : (need == 2) ? (c & 0x0F) : (c & 0x07);
	for (int i=1; i<=need; ++i)
	{
		if ((s[i] & 0xC0) != 0x80) {len = 1; return 0xFFFD;}
		cp = (cp << 6) | (s[i] & 0x3F);
	}
	if ((need == 2 && cp < 0x800) || (need == 3 && (cp < 0x10000 || cp > CP_MAX)) || (cp >= 0xD800 && cp <= 0xDFFF)) {len = 1; return 0xFFFD;}
	len = need + 1;
	return cp;
}

class Parser
{
public:
	Parser( const std::string& src_, unsigned int flags_)
		:m_src(src_),m_pos(0),m_flags(flags_),m_ncap(0){}

	NodeRef parse()
	{
		NodeRef rt = parseAlt();
		if (m_pos != m_src.size()) fail( "unbalanced ')'");
		return rt;
	}
	int nofCaptures() const {return m_ncap;}

private:
	[[noreturn]] void fail( const char* msg) const
	{
		char buf[ 256];
		snprintf( buf, sizeof(buf), "%s at offset %u", msg, (unsigned int)m_pos);
		throw std::runtime_error( buf);
	}
	bool eof() const {return m_pos >= m_src.size();}
	unsigned char peek() const {return (unsigned char)m_src[ m_pos];}
	uint32_t nextCp()
	{
		std::size_t len = 0;
		uint32_t cp = decodeUtf8( (const unsigned char*)m_src.data() + m_pos, m_src.size() - m_pos, len);
		m_pos += len;
		return cp;
	}

	NodeRef parseAlt()
	{
		std::vector<NodeRef> alts;
		alts.push_back( parseCat());
		while (!eof() && peek() == '|') {++m_pos; alts.push_back( parseCat());}
		if (alts.size() == 1) return alts[0];
		NodeRef rt( new Node( N_ALT));
		rt->kids = alts;
		return rt;
	}
	NodeRef parseCat()
	{
		NodeRef rt( new Node( N_CAT));
		while (!eof() && peek() != '|' && peek() != ')')
		{
			NodeRef atom = parseAtom();
			atom = parseQuantifiers( atom);
			rt->kids.push_back( atom);
		}
		if (rt->kids.empty()) return NodeRef( new Node( N_EMPTY));
		if (rt->kids.size() == 1) return rt->kids[0];
		return rt;
	}
	NodeRef parseQuantifiers( NodeRef atom)
	{
		for (;;)
		{
			if (eof()) return atom;
			int rmin, rmax;
			unsigned char c = peek();
			if (c == '*') {rmin = 0; rmax = -1; ++m_pos;}
			else if (c == '+') {rmin = 1; rmax = -1; ++m_pos;}
			else if (c == '?') {rmin = 0; rmax = 1; ++m_pos;}
			else if (c == '{')
			{
				std::size_t save = m_pos;
				if (!parseBounds( rmin, rmax)) {m_pos = save; return atom;}	// literal '{' handled by caller as atom next round
			}
			else return atom;
			if (atom->type == N_ASSERT) fail( "quantifier on an assertion");
			// lazy / possessive suffix: laziness is irrelevant under all-matches semantics
			if (!eof() && peek() == '?') ++m_pos;
			else if (!eof() && peek() == '+') fail( "possessive quantifiers are not supported");
			NodeRef rep( new Node( N_REPEAT));
			rep->rmin = rmin; rep->rmax = rmax;
			rep->kids.push_back( atom);
			atom = rep;
		}
	}
	bool parseBounds( int& rmin, int& rmax)
	{
		// at '{'
		std::size_t p = m_pos + 1;
		auto number = [&]( int& v) -> bool {
			std::size_t b = p; long n = 0;
			while (p < m_src.size() && m_src[p] >= '0' && m_src[p] <= '9') {n = n*10 + (m_src[p]-'0'); if (n > 100000) fail( "repeat bound too large"); ++p;}
			v = (int)n; return p > b;
		};
		if (!number( rmin)) return false;
		if (p < m_src.size() && m_src[p] == '}') {rmax = rmin; m_pos = p+1; return true;}
		if (p >= m_src.size() || m_src[p] != ',') return false;
		++p;
		if (p < m_src.size() && m_src[p] == '}') {rmax = -1; m_pos = p+1; return true;}
		if (!number( rmax)) return false;
		if (p >= m_src.size() || m_src[p] != '}') return false;
		if (rmax < rmin) fail( "repeat bounds out of order");
		m_pos = p+1;
		return true;
	}
	NodeRef setNode( const CpSet& s)
	{
		NodeRef rt( new Node( N_SET));
		rt->set = s;
		rt->set.normalize();
		return rt;
	}
	NodeRef literal( uint32_t cp)
	{
		CpSet s; s.add( cp);
		if (m_flags & FLAG_CASELESS) addCaseVariants( s);
		return setNode( s);
	}
	NodeRef assertion( AssertKind k)
	{
		NodeRef rt( new Node( N_ASSERT));
		rt->akind = k;
		return rt;
	}
	NodeRef parseAtom()
	{
		unsigned char c = peek();
		switch (c)
		{
			case '(': return parseGroup();
			case '[': return parseClass();
			case '.':
			{
				++m_pos;
				CpSet s;
				if (m_flags & FLAG_DOTALL) s.add( 0, CP_MAX);
				else {s.add( 0, '\n'-1); s.add( '\n'+1, CP_MAX);}
				return setNode( s);
			}
			case '^': ++m_pos; return assertion( (m_flags & FLAG_MULTILINE) ? A_BEGIN_LINE : A_BEGIN_TEXT);
			case '$': ++m_pos; return assertion( (m_flags & FLAG_MULTILINE) ? A_END_LINE : A_END_TEXT_OPT_LF);
			case '\\': return parseEscape();
			case '*': case '+': case '?': fail( "quantifier without operand");
			default: return literal( nextCp());
		}
	}
	NodeRef parseGroup()
	{
		++m_pos;	// '('
		int capidx = 0;
		unsigned int savedflags = m_flags;
		bool scoped = false;
		if (!eof() && peek() == '?')
		{
			++m_pos;
			if (eof()) fail( "incomplete group");
			if (peek() == ':') {++m_pos;}
			else if (peek() == '#')
			{
				while (!eof() && peek() != ')') ++m_pos;
				if (eof()) fail( "unterminated comment");
				++m_pos;
				return NodeRef( new Node( N_EMPTY));
			}
			else if (peek() == '=' || peek() == '!' || peek() == '<' || peek() == '>' || peek() == 'P' || peek() == '\'')
			{
				fail( "look-around, atomic and named groups are not supported");
			}
			else
			{
				// inline flags (?i) (?s) (?m) (?-i) (?i:...)
				bool on = true;
				for (;;)
				{
					if (eof()) fail( "incomplete inline flags");
					unsigned char f = peek();
					if (f == '-') {on = false; ++m_pos; continue;}
					unsigned int bit = 0;
					if (f == 'i') bit = FLAG_CASELESS; else if (f == 's') bit = FLAG_DOTALL; else if (f == 'm') bit = FLAG_MULTILINE;
					else if (f == ')' || f == ':') break;
					else fail( "unsupported inline flag");
					if (on) m_flags |= bit; else m_flags &= ~bit;
					++m_pos;
				}
				if (peek() == ')') {++m_pos; return NodeRef( new Node( N_EMPTY));}	// flags stay set until the enclosing group ends
				++m_pos; scoped = true;
			}
		}
		else
		{
			capidx = ++m_ncap;
		}
		(void)scoped;
		NodeRef body = parseAlt();
		if (eof() || peek() != ')') fail( "missing ')'");
		++m_pos;
		m_flags = savedflags;	// inline flags set inside a group end with the group
		NodeRef rt( new Node( N_GROUP));
		rt->capidx = capidx;
		rt->kids.push_back( body);
		return rt;
	}
	bool classEscape( uint32_t e, CpSet& s)
	{
		switch (e)
		{
			case 'd': s = digitSet( m_flags); return true;
			case 'D': s = digitSet( m_flags); s.negate(); return true;
			case 'w': s = wordSet( m_flags); return true;
			case 'W': s = wordSet( m_flags); s.negate(); return true;
			case 's': s = spaceSet( m_flags); return true;
			case 'S': s = spaceSet( m_flags); s.negate(); return true;
			case 'h': s.add( 0x09); s.add( 0x20); s.add( 0xA0); s.add( 0x1680); s.add( 0x180E); s.add( 0x2000, 0x200A); s.add( 0x202F); s.add( 0x205F); s.add( 0x3000); s.normalize(); return true;
			case 'v': s.add( 0x0A, 0x0D); s.add( 0x85); s.add( 0x2028, 0x2029); s.normalize(); return true;
			default: return false;
		}
	}
	void parseProperty( bool negated, CpSet& s)
	{
		// at the char after \p or \P
		std::string name;
		if (eof()) fail( "incomplete \\p");
		if (peek() == '{')
		{
			++m_pos;
			if (!eof() && peek() == '^') {negated = !negated; ++m_pos;}
			while (!eof() && peek() != '}') name.push_back( (char)m_src[ m_pos++]);
			if (eof()) fail( "unterminated \\p{");
			++m_pos;
		}
		else name.push_back( (char)m_src[ m_pos++]);
		if (name == "Any") s.add( 0, CP_MAX);
		else if (name == "L&") {addCategory( s, "Lu"); addCategory( s, "Ll"); addCategory( s, "Lt");}
		else
		{
			if (name.size() < 1 || name.size() > 2) fail( "unsupported unicode property (scripts are not supported)");
			addCategory( s, name.c_str());
		}
		s.normalize();
		if (negated) s.negate();
	}
	uint32_t parseHexEscape()
	{
		// after \x
		uint32_t v = 0;
		if (!eof() && peek() == '{')
		{
			++m_pos;
			int nd = 0;
			while (!eof() && peek() != '}')
			{
				int d = hexval( peek());
				if (d < 0) fail( "bad hex escape");
				v = v*16 + d; ++nd; ++m_pos;
				if (v > CP_MAX) fail( "hex escape out of range");
			}
			if (eof() || nd == 0) fail( "bad hex escape");
			++m_pos;
			return v;
		}
		int nd = 0;
		while (nd < 2 && !eof() && hexval( peek()) >= 0) {v = v*16 + hexval( peek()); ++m_pos; ++nd;}
		return v;
	}
	static int hexval( unsigned char c)
	{
		if (c >= '0' && c <= '9') return c - '0';
		if (c >= 'a' && c <= 'f') return c - 'a' + 10;
		if (c >= 'A' && c <= 'F') return c - 'A' + 10;
		return -1;
	}
	bool simpleEscape( uint32_t e, uint32_t& cp)
	{
		switch (e)
		{
			case 't': cp = 0x09; return true;
			case 'n': cp = 0x0A; return true;
			case 'r': cp = 0x0D; return true;
			case 'f': cp = 0x0C; return true;
			case 'a': cp = 0x07; return true;
			case 'e': cp = 0x1B; return true;
			case '0': cp = 0; return true;
			default: return false;
		}
	}
	NodeRef parseEscape()
	{
		++m_pos;	// backslash
		if (eof()) fail( "trailing backslash");
		uint32_t e = nextCp();
		CpSet s;
		uint32_t cp;
		if (classEscape( e, s)) return setNode( s);
		if (simpleEscape( e, cp)) return literal( cp);
		switch (e)
		{
			case 'b': return assertion( A_WORDB);
			case 'B': return assertion( A_NWORDB);
			case 'A': return assertion( A_BEGIN_TEXT);
			case 'z': return assertion( A_END_TEXT);
			case 'Z': return assertion( A_END_TEXT_OPT_LF);
			case 'p': parseProperty( false, s); return setNode( s);
			case 'P': parseProperty( true, s); return setNode( s);
			case 'x': return literal( parseHexEscape());
			case 'Q':
			{
				NodeRef cat( new Node( N_CAT));
				while (!eof())
				{
					if (peek() == '\\' && m_pos+1 < m_src.size() && m_src[m_pos+1] == 'E') {m_pos += 2; break;}
					cat->kids.push_back( literal( nextCp()));
				}
				if (cat->kids.empty()) return NodeRef( new Node( N_EMPTY));
				return cat;
			}
			default: break;
		}
		if (e >= '1' && e <= '9') fail( "back references are not supported");
		if ((e >= 'a' && e <= 'z') || (e >= 'A' && e <= 'Z')) fail( "unsupported escape sequence");
		return literal( e);
	}
	NodeRef parseClass()
	{
		++m_pos;	// '['
		bool negated = false;
		if (!eof() && peek() == '^') {negated = true; ++m_pos;}
		CpSet s;
		bool first = true;
		for (;;)
		{
			if (eof()) fail( "unterminated character class");
			unsigned char c = peek();
			if (c == ']' && !first) {++m_pos; break;}
			first = false;
			uint32_t lo;
			bool isset = false;
			CpSet sub;
			if (c == '[' && m_pos+1 < m_src.size() && m_src[m_pos+1] == ':')
			{
				std::size_t e = m_src.find( ":]", m_pos+2);
				if (e == std::string::npos) fail( "unterminated posix class");
				std::string name = m_src.substr( m_pos+2, e-(m_pos+2));
				m_pos = e+2;
				bool neg = false;
				if (!name.empty() && name[0] == '^') {neg = true; name.erase( 0, 1);}
				posixClass( name, sub);
				if (neg) sub.negate();
				isset = true;
			}
			else if (c == '\\')
			{
				++m_pos;
				if (eof()) fail( "trailing backslash");
				uint32_t e = nextCp();
				if (classEscape( e, sub)) isset = true;
				else if (e == 'p') {parseProperty( false, sub); isset = true;}
				else if (e == 'P') {parseProperty( true, sub); isset = true;}
				else if (e == 'b') lo = 0x08;
				else if (e == 'x') lo = parseHexEscape();
				else if (simpleEscape( e, lo)) {}
				else if ((e >= 'a' && e <= 'z') || (e >= 'A' && e <= 'Z')) fail( "unsupported escape sequence in class");
				else lo = e;
			}
			else lo = nextCp();
			if (isset) {s.add( sub); continue;}
			uint32_t hi = lo;
			if (m_pos+1 < m_src.size() && peek() == '-' && m_src[m_pos+1] != ']')
			{
				++m_pos;
				if (peek() == '\\')
				{
					++m_pos;
					if (eof()) fail( "trailing backslash");
					uint32_t e = nextCp();
					if (e == 'x') hi = parseHexEscape();
					else if (simpleEscape( e, hi)) {}
					else if (e == 'b') hi = 0x08;
					else if ((e >= 'a' && e <= 'z') || (e >= 'A' && e <= 'Z')) fail( "bad range end in class");
					else hi = e;
				}
				else hi = nextCp();
				if (hi < lo) fail( "range out of order in character class");
			}
			s.add( lo, hi);
		}
		s.normalize();
		if (m_flags & FLAG_CASELESS) addCaseVariants( s);
		if (negated) s.negate();
		return setNode( s);
	}
	void posixClass( const std::string& name, CpSet& s)
	{
		if (name == "alpha") {s.add('a','z'); s.add('A','Z');}
		else if (name == "digit") s.add('0','9');
		else if (name == "alnum") {s.add('a','z'); s.add('A','Z'); s.add('0','9');}
		else if (name == "upper") s.add('A','Z');
		else if (name == "lower") s.add('a','z');
		else if (name == "space") {s.add(0x09,0x0D); s.add(0x20);}
		else if (name == "blank") {s.add(0x09); s.add(0x20);}
		else if (name == "punct") {s.add(0x21,0x2F); s.add(0x3A,0x40); s.add(0x5B,0x60); s.add(0x7B,0x7E);}
		else if (name == "xdigit") {s.add('0','9'); s.add('a','f'); s.add('A','F');}
		else if (name == "word") {s.add('a','z'); s.add('A','Z'); s.add('0','9'); s.add('_');}
		else if (name == "cntrl") {s.add(0,0x1F); s.add(0x7F);}
		else if (name == "print") s.add(0x20,0x7E);
		else if (name == "graph") s.add(0x21,0x7E);
		else if (name == "ascii") s.add(0,0x7F);
		else fail( "unknown posix class");
		s.normalize();
	}

	std::string m_src;
	std::size_t m_pos;
	unsigned int m_flags;
	int m_ncap;
};

// ---------------------------------------------------------------- Thompson NFA over code points
struct NfaState
{
	enum Kind {Set, Eps, Assert, Accept} kind;
	int setidx;		// Set: index into Nfa::sets
	AssertKind akind;	// Assert
	int out1, out2;		// successors (-1 = none)
};

struct Nfa
{
	std::vector<NfaState> st;
	std::vector<CpSet> sets;
	int start;
	bool nullable;		// accepts the empty string in some context

	int add( NfaState::Kind k) {NfaState s; s.kind = k; s.setidx = -1; s.akind = A_WORDB; s.out1 = s.out2 = -1; st.push_back( s); return (int)st.size()-1;}
};

struct Frag {int start; std::vector<std::pair<int,int> > outs;};	// dangling (state, which out)

class NfaBuilder
{
public:
	explicit NfaBuilder( Nfa& nfa_) :m_nfa(nfa_){}

	void build( const NodeRef& root)
	{
		Frag f = gen( root);
		int acc = m_nfa.add( NfaState::Accept);
		patch( f, acc);
		m_nfa.start = f.start;
		if (m_nfa.st.size() > 200000) throw std::runtime_error( "pattern too large");
	}
private:
	void patch( Frag& f, int target)
	{
		for (std::size_t i=0; i<f.outs.size(); ++i)
		{
			if (f.outs[i].second == 1) m_nfa.st[ f.outs[i].first].out1 = target; else m_nfa.st[ f.outs[i].first].out2 = target;
		}
		f.outs.clear();
	}
	Frag eps()
	{
		Frag f; f.start = m_nfa.add( NfaState::Eps);
		f.outs.push_back( std::make_pair( f.start, 1));
		return f;
	}
	Frag cat( Frag a, Frag b)
	{
		patch( a, b.start);
		a.outs = b.outs;
		return a;
	}
	Frag gen( const NodeRef& n)
	{
		switch (n->type)
		{
			case N_EMPTY: return eps();
			case N_SET:
			{
				Frag f; f.start = m_nfa.add( NfaState::Set);
				m_nfa.sets.push_back( n->set);
				m_nfa.st[ f.start].setidx = (int)m_nfa.sets.size()-1;
				f.outs.push_back( std::make_pair( f.start, 1));
				return f;
			}
			case N_ASSERT:
			{
				Frag f; f.start = m_nfa.add( NfaState::Assert);
				m_nfa.st[ f.start].akind = n->akind;
				f.outs.push_back( std::make_pair( f.start, 1));
				return f;
			}
			case N_GROUP: return gen( n->kids[0]);
			case N_CAT:
			{
				Frag f = gen( n->kids[0]);
				for (std::size_t i=1; i<n->kids.size(); ++i) f = cat( f, gen( n->kids[i]));
				return f;
			}
			case N_ALT:
			{
				Frag f; f.start = -1;
				for (std::size_t i=n->kids.size(); i>0; --i)
				{
					Frag k = gen( n->kids[i-1]);
					if (f.start < 0) {f = k; continue;}
					int s = m_nfa.add( NfaState::Eps);
					m_nfa.st[s].out1 = k.start; m_nfa.st[s].out2 = f.start;
					Frag nf; nf.start = s; nf.outs = k.outs;
					nf.outs.insert( nf.outs.end(), f.outs.begin(), f.outs.end());
					f = nf;
				}
				return f;
			}
			case N_REPEAT:
			{
				const NodeRef& kid = n->kids[0];
				Frag f = eps();
				for (int i=0; i<n->rmin; ++i) f = cat( f, gen( kid));
				if (n->rmax < 0)
				{
					// kid*
					int s = m_nfa.add( NfaState::Eps);
					Frag k = gen( kid);
					m_nfa.st[s].out1 = k.start;
					patch( k, s);
					Frag loop; loop.start = s; loop.outs.push_back( std::make_pair( s, 2));
					f = cat( f, loop);
				}
				else
				{
					// (kid(kid(kid)?)?)? for the optional part, built from the inside out
					int opt = n->rmax - n->rmin;
					Frag tail; tail.start = -1;
					for (int i=0; i<opt; ++i)
					{
						Frag k = gen( kid);
						if (tail.start >= 0) k = cat( k, tail);
						int s = m_nfa.add( NfaState::Eps);
						m_nfa.st[s].out1 = k.start;
						Frag q; q.start = s; q.outs = k.outs; q.outs.push_back( std::make_pair( s, 2));
						tail = q;
					}
					if (tail.start >= 0) f = cat( f, tail);
				}
				return f;
			}
		}
		throw std::logic_error( "bad node");
	}
	Nfa& m_nfa;
};

// A document decoded to code points with the byte offset of every code point (ofs[n] = byte length).
struct DecodedText
{
	std::vector<uint32_t> cp;
	std::vector<uint32_t> ofs;

	void init( const unsigned char* src, std::size_t srclen)
	{
		cp.clear(); ofs.clear();
		std::size_t p = 0;
		while (p < srclen)
		{
			std::size_t len = 0;
			cp.push_back( decodeUtf8( src+p, srclen-p, len));
			ofs.push_back( (uint32_t)p);
			p += len;
		}
		ofs.push_back( (uint32_t)srclen);
	}
	std::size_t size() const {return cp.size();}
};

struct MatchContext
{
	const DecodedText* text;
	const CpSet* word;

	bool isWord( std::size_t i, bool next) const
	{
		// next=false: code point before position i, next=true: code point at position i
		if (next) return i < text->size() && word->contains( text->cp[i]);
		return i > 0 && word->contains( text->cp[i-1]);
	}
	bool holds( AssertKind k, std::size_t i) const
	{
		std::size_t n = text->size();
		switch (k)
		{
			case A_WORDB: return isWord( i, false) != isWord( i, true);
			case A_NWORDB: return isWord( i, false) == isWord( i, true);
			case A_BEGIN_TEXT: return i == 0;
			case A_END_TEXT: return i == n;
			case A_END_TEXT_OPT_LF: return i == n || (i+1 == n && text->cp[i] == '\n');
			case A_BEGIN_LINE: return i == 0 || text->cp[i-1] == '\n';
			case A_END_LINE: return i == n || text->cp[i] == '\n';
		}
		return false;
	}
};

// NFA simulation helper with generation-stamped visited marks
class NfaRunner
{
public:
	explicit NfaRunner( const Nfa& nfa_) :m_nfa(nfa_),m_mark(nfa_.st.size(),0),m_gen(0){}

	// closure of the states in `seeds` at text position i; fills `out` with Set states, returns accept flag
	bool closure( const std::vector<int>& seeds, std::size_t i, const MatchContext& ctx, std::vector<int>& out)
	{
		++m_gen;
		out.clear();
		bool accept = false;
		m_stack.clear();
		for (std::size_t k=0; k<seeds.size(); ++k) push( seeds[k]);
		while (!m_stack.empty())
		{
			int s = m_stack.back(); m_stack.pop_back();
			const NfaState& st = m_nfa.st[s];
			switch (st.kind)
			{
				case NfaState::Set: out.push_back( s); break;
				case NfaState::Accept: accept = true; break;
				case NfaState::Eps: push( st.out1); push( st.out2); break;
				case NfaState::Assert: if (ctx.holds( st.akind, i)) push( st.out1); break;
			}
		}
		return accept;
	}
	void step( const std::vector<int>& cur, uint32_t cp, std::vector<int>& seeds) const
	{
		seeds.clear();
		for (std::size_t k=0; k<cur.size(); ++k)
		{
			const NfaState& st = m_nfa.st[ cur[k]];
			if (m_nfa.sets[ st.setidx].contains( cp)) seeds.push_back( st.out1);
		}
	}
private:
	void push( int s)
	{
		if (s < 0 || m_mark[s] == m_gen) return;
		m_mark[s] = m_gen;
		m_stack.push_back( s);
	}
	const Nfa& m_nfa;
	std::vector<uint32_t> m_mark;
	uint32_t m_gen;
	std::vector<int> m_stack;
};

static void compileNfa( const NodeRef& ast, Nfa& nfa)
{
	NfaBuilder builder( nfa);
	builder.build( ast);
}

// true if [a,b) (code point positions) matches the NFA exactly, assertions seeing the real neighbours
static bool matchesSpan( const Nfa& nfa, NfaRunner& runner, const MatchContext& ctx, std::size_t a, std::size_t b)
{
	std::vector<int> seeds( 1, nfa.start), cur;
	for (std::size_t i=a;; ++i)
	{
		bool acc = runner.closure( seeds, i, ctx, cur);
		if (i == b) return acc;
		if (cur.empty()) return false;
		runner.step( cur, ctx.text->cp[i], seeds);
		if (seeds.empty()) return false;
	}
}

// ---------------------------------------------------------------- lexer definition
struct PatternDef
{
	std::string expression;
	uint32_t id;
	uint8_t posbind;
	uint8_t level;
	uint8_t resultidx;
	uint8_t editdist;
	uint8_t symtabref;

	Nfa nfa;			// whole expression, flags of the instance
	std::vector<bool> firstAscii;	// conservative filter: ASCII code points that can start a match (size 128)
	bool firstNonAscii;
	// capture selection (resultidx > 0): expression = A (G) B at top level, compiled WITHOUT the instance's
	// option flags (the reference compiles the libtre expression with REG_EXTENDED only, src/patternLexer.cpp:474)
	Nfa capWhole, capA, capGB, capG, capB;
};

struct MatchEvent
{
	uint32_t id;
	uint8_t level;
	uint8_t posbind;
	uint16_t origsize;
	uint32_t origpos;

	MatchEvent() :id(0),level(0),posbind(0),origsize(0),origpos(0){}
	MatchEvent( uint32_t id_, uint8_t level_, uint8_t posbind_, uint32_t origpos_, uint32_t origsize_)
		:id(id_),level(level_),posbind(posbind_),origsize((uint16_t)origsize_),origpos(origpos_){}
};

struct RawEvent {uint32_t idx, from, to;};	// idx = 1-based definition index (src/patternLexer.cpp:390)

struct SymbolTableDef
{
	std::unordered_map<std::string,uint32_t> map;	// name -> symbol id
};

static unsigned int extractEditDist( std::string& expr)
{
	// restates src/patternLexer.cpp:605-626: optional trailing "~N" (blanks allowed around it)
	if (expr.empty()) return 0;
	long se = (long)expr.size() - 1;
	while (se >= 0 && (unsigned char)expr[se] <= 32) --se;
	long de = se;
	while (se >= 0 && expr[se] >= '0' && expr[se] <= '9') --se;
	if (de == se) return 0;
	std::string digits = expr.substr( se+1, de-se);
	while (se >= 0 && (unsigned char)expr[se] <= 32) --se;
	if (se >= 0 && expr[se] == '~')
	{
		unsigned int rt = (unsigned int)atoi( digits.c_str());
		while (se > 0 && (unsigned char)expr[se-1] <= 32) --se;
		expr.resize( se);
		return rt;
	}
	return 0;
}

static void computeFirstSet( PatternDef& def)
{
	// conservative: which code points can be consumed first, ignoring assertions
	def.firstAscii.assign( 128, false);
	def.firstNonAscii = false;
	std::vector<int> stack( 1, def.nfa.start);
	std::vector<bool> seen( def.nfa.st.size(), false);
	while (!stack.empty())
	{
		int s = stack.back(); stack.pop_back();
		if (s < 0 || seen[s]) continue;
		seen[s] = true;
		const NfaState& st = def.nfa.st[s];
		switch (st.kind)
		{
			case NfaState::Set:
			{
				const CpSet& set = def.nfa.sets[ st.setidx];
				for (std::size_t ri=0; ri<set.r.size(); ++ri)
				{
					for (uint32_t c=set.r[ri].first; c<=set.r[ri].second && c<128; ++c) def.firstAscii[c] = true;
					if (set.r[ri].second >= 128) def.firstNonAscii = true;
				}
				break;
			}
			case NfaState::Accept: break;
			case NfaState::Eps: stack.push_back( st.out1); stack.push_back( st.out2); break;
			case NfaState::Assert: stack.push_back( st.out1); break;
		}
	}
}

static bool nullableNfa( const Nfa& nfa)
{
	// reachable accept through eps/assert edges only (any context)
	std::vector<int> stack( 1, nfa.start);
	std::vector<bool> seen( nfa.st.size(), false);
	while (!stack.empty())
	{
		int s = stack.back(); stack.pop_back();
		if (s < 0 || seen[s]) continue;
		seen[s] = true;
		const NfaState& st = nfa.st[s];
		if (st.kind == NfaState::Accept) return true;
		if (st.kind == NfaState::Eps) {stack.push_back( st.out1); stack.push_back( st.out2);}
		if (st.kind == NfaState::Assert) stack.push_back( st.out1);
	}
	return false;
}

struct LexerOracle
{
	std::vector<PatternDef> defs;
	std::map<uint32_t,uint8_t> idsymtabmap;
	std::vector<SymbolTableDef> symtabs;
	unsigned int flags;
	bool bytechar;
	bool compiled;
	CpSet word;
	std::string lasterror;

	// last batch output
	std::vector<sp_lexem> lexems;
	std::vector<uint64_t> offsets;
	std::vector<uint32_t> status;
	std::vector<RawEvent> rawevents;
	std::vector<uint64_t> rawoffsets;
	uint64_t nofRawEvents;

	LexerOracle() :flags(0),bytechar(false),compiled(false),nofRawEvents(0){}

	void defineLexem( uint32_t id, const std::string& expression_, unsigned int resultIndex, unsigned int level, unsigned int posbind)
	{
		if (compiled) throw std::runtime_error( "called define pattern after calling 'compile'");
		std::string expression( expression_);
		unsigned int editdist = extractEditDist( expression);
		if (id > MaxPatternId) throw std::runtime_error( "pattern id out of range");
		if (level > 255) throw std::runtime_error( "level out of range");
		if (resultIndex > 255) throw std::runtime_error( "result index out of range");
		if (editdist > 255) throw std::runtime_error( "edit distance out of range");
		PatternDef def;
		def.expression = expression; def.id = id; def.posbind = (uint8_t)posbind; def.level = (uint8_t)level;
		def.resultidx = (uint8_t)resultIndex; def.editdist = (uint8_t)editdist; def.symtabref = 0;
		def.firstNonAscii = false;
		defs.push_back( def);
	}
	void defineSymbol( uint32_t symbolid, uint32_t patternid, const std::string& name)
	{
		if (compiled) throw std::runtime_error( "called define pattern after calling 'compile'");
		if (patternid > MaxPatternId) throw std::runtime_error( "pattern id out of range");
		if (symbolid > MaxPatternId) throw std::runtime_error( "symbol id out of range");
		uint8_t ref;
		std::map<uint32_t,uint8_t>::const_iterator yi = idsymtabmap.find( patternid);
		if (yi == idsymtabmap.end())
		{
			if (symtabs.size() >= 255) throw std::runtime_error( "too many pattern symbol tables defined");
			symtabs.push_back( SymbolTableDef());
			ref = idsymtabmap[ patternid] = (uint8_t)symtabs.size();
		}
		else ref = yi->second;
		SymbolTableDef& st = symtabs[ ref-1];
		if (st.map.find( name) != st.map.end()) throw std::runtime_error( "symbol defined twice: '" + name + "'");
		st.map[ name] = symbolid;
	}
	void compile()
	{
		word = wordSet( flags);
		for (std::size_t di=0; di<defs.size(); ++di)
		{
			PatternDef& def = defs[di];
			if (def.editdist || bytechar) throw std::runtime_error( "unsupported: approximate matching (~N) and BYTECHAR are not restated by this oracle");
			try
			{
				Parser parser( def.expression, flags);
				NodeRef ast = parser.parse();
				def.nfa = Nfa();
				compileNfa( ast, def.nfa);
				if (!(flags & FLAG_ALLOWEMPTY) && nullableNfa( def.nfa)) throw std::runtime_error( "pattern matches empty buffer; use ALLOWEMPTY");
				computeFirstSet( def);
				if (def.resultidx) compileCapture( def);
			}
			catch (const std::runtime_error& e)
			{
				throw std::runtime_error( "failed to compile pattern \"" + def.expression + "\": " + e.what());
			}
			std::map<uint32_t,uint8_t>::const_iterator ti = idsymtabmap.find( def.id);
			if (ti != idsymtabmap.end()) def.symtabref = ti->second;
		}
		compiled = true;
	}
	static NodeRef catOf( const std::vector<NodeRef>& kids, std::size_t b, std::size_t e)
	{
		if (b >= e) return NodeRef( new Node( N_EMPTY));
		NodeRef rt( new Node( N_CAT));
		rt->kids.assign( kids.begin()+b, kids.begin()+e);
		return rt;
	}
	void compileCapture( PatternDef& def)
	{
		Parser parser( def.expression, 0/*no option flags reach libtre*/ | FLAG_DOTALL);
		NodeRef ast = parser.parse();
		if (def.resultidx > parser.nofCaptures()) throw std::runtime_error( "sub expression selection index out of range");
		std::vector<NodeRef> kids;
		if (ast->type == N_CAT) kids = ast->kids; else kids.push_back( ast);
		std::size_t gi = 0;
		for (; gi<kids.size(); ++gi) if (kids[gi]->type == N_GROUP && kids[gi]->capidx == def.resultidx) break;
		if (gi == kids.size()) throw std::runtime_error( "unsupported: the selected sub expression must be a top level group of the expression");
		compileNfa( ast, def.capWhole);
		compileNfa( catOf( kids, 0, gi), def.capA);
		compileNfa( catOf( kids, gi, kids.size()), def.capGB);
		compileNfa( catOf( kids, gi, gi+1), def.capG);
		compileNfa( catOf( kids, gi+1, kids.size()), def.capB);
	}

	// src/patternLexer.cpp:488-507 with libtre's POSIX leftmost-longest rules: the overall match starts at
	// `from` and is the longest one; the selected group starts as early and then ends as late as possible.
	bool matchSubExpression( const PatternDef& def, const MatchContext& ctx, std::size_t& fromi, std::size_t& toi) const
	{
		const DecodedText& text = *ctx.text;
		NfaRunner rw( def.capWhole), ra( def.capA), rgb( def.capGB), rg( def.capG), rb( def.capB);
		// longest overall match starting at fromi (the subject is the NUL terminated rest of the document)
		std::size_t best = NONE;
		{
			std::vector<int> seeds( 1, def.capWhole.start), cur;
			for (std::size_t i=fromi;; ++i)
			{
				bool acc = rw.closure( seeds, i, ctx, cur);
				if (acc) best = i;
				if (i == text.size() || cur.empty() || text.cp[i] == 0) break;
				rw.step( cur, text.cp[i], seeds);
				if (seeds.empty()) break;
			}
		}
		if (best == NONE) return false;
		std::size_t gs = NONE, ge = NONE;
		for (std::size_t s=fromi; s<=best && gs == NONE; ++s)
		{
			if (matchesSpan( def.capA, ra, ctx, fromi, s) && matchesSpan( def.capGB, rgb, ctx, s, best)) gs = s;
		}
		if (gs == NONE) return false;
		for (std::size_t e=best+1; e>gs && ge == NONE; --e)
		{
			if (matchesSpan( def.capG, rg, ctx, gs, e-1) && matchesSpan( def.capB, rb, ctx, e-1, best)) ge = e-1;
		}
		if (ge == NONE)
		{
			if (matchesSpan( def.capG, rg, ctx, gs, gs) && matchesSpan( def.capB, rb, ctx, gs, best)) ge = gs; else return false;
		}
		// mt.rm_eo > to - from -> reject (byte offsets relative to from)
		if (text.ofs[ge] - text.ofs[fromi] > text.ofs[toi] - text.ofs[fromi]) return false;
		fromi = gs; toi = ge;
		return true;
	}

	// The engine contract: all (idx, leftmost from, to) of one document in (to, idx) order
	void collectRawEvents( const DecodedText& text, std::vector<RawEvent>& events, std::vector<uint32_t>& fromOfEnd) const
	{
		MatchContext ctx; ctx.text = &text; ctx.word = &word;
		std::size_t n = text.size();
		fromOfEnd.assign( n+1, NONE);
		std::vector<int> seeds, cur;
		std::vector<std::size_t> touched;
		for (std::size_t di=0; di<defs.size(); ++di)
		{
			const PatternDef& def = defs[di];
			NfaRunner runner( def.nfa);
			touched.clear();
			for (std::size_t s=0; s<n; ++s)
			{
				uint32_t c0 = text.cp[s];
				if (c0 < 128 ? !def.firstAscii[c0] : !def.firstNonAscii) continue;
				seeds.assign( 1, def.nfa.start);
				for (std::size_t i=s;; ++i)
				{
					bool acc = runner.closure( seeds, i, ctx, cur);
					if (acc && i > s && fromOfEnd[i] == NONE) {fromOfEnd[i] = (uint32_t)s; touched.push_back( i);}
					if (i == n || cur.empty()) break;
					runner.step( cur, text.cp[i], seeds);
					if (seeds.empty()) break;
				}
			}
			for (std::size_t ti=0; ti<touched.size(); ++ti)
			{
				RawEvent ev; ev.idx = (uint32_t)di+1; ev.from = fromOfEnd[ touched[ti]]; ev.to = (uint32_t)touched[ti];
				events.push_back( ev);
				fromOfEnd[ touched[ti]] = NONE;
			}
		}
		std::sort( events.begin(), events.end(), []( const RawEvent& a, const RawEvent& b) {return a.to != b.to ? a.to < b.to : a.idx < b.idx;});
	}

	// src/patternLexer.cpp:717-826, one call per raw event. Returns false on the "size out of range" error.
	bool handleEvent( const RawEvent& rev, const DecodedText& text, const unsigned char* src, std::vector<MatchEvent>& ar) const
	{
		MatchContext ctx; ctx.text = &text; ctx.word = &word;
		const PatternDef& def = defs[ rev.idx-1];
		std::size_t fromi = rev.from, toi = rev.to;
		if (text.ofs[toi] - text.ofs[fromi] >= 65535) return false;
		if (def.resultidx)
		{
			if (!matchSubExpression( def, ctx, fromi, toi)) return true;
		}
		uint32_t from = text.ofs[fromi], to = text.ofs[toi];
		uint32_t patternid = def.id;
		if (def.symtabref)
		{
			const SymbolTableDef& st = symtabs[ def.symtabref-1];
			std::unordered_map<std::string,uint32_t>::const_iterator si = st.map.find( std::string( (const char*)src + from, to - from));
			if (si != st.map.end() && si->second) patternid = si->second;
		}
		MatchEvent ev( def.id, def.level, def.posbind, from, to - from);
		MatchEvent symev( patternid, def.level, def.posbind, from, to - from);
		bool withSymbol = (patternid != def.id);
		if (ar.empty())
		{
			ar.push_back( ev);
			if (withSymbol) ar.push_back( symev);
			return true;
		}
		// supersede: drop older elements of the tail that are replaced by or covered by the new one
		uint32_t evEnd = ev.origpos + ev.origsize;
		std::size_t nofDeletes = 0;
		for (std::size_t k=ar.size(); k>0 && ar[k-1].origpos >= ev.origpos; --k)
		{
			const MatchEvent& old = ar[k-1];
			bool sameElement = (old.id == ev.id && old.origpos == ev.origpos && old.level == ev.level);
			bool coveredByHigher = (old.level < ev.level && old.origpos + old.origsize <= evEnd);
			if (sameElement || coveredByHigher)
			{
				ar.erase( ar.begin() + (k-1));
				++nofDeletes;
			}
		}
		if (!nofDeletes)
		{
			// the new element is ignored if an element of higher level in the tail covers it; the walk ends
			// at the first element (from the back) that ends before the new one
			for (std::size_t k=ar.size(); k>0 && ar[k-1].origpos + ar[k-1].origsize >= evEnd; --k)
			{
				if (ar[k-1].level > ev.level && ar[k-1].origpos <= ev.origpos) return true;
			}
		}
		if (!withSymbol)
		{
			// stable insert behind the last element starting at or before the new one
			std::size_t k = ar.size();
			ar.push_back( MatchEvent());
			while (k > 0 && ar[k-1].origpos > ev.origpos) {ar[k] = ar[k-1]; --k;}
			ar[k] = ev;
		}
		else
		{
			// insert of two elements; QUIRK kept: the first later-starting element moves by two places,
			// every further one only by one (src/patternLexer.cpp:809-820)
			std::size_t oldsize = ar.size();
			ar.resize( oldsize + 2);
			std::size_t dst = oldsize + 1;		// index written next
			std::size_t k = oldsize;		// elements [0,k) not yet visited
			while (k > 0 && ar[k-1].origpos > ev.origpos)
			{
				ar[dst] = ar[k-1];
				dst = k-1;
				--k;
			}
			ar[k] = ev;
			ar[k+1] = symev;
		}
		return true;
	}

	// src/patternLexer.cpp:893-945
	static void assignOrdpos( const std::vector<MatchEvent>& ar, std::vector<sp_lexem>& out)
	{
		std::size_t base = out.size();
		std::size_t mi = 0, me = ar.size();
		uint32_t ordpos = 0, origpos = 0;
		uint8_t lastposbind = SP_BIND_CONTENT;
		auto emit = [&]( const MatchEvent& e, uint32_t pos) {
			sp_lexem lx; lx.id = e.id; lx.ordpos = pos; lx.origpos = e.origpos; lx.origsize = e.origsize;
			out.push_back( lx);
		};
		for (; mi != me; ++mi)
		{
			lastposbind = ar[mi].posbind;
			if (ar[mi].posbind == SP_BIND_UNIQUE || ar[mi].posbind == SP_BIND_CONTENT)
			{
				ordpos = 1; origpos = ar[mi].origpos;
				emit( ar[mi], 1);
				++mi;
				break;
			}
			else if (ar[mi].posbind == SP_BIND_SUCCESSOR) emit( ar[mi], 1);
		}
		if (ordpos == 0) out.resize( base);
		for (; mi != me; ++mi)
		{
			switch (ar[mi].posbind)
			{
				case SP_BIND_UNIQUE:
					if (lastposbind == SP_BIND_UNIQUE) break;
					/*no break here!*/
				case SP_BIND_CONTENT:
					if (ar[mi].origpos > origpos) {origpos = ar[mi].origpos; ++ordpos;}
					emit( ar[mi], ordpos);
					break;
				case SP_BIND_SUCCESSOR: emit( ar[mi], ordpos+1); break;
				case SP_BIND_PREDECESSOR: emit( ar[mi], ordpos); break;
			}
			lastposbind = ar[mi].posbind;
		}
	}

	// PatternLexerContext::match, src/patternLexer.cpp:858-950
	uint32_t matchDocument( const unsigned char* src, std::size_t srclen, std::vector<sp_lexem>& out, std::vector<RawEvent>* rawout, uint64_t& nraw) const
	{
		if (srclen >= 0xFFFFFFFFull) return (uint32_t)-SP_ERR_RUNTIME;
		DecodedText text; text.init( src, srclen);
		std::vector<RawEvent> events;
		std::vector<uint32_t> scratch;
		collectRawEvents( text, events, scratch);
		nraw += events.size();
		if (rawout)
		{
			for (std::size_t i=0; i<events.size(); ++i)
			{
				RawEvent e; e.idx = events[i].idx; e.from = text.ofs[ events[i].from]; e.to = text.ofs[ events[i].to];
				rawout->push_back( e);
			}
		}
		std::vector<MatchEvent> ar;
		ar.reserve( srclen / 4 + 10);
		for (std::size_t i=0; i<events.size(); ++i)
		{
			if (!handleEvent( events[i], text, src, ar)) return (uint32_t)-SP_ERR_RUNTIME;
		}
		assignOrdpos( ar, out);
		return 0;
	}
};

struct DocResult {std::vector<sp_lexem> lexems; std::vector<RawEvent> raw; uint32_t status; uint64_t nraw;};

} // namespace

extern "C" {

void* lxo_create() {return new LexerOracle();}
void lxo_destroy( void* h) {delete (LexerOracle*)h;}
const char* lxo_last_error( void* h) {return ((LexerOracle*)h)->lasterror.c_str();}

int lxo_define_option( void* h, const char* name, double)
{
	LexerOracle* lx = (LexerOracle*)h;
	std::string n( name);
	for (std::size_t i=0; i<n.size(); ++i) n[i] = (char)toupper( (unsigned char)n[i]);
	if (n == "CASELESS") lx->flags |= FLAG_CASELESS;
	else if (n == "DOTALL") lx->flags |= FLAG_DOTALL;
	else if (n == "MULTILINE") lx->flags |= FLAG_MULTILINE;
	else if (n == "ALLOWEMPTY") lx->flags |= FLAG_ALLOWEMPTY;
	else if (n == "UCP") lx->flags |= FLAG_UCP;
	else if (n == "BYTECHAR") lx->bytechar = true;
	else {lx->lasterror = std::string("unknown option '") + name + "'"; return -1;}
	return 0;
}
int lxo_define_lexem( void* h, uint32_t id, const char* expr, size_t exprlen, uint32_t resultIndex, uint32_t level, uint32_t posbind)
{
	LexerOracle* lx = (LexerOracle*)h;
	try {lx->defineLexem( id, std::string( expr, exprlen), resultIndex, level, posbind); return 0;}
	catch (const std::exception& e) {lx->lasterror = e.what(); return -1;}
}
int lxo_define_symbol( void* h, uint32_t symbolid, uint32_t lexemid, const char* name, size_t namelen)
{
	LexerOracle* lx = (LexerOracle*)h;
	try {lx->defineSymbol( symbolid, lexemid, std::string( name, namelen)); return 0;}
	catch (const std::exception& e) {lx->lasterror = e.what(); return -1;}
}
int lxo_compile( void* h)
{
	LexerOracle* lx = (LexerOracle*)h;
	try {lx->compile(); return 0;}
	catch (const std::exception& e) {lx->lasterror = e.what(); return -1;}
}
// keep_raw != 0: also keep the raw engine events (idx, from, to) of every document
int lxo_scan_batch( void* h, const uint8_t* text, const uint64_t* doc_offsets, uint32_t ndocs, int nthreads, int keep_raw)
{
	LexerOracle* lx = (LexerOracle*)h;
	if (!lx->compiled) {lx->lasterror = "called scan without calling 'compile'"; return -1;}
	if (nthreads < 1) nthreads = 1;
	std::vector<DocResult> res( ndocs);
	std::vector<std::thread> threads;
	// interleaved assignment: documents differ in size
	for (int t=0; t<nthreads; ++t)
	{
		threads.push_back( std::thread( [=,&res]() {
			for (uint32_t d=t; d<ndocs; d+=nthreads)
			{
				res[d].nraw = 0;
				res[d].status = lx->matchDocument( text + doc_offsets[d], doc_offsets[d+1]-doc_offsets[d], res[d].lexems, keep_raw?&res[d].raw:0, res[d].nraw);
				if (res[d].status) res[d].lexems.clear();
			}
		}));
	}
	for (std::size_t t=0; t<threads.size(); ++t) threads[t].join();
	lx->lexems.clear(); lx->offsets.assign( 1, 0); lx->status.clear();
	lx->rawevents.clear(); lx->rawoffsets.assign( 1, 0); lx->nofRawEvents = 0;
	for (uint32_t d=0; d<ndocs; ++d)
	{
		lx->lexems.insert( lx->lexems.end(), res[d].lexems.begin(), res[d].lexems.end());
		lx->offsets.push_back( lx->lexems.size());
		lx->status.push_back( res[d].status);
		lx->rawevents.insert( lx->rawevents.end(), res[d].raw.begin(), res[d].raw.end());
		lx->rawoffsets.push_back( lx->rawevents.size());
		lx->nofRawEvents += res[d].nraw;
	}
	return 0;
}
uint64_t lxo_result_count( void* h) {return ((LexerOracle*)h)->lexems.size();}
uint64_t lxo_raw_event_count( void* h) {return ((LexerOracle*)h)->nofRawEvents;}
int lxo_fetch( void* h, sp_lexem* lexems, uint64_t* offsets, uint32_t* status)
{
	LexerOracle* lx = (LexerOracle*)h;
	if (lexems && !lx->lexems.empty()) std::memcpy( lexems, lx->lexems.data(), lx->lexems.size()*sizeof(sp_lexem));
	if (offsets) std::memcpy( offsets, lx->offsets.data(), lx->offsets.size()*sizeof(uint64_t));
	if (status && !lx->status.empty()) std::memcpy( status, lx->status.data(), lx->status.size()*sizeof(uint32_t));
	return 0;
}
// raw events as (idx, from, to) uint32 triples with byte offsets; offsets[ndocs+1]
int lxo_fetch_raw( void* h, uint32_t* triples, uint64_t* offsets)
{
	LexerOracle* lx = (LexerOracle*)h;
	if (triples && !lx->rawevents.empty()) std::memcpy( triples, lx->rawevents.data(), lx->rawevents.size()*sizeof(RawEvent));
	if (offsets) std::memcpy( offsets, lx->rawoffsets.data(), lx->rawoffsets.size()*sizeof(uint64_t));
	return 0;
}

} // extern "C"
