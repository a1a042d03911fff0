// Back end of the B200 lexer compiler: syntax trees -> byte level NFA (UTF-8 decode folded in) -> DFA tables.
//
// Replaces the automaton construction of hs_compile_ext_multi (reference call site
// src/patternLexer.cpp:1083-1086) with flags HS_FLAG_UTF8|HS_FLAG_SOM_LEFTMOST (:403):
//  * code point sets are expanded into UTF-8 byte sequence automata, so the device scans raw bytes and no
//    separate decode pass exists (the decode of src/unicodeUtils.cpp is "fused in");
//  * a FORWARD dfa over all lexem definitions of a group reports (definition, end offset) for every end
//    offset (Hyperscan block mode semantics);
//  * one REVERSE dfa per definition, run backwards from an end offset, yields the leftmost start
//    (HS_FLAG_SOM_LEFTMOST);
//  * assertions (\b \B ^ $ \A \z) are resolved inside the subset construction: a dfa state carries the kind
//    of the previously consumed byte, and a match ending before byte i is reported on the transition that
//    consumes byte i (or the virtual end-of-text symbol), when the kind of the next byte is known.
#ifndef SP_AUTOMATON_HPP_INCLUDED
#define SP_AUTOMATON_HPP_INCLUDED
#include "regex.hpp"
#include <map>
#include <unordered_map>
#include <set>

namespace sp {

// ------------------------------------------------------------------ UTF-8 byte sequences of a code point set
struct ByteSeq {int n; uint8_t lo[4], hi[4];};

inline void utf8Encode( uint32_t cp, int n, uint8_t* out)
{
	switch (n)
	{
		case 1: out[0] = (uint8_t)cp; break;
		case 2: out[0] = 0xC0 | (cp >> 6); out[1] = 0x80 | (cp & 0x3F); break;
		case 3: out[0] = 0xE0 | (cp >> 12); out[1] = 0x80 | ((cp >> 6) & 0x3F); out[2] = 0x80 | (cp & 0x3F); break;
		default: out[0] = 0xF0 | (cp >> 18); out[1] = 0x80 | ((cp >> 12) & 0x3F); out[2] = 0x80 | ((cp >> 6) & 0x3F); out[3] = 0x80 | (cp & 0x3F); break;
	}
}
inline void utf8SplitRange( uint32_t lo, uint32_t hi, int n, std::vector<ByteSeq>& out)
{
	// all code points of [lo,hi] have n bytes; split until every byte position is an independent range
	for (int i=1; i<n; ++i)
	{
		uint32_t m = (1u << (6*i)) - 1;
		if ((lo & ~m) != (hi & ~m))
		{
			if ((lo & m) != 0) {utf8SplitRange( lo, lo|m, n, out); utf8SplitRange( (lo|m)+1, hi, n, out); return;}
			if ((hi & m) != m) {utf8SplitRange( lo, (hi & ~m)-1, n, out); utf8SplitRange( hi & ~m, hi, n, out); return;}
		}
	}
	ByteSeq s; s.n = n;
	utf8Encode( lo, n, s.lo); utf8Encode( hi, n, s.hi);
	out.push_back( s);
}
inline std::vector<ByteSeq> utf8Sequences( const CodeSet& set)
{
	static const uint32_t lim[5] = {0, 0x7F, 0x7FF, 0xFFFF, 0x10FFFF};
	std::vector<ByteSeq> out;
	for (const CodeSet::Iv& iv: set.intervals())
	{
		for (int n=1; n<=4; ++n)
		{
			uint32_t lo = std::max( iv.lo, n == 1 ? 0u : lim[n-1]+1), hi = std::min( iv.hi, lim[n]);
			if (lo > hi) continue;
			if (n == 3)
			{
				// surrogates have no UTF-8 encoding
				if (lo <= 0xD7FF) utf8SplitRange( lo, std::min( hi, 0xD7FFu), n, out);
				if (hi >= 0xE000) utf8SplitRange( std::max( lo, 0xE000u), hi, n, out);
			}
			else utf8SplitRange( lo, hi, n, out);
		}
	}
	return out;
}

// ------------------------------------------------------------------ byte level NFA
enum ByteKind {KindOther=0, KindWord=1, KindNewline=2, KindEdge=3};	// KindEdge: begin (as previous) / end (as next) of the text
inline ByteKind kindOfByte( unsigned int b)
{
	if (b == '\n') return KindNewline;
	if ((b >= '0' && b <= '9') || (b >= 'A' && b <= 'Z') || (b >= 'a' && b <= 'z') || b == '_') return KindWord;
	return KindOther;
}
inline bool assertionHolds( AssertKind a, int prev, int next)
{
	switch (a)
	{
		case AsWordBoundary: return (prev == KindWord) != (next == KindWord);
		case AsNotWordBoundary: return (prev == KindWord) == (next == KindWord);
		case AsTextBegin: return prev == KindEdge;
		case AsTextEnd: return next == KindEdge;
		case AsLineBegin: return prev == KindEdge || prev == KindNewline;
		case AsLineEnd: return next == KindEdge || next == KindNewline;
		case AsTextEndOptNewline: break;
	}
	return false;
}

struct ByteNfa
{
	struct Edge {uint8_t lo, hi; int to;};
	struct State
	{
		std::vector<Edge> edges;
		std::vector<int> eps;
		int assertion;		// AssertKind or -1
		int assertTo;
		int accept;		// definition index or -1
		State() :assertion(-1),assertTo(-1),accept(-1){}
	};
	std::vector<State> st;

	int add() {st.push_back( State()); if (st.size() > 400000) throw UnsupportedError( "expression too large"); return (int)st.size()-1;}
};

// Acyclic automaton of one code point set, shared by all its occurrences; state 0 = source, state 1 = sink.
struct ClassGraph
{
	struct Node {std::vector<ByteNfa::Edge> edges;};
	std::vector<Node> nodes;
};

class ClassGraphCache
{
public:
	const ClassGraph& get( const CodeSet& set, bool reversed)
	{
		std::map<CodeSet,ClassGraph>& cache = reversed ? m_rev : m_fwd;
		std::map<CodeSet,ClassGraph>::iterator it = cache.find( set);
		if (it != cache.end()) return it->second;
		return cache[ set] = build( set, reversed);
	}
private:
	struct Trie {std::vector<std::vector<std::pair<std::pair<int,int>,int> > > kids;};	// node -> [((lo,hi), child)]

	static int canonical( const Trie& t, int node, std::map<std::vector<int>,int>& sigmap, ClassGraph& g)
	{
		if (t.kids[node].empty()) return 1;
		std::vector<int> sig;
		std::vector<ByteNfa::Edge> edges;
		for (std::size_t k=0; k<t.kids[node].size(); ++k)
		{
			int child = canonical( t, t.kids[node][k].second, sigmap, g);
			ByteNfa::Edge e; e.lo = (uint8_t)t.kids[node][k].first.first; e.hi = (uint8_t)t.kids[node][k].first.second; e.to = child;
			edges.push_back( e);
		}
		std::sort( edges.begin(), edges.end(), []( const ByteNfa::Edge& a, const ByteNfa::Edge& b) {return a.lo != b.lo ? a.lo < b.lo : a.hi != b.hi ? a.hi < b.hi : a.to < b.to;});
		// merge adjacent ranges that lead to the same node
		std::vector<ByteNfa::Edge> merged;
		for (const ByteNfa::Edge& e: edges)
		{
			if (!merged.empty() && merged.back().to == e.to && (int)merged.back().hi + 1 >= (int)e.lo) merged.back().hi = std::max( merged.back().hi, e.hi);
			else merged.push_back( e);
		}
		for (const ByteNfa::Edge& e: merged) {sig.push_back( e.lo); sig.push_back( e.hi); sig.push_back( e.to);}
		if (node == 0) {g.nodes[0].edges = merged; return 0;}
		std::map<std::vector<int>,int>::iterator it = sigmap.find( sig);
		if (it != sigmap.end()) return it->second;
		g.nodes.push_back( ClassGraph::Node());
		g.nodes.back().edges = merged;
		return sigmap[ sig] = (int)g.nodes.size()-1;
	}
	static ClassGraph build( const CodeSet& set, bool reversed)
	{
		std::vector<ByteSeq> seqs = utf8Sequences( set);
		Trie t; t.kids.resize( 1);
		for (const ByteSeq& s: seqs)
		{
			int node = 0;
			for (int k=0; k<s.n; ++k)
			{
				int p = reversed ? s.n-1-k : k;
				std::pair<int,int> key( s.lo[p], s.hi[p]);
				int next = -1;
				for (auto& kid: t.kids[node]) if (kid.first == key) {next = kid.second; break;}
				if (next < 0)
				{
					next = (int)t.kids.size();
					t.kids.push_back( std::vector<std::pair<std::pair<int,int>,int> >());
					t.kids[node].push_back( std::make_pair( key, next));
				}
				node = next;
			}
		}
		ClassGraph g; g.nodes.resize( 2);
		std::map<std::vector<int>,int> sigmap;
		canonical( t, 0, sigmap, g);
		return g;
	}
	std::map<CodeSet,ClassGraph> m_fwd, m_rev;
};

inline bool containsAssertion( const RxRef& rx, AssertKind* which=0)
{
	if (rx->kind == RxAssert) {if (which) *which = rx->assertion; return true;}
	for (const RxRef& s: rx->sub) if (containsAssertion( s, which)) return true;
	return false;
}
inline bool containsWordAssertion( const RxRef& rx)
{
	if (rx->kind == RxAssert && (rx->assertion == AsWordBoundary || rx->assertion == AsNotWordBoundary)) return true;
	for (const RxRef& s: rx->sub) if (containsWordAssertion( s)) return true;
	return false;
}

// Thompson construction; every fragment has one entry and one exit state.
class NfaBuilder
{
public:
	NfaBuilder( ByteNfa& nfa, ClassGraphCache& cache, bool reversed) :m_nfa(nfa),m_cache(cache),m_reversed(reversed){}

	// adds the automaton of `rx`; returns its entry state; the exit state is marked accepting for `accept`
	int addPattern( const RxRef& rx, int accept)
	{
		Frag f = gen( rx);
		m_nfa.st[ f.out].accept = accept;
		return f.in;
	}
private:
	struct Frag {int in, out;};

	Frag gen( const RxRef& rx)
	{
		Frag f;
		switch (rx->kind)
		{
			case RxEmpty:
				f.in = f.out = m_nfa.add();
				return f;
			case RxSet:
			{
				f.in = m_nfa.add(); f.out = m_nfa.add();
				const ClassGraph& g = m_cache.get( rx->set, m_reversed);
				std::vector<int> map( g.nodes.size(), -1);
				map[0] = f.in; map[1] = f.out;
				for (std::size_t n=2; n<g.nodes.size(); ++n) map[n] = m_nfa.add();
				for (std::size_t n=0; n<g.nodes.size(); ++n)
				{
					for (const ByteNfa::Edge& e: g.nodes[n].edges)
					{
						ByteNfa::Edge ne = e; ne.to = map[ e.to];
						m_nfa.st[ map[n]].edges.push_back( ne);
					}
				}
				return f;
			}
			case RxAssert:
			{
				AssertKind a = rx->assertion;
				if (a == AsTextEndOptNewline) throw UnsupportedError( "'$' without MULTILINE and \\Z are not supported (use \\z)");
				if (m_reversed)
				{
					if (a == AsTextBegin) a = AsTextEnd; else if (a == AsTextEnd) a = AsTextBegin;
					else if (a == AsLineBegin) a = AsLineEnd; else if (a == AsLineEnd) a = AsLineBegin;
				}
				f.in = m_nfa.add(); f.out = m_nfa.add();
				m_nfa.st[ f.in].assertion = a; m_nfa.st[ f.in].assertTo = f.out;
				return f;
			}
			case RxGroup:
				return gen( rx->sub[0]);
			case RxConcat:
			{
				std::size_t n = rx->sub.size();
				f = gen( rx->sub[ m_reversed ? n-1 : 0]);
				for (std::size_t k=1; k<n; ++k)
				{
					Frag g = gen( rx->sub[ m_reversed ? n-1-k : k]);
					m_nfa.st[ f.out].eps.push_back( g.in);
					f.out = g.out;
				}
				return f;
			}
			case RxChoice:
			{
				f.in = m_nfa.add(); f.out = m_nfa.add();
				for (const RxRef& alt: rx->sub)
				{
					Frag g = gen( alt);
					m_nfa.st[ f.in].eps.push_back( g.in);
					m_nfa.st[ g.out].eps.push_back( f.out);
				}
				return f;
			}
			case RxLoop:
			{
				f.in = f.out = m_nfa.add();
				for (int k=0; k<rx->lo; ++k)
				{
					Frag g = gen( rx->sub[0]);
					m_nfa.st[ f.out].eps.push_back( g.in);
					f.out = g.out;
				}
				if (rx->hi < 0)
				{
					Frag g = gen( rx->sub[0]);
					int hub = m_nfa.add(), exit = m_nfa.add();
					m_nfa.st[ f.out].eps.push_back( hub);
					m_nfa.st[ hub].eps.push_back( g.in);
					m_nfa.st[ hub].eps.push_back( exit);
					m_nfa.st[ g.out].eps.push_back( hub);
					f.out = exit;
				}
				else if (rx->hi > rx->lo)
				{
					int exit = m_nfa.add();
					for (int k=rx->lo; k<rx->hi; ++k)
					{
						Frag g = gen( rx->sub[0]);
						m_nfa.st[ f.out].eps.push_back( exit);
						m_nfa.st[ f.out].eps.push_back( g.in);
						f.out = g.out;
					}
					m_nfa.st[ f.out].eps.push_back( exit);
					f.out = exit;
				}
				return f;
			}
		}
		throw std::logic_error( "bad syntax tree");
	}
	ByteNfa& m_nfa;
	ClassGraphCache& m_cache;
	bool m_reversed;
};

// true if the expression can match the empty string in some context
inline bool nfaNullable( const ByteNfa& nfa, int entry)
{
	std::vector<int> stack( 1, entry);
	std::vector<char> seen( nfa.st.size(), 0);
	while (!stack.empty())
	{
		int s = stack.back(); stack.pop_back();
		if (seen[s]) continue;
		seen[s] = 1;
		const ByteNfa::State& st = nfa.st[s];
		if (st.accept >= 0) return true;
		for (int e: st.eps) stack.push_back( e);
		if (st.assertion >= 0) stack.push_back( st.assertTo);
	}
	return false;
}

// ------------------------------------------------------------------ subset construction
struct Dfa
{
	uint32_t nstates;
	uint32_t nclasses;			// real byte classes; the virtual end-of-text symbol is class `nclasses`
	uint8_t byteclass[256];
	std::vector<uint32_t> next;		// [nstates][nclasses+1]
	std::vector<std::vector<uint32_t> > accepts;	// per state: definitions whose match ended in front of the byte that led here
	uint32_t init[4];			// initial state by kind of the previous byte (ByteKind)
	uint32_t dead;				// anchored mode: state with no future (else 0xFFFFFFFF)
	std::vector<uint8_t> cut;		// unanchored mode, per state: 1 if no thread that was alive in front of the byte that
						// led here survived it, i.e. every match reported later starts at that byte or behind it

	uint32_t stride() const {return nclasses+1;}
};

struct DfaOptions
{
	bool unanchored;	// start states are re-entered at every position (forward multi pattern scan)
	bool allowEmpty;	// report matches of length 0 (anchored helper automata of the capture selection)
	bool markCuts;		// unanchored: distinguish states reached by a transition no older thread survived (Dfa::cut)
	uint32_t maxStates;
	DfaOptions() :unanchored(true),allowEmpty(false),markCuts(false),maxStates(60000){}
};

struct DfaTooLarge :public std::runtime_error
{
	DfaTooLarge() :std::runtime_error("automaton has too many states"){}
};

class DfaBuilder
{
public:
	DfaBuilder( const ByteNfa& nfa, const std::vector<int>& entries, const DfaOptions& opt)
		:m_nfa(nfa),m_entries(entries),m_opt(opt),m_stamp(nfa.st.size(),0),m_gen(0){}

	Dfa build()
	{
		Dfa dfa;
		makeRuns();
		dfa.dead = 0xFFFFFFFFu;
		std::vector<int> key;
		if (m_opt.unanchored)
		{
			key.assign( 3, 0); key[0] = KindEdge;	// [prevkind, initial flag, #accepts]
			dfa.init[0] = dfa.init[1] = dfa.init[2] = dfa.init[3] = intern( key);
		}
		else
		{
			key.assign( 3, 0); key[0] = 0; key[1] = 0;		// dead state first
			dfa.dead = intern( key);
			for (int pk=0; pk<4; ++pk)
			{
				key.assign( 3, 0); key[0] = pk; key[1] = 1;
				for (int e: m_entries) key.push_back( e);
				std::sort( key.begin()+3, key.end());
				dfa.init[pk] = intern( key);
			}
		}
		std::vector<uint32_t> rows;	// [state][run]
		std::vector<int> closure, target, acc;
		for (uint32_t s=0; s<m_keys.size(); ++s)
		{
			const std::vector<int> cur = m_keys[s];	// copy: m_keys grows
			int pk = cur[0]; bool initial = cur[1] == 1; int nacc = cur[2];
			std::vector<int> base( cur.begin()+3+nacc, cur.end());
			for (std::size_t r=0; r<=m_runs.size(); ++r)
			{
				bool eot = (r == m_runs.size());
				int nk = eot ? (int)KindEdge : (int)kindOfByte( m_runs[r].first);
				acc.clear();
				expand( base, pk, nk, closure, (initial && !m_opt.allowEmpty) ? 0 : &acc);
				std::size_t nold = closure.size();
				if (m_opt.unanchored && !eot)
				{
					std::vector<int> fresh;
					expand( m_entries, pk, nk, fresh, m_opt.allowEmpty ? &acc : 0);
					closure.insert( closure.end(), fresh.begin(), fresh.end());
				}
				target.clear();
				bool oldSurvives = false;
				if (!eot)
				{
					unsigned int b = m_runs[r].first;
					for (std::size_t ci=0; ci<closure.size(); ++ci)
					{
						for (const ByteNfa::Edge& e: m_nfa.st[ closure[ci]].edges) if (e.lo <= b && b <= e.hi) {target.push_back( e.to); if (ci < nold) oldSurvives = true;}
					}
					std::sort( target.begin(), target.end());
					target.erase( std::unique( target.begin(), target.end()), target.end());
				}
				std::sort( acc.begin(), acc.end());
				acc.erase( std::unique( acc.begin(), acc.end()), acc.end());
				key.clear();
				bool noFuture = target.empty() && (!m_opt.unanchored || eot);
				key.push_back( (noFuture && acc.empty()) ? 0 : (eot ? (int)KindEdge : nk));
				key.push_back( (m_opt.unanchored && m_opt.markCuts && !eot && !oldSurvives) ? 2 : 0);
				key.push_back( (int)acc.size());
				key.insert( key.end(), acc.begin(), acc.end());
				key.insert( key.end(), target.begin(), target.end());
				if (noFuture && !acc.empty()) key[0] = 0;	// kind is irrelevant for a state without successors
				rows.push_back( intern( key));
			}
			if (m_keys.size() > m_opt.maxStates) throw DfaTooLarge();
		}
		// merge byte runs with identical columns into classes
		uint32_t nruns = (uint32_t)m_runs.size(), nstates = (uint32_t)m_keys.size();
		std::map<std::vector<uint32_t>,uint32_t> colmap;
		std::vector<uint32_t> runclass( nruns);
		std::vector<uint32_t> repr;
		for (uint32_t r=0; r<nruns; ++r)
		{
			std::vector<uint32_t> col( nstates);
			for (uint32_t s=0; s<nstates; ++s) col[s] = rows[ (std::size_t)s*(nruns+1) + r];
			std::map<std::vector<uint32_t>,uint32_t>::iterator it = colmap.find( col);
			if (it == colmap.end()) {it = colmap.insert( std::make_pair( col, (uint32_t)repr.size())).first; repr.push_back( r);}
			runclass[r] = it->second;
		}
		dfa.nstates = nstates;
		dfa.nclasses = (uint32_t)repr.size();
		for (uint32_t r=0; r<nruns; ++r)
		{
			for (unsigned int b=m_runs[r].first; b<=m_runs[r].second; ++b) dfa.byteclass[b] = (uint8_t)runclass[r];
		}
		if (dfa.nclasses > 255) throw UnsupportedError( "too many byte classes");
		dfa.next.resize( (std::size_t)nstates * dfa.stride());
		for (uint32_t s=0; s<nstates; ++s)
		{
			for (uint32_t c=0; c<dfa.nclasses; ++c) dfa.next[ (std::size_t)s*dfa.stride() + c] = rows[ (std::size_t)s*(nruns+1) + repr[c]];
			dfa.next[ (std::size_t)s*dfa.stride() + dfa.nclasses] = rows[ (std::size_t)s*(nruns+1) + nruns];
		}
		dfa.accepts.resize( nstates);
		dfa.cut.assign( nstates, 0);
		for (uint32_t s=0; s<nstates; ++s)
		{
			const std::vector<int>& k = m_keys[s];
			dfa.accepts[s].assign( k.begin()+3, k.begin()+3+k[2]);
			dfa.cut[s] = (k[1] == 2);
		}
		return dfa;
	}

private:
	void makeRuns()
	{
		bool cut[ 257];
		std::memset( cut, 0, sizeof(cut));
		cut[0] = true;
		for (unsigned int b=1; b<256; ++b) if (kindOfByte( b) != kindOfByte( b-1)) cut[b] = true;
		for (const ByteNfa::State& st: m_nfa.st)
		{
			for (const ByteNfa::Edge& e: st.edges) {cut[ e.lo] = true; cut[ (unsigned int)e.hi+1] = true;}
		}
		for (unsigned int b=0; b<256;)
		{
			unsigned int e = b+1;
			while (e < 256 && !cut[e]) ++e;
			m_runs.push_back( std::make_pair( b, e-1));
			b = e;
		}
	}
	uint32_t intern( const std::vector<int>& key)
	{
		std::map<std::vector<int>,uint32_t>::iterator it = m_index.find( key);
		if (it != m_index.end()) return it->second;
		uint32_t id = (uint32_t)m_keys.size();
		m_keys.push_back( key);
		m_index[ key] = id;
		return id;
	}
	// epsilon/assertion closure of `seeds` between a previous byte of kind `pk` and a next byte of kind `nk`:
	// `out` receives the states with byte edges, `acc` (if given) the accepting definitions reached
	void expand( const std::vector<int>& seeds, int pk, int nk, std::vector<int>& out, std::vector<int>* acc)
	{
		++m_gen;
		out.clear();
		m_stack.clear();
		for (int s: seeds) if (m_stamp[s] != m_gen) {m_stamp[s] = m_gen; m_stack.push_back( s);}
		while (!m_stack.empty())
		{
			int s = m_stack.back(); m_stack.pop_back();
			const ByteNfa::State& st = m_nfa.st[s];
			if (!st.edges.empty()) out.push_back( s);
			if (st.accept >= 0 && acc) acc->push_back( st.accept);
			for (int e: st.eps) if (m_stamp[e] != m_gen) {m_stamp[e] = m_gen; m_stack.push_back( e);}
			if (st.assertion >= 0 && assertionHolds( (AssertKind)st.assertion, pk, nk) && m_stamp[ st.assertTo] != m_gen)
			{
				m_stamp[ st.assertTo] = m_gen; m_stack.push_back( st.assertTo);
			}
		}
	}

	const ByteNfa& m_nfa;
	std::vector<int> m_entries;
	DfaOptions m_opt;
	std::vector<std::pair<unsigned int,unsigned int> > m_runs;
	std::vector<std::vector<int> > m_keys;
	std::map<std::vector<int>,uint32_t> m_index;
	std::vector<uint32_t> m_stamp;
	uint32_t m_gen;
	std::vector<int> m_stack;
};

} // namespace sp
#endif
