// Host side lexer compiler: lexem definitions -> flat tables (lexer_tables.h) ready for upload to HBM.
//
// Mirrors the definition phase of the reference's PatternLexerInstance:
//   PatternTable::definePattern   src/patternLexer.cpp:245-262  (" ~N" suffix :605-626, range checks :76-99)
//   PatternTable::defineSymbol    src/patternLexer.cpp:264-311
//   PatternTable::complete        src/patternLexer.cpp:333-412  (engine id = definition index)
//   PatternLexerInstance::compile src/patternLexer.cpp:1068-1118
#ifndef SP_LEXER_COMPILE_HPP_INCLUDED
#define SP_LEXER_COMPILE_HPP_INCLUDED
#include "automaton.hpp"
#include "lexer_tables.h"
#include <cstring>
#include <cstdlib>

namespace sp {

struct LexemDefinition
{
	std::string expression;
	uint32_t id;
	uint8_t level, posbind, resultidx, editdist, symtab;
};

// All table arrays as host vectors; `view` points into them (or into their device copies after upload).
struct LexerTablesHost
{
	std::vector<uint16_t> fwd_next;
	std::vector<uint8_t> fwd_class;
	std::vector<uint32_t> fwd_acc_off;
	std::vector<SpAccSpan> fwd_acc_span;
	std::vector<uint16_t> fwd_acc_pat;
	std::vector<uint32_t> rev_next;
	std::vector<uint8_t> rev_classes;
	std::vector<uint32_t> cap_next;
	std::vector<uint8_t> cap_classes;
	std::vector<SpAnchoredDfa> cap;
	std::vector<SpPatternInfo> pat;
	std::vector<SpSymbolSlot> sym_slots;
	std::vector<uint8_t> sym_pool;
	SpLexerTables view;

	void pointToHost()
	{
		view.fwd_next = fwd_next.data(); view.fwd_class = fwd_class.data();
		view.fwd_acc_off = fwd_acc_off.data(); view.fwd_acc_span = fwd_acc_span.data(); view.fwd_acc_pat = fwd_acc_pat.data();
		view.rev_next = rev_next.data(); view.rev_classes = rev_classes.data();
		view.cap_next = cap_next.data(); view.cap_classes = cap_classes.data(); view.cap = cap.data();
		view.pat = pat.data(); view.sym_slots = sym_slots.data(); view.sym_pool = sym_pool.data();
	}
};

class LexerProgram
{
public:
	enum {MaxPatternId = (1<<30)-1};

	LexerProgram() :m_flags(0),m_bytechar(false),m_compiled(false),m_fwdStates(0),m_revStates(0){}

	bool compiled() const {return m_compiled;}
	unsigned int flags() const {return m_flags;}
	const std::vector<LexemDefinition>& definitions() const {return m_defs;}
	const LexerTablesHost& tables() const {return m_tables;}
	LexerTablesHost& tables() {return m_tables;}
	uint32_t forwardStates() const {return m_fwdStates;}
	uint32_t reverseStates() const {return m_revStates;}

	// ---- compiled program <-> flat byte image (the reference recompiles its Hyperscan database on every
	// start, hs_serialize_database is unused; compiling the C2 set takes seconds here, loading the image milliseconds).
	// Image: header {magic, version, layout hash, payload bytes, payload FNV-1a} + payload (scalars, definitions, symbols,
	// every table array as {count, raw elements}). It is tied to this build's table layout (layout hash) and byte order.
	enum {ImageMagic = 0x584C5053u /*"SPLX"*/, ImageVersion = 1, NofImageScalars = 10};

	void serialize( std::vector<uint8_t>& out) const
	{
		if (!m_compiled) throw std::runtime_error( "serialize called before compile");
		ImageWriter w;
		w.pod( m_flags); w.pod( (uint32_t)m_bytechar); w.pod( m_fwdStates); w.pod( m_revStates);
		const SpLexerTables& V = m_tables.view;
		const uint32_t scalars[ NofImageScalars] = {V.fwd_states, V.fwd_stride, V.fwd_init, V.fwd_acc_base, V.fwd_cut_lo, V.fwd_cut_n, V.fwd_max_acc, V.npat, V.sym_mask, V.has_symbols};
		w.pod( scalars);
		w.pod( (uint64_t)m_defs.size());
		for (const LexemDefinition& d: m_defs)
		{
			w.str( d.expression); w.pod( d.id); w.pod( d.level); w.pod( d.posbind); w.pod( d.resultidx); w.pod( d.editdist); w.pod( d.symtab);
		}
		w.pod( (uint64_t)m_symtabOfLexem.size());
		for (const auto& kv: m_symtabOfLexem) {w.pod( kv.first); w.pod( kv.second);}
		w.pod( (uint64_t)m_symbols.size());
		for (const auto& kv: m_symbols) {w.pod( kv.first.first); w.str( kv.first.second); w.pod( kv.second);}
		const LexerTablesHost& T = m_tables;
		w.vec( T.fwd_next); w.vec( T.fwd_class); w.vec( T.fwd_acc_off); w.vec( T.fwd_acc_span); w.vec( T.fwd_acc_pat);
		w.vec( T.rev_next); w.vec( T.rev_classes); w.vec( T.cap_next); w.vec( T.cap_classes); w.vec( T.cap);
		w.vec( T.pat); w.vec( T.sym_slots); w.vec( T.sym_pool);
		const uint32_t header[ 4] = {(uint32_t)ImageMagic, (uint32_t)ImageVersion, layoutHash(), fnv1a( w.bytes.data(), w.bytes.size())};
		const uint64_t payload = w.bytes.size();
		out.clear();
		out.insert( out.end(), (const uint8_t*)header, (const uint8_t*)header + sizeof(header));
		out.insert( out.end(), (const uint8_t*)&payload, (const uint8_t*)&payload + sizeof(payload));
		out.insert( out.end(), w.bytes.begin(), w.bytes.end());
	}

	void deserialize( const uint8_t* image, std::size_t size)
	{
		if (m_compiled || !m_defs.empty()) throw std::runtime_error( "a compiled lexer image can only be loaded into an empty lexer");
		uint32_t header[ 4]; uint64_t payload;
		if (size < sizeof(header) + sizeof(payload)) throw std::runtime_error( "lexer image truncated");
		std::memcpy( header, image, sizeof(header)); std::memcpy( &payload, image + sizeof(header), sizeof(payload));
		if (header[0] != (uint32_t)ImageMagic) throw std::runtime_error( "not a lexer image");
		if (header[1] != (uint32_t)ImageVersion || header[2] != layoutHash()) throw std::runtime_error( "lexer image of another version or table layout: compile again");
		const uint8_t* body = image + sizeof(header) + sizeof(payload);
		if (payload != size - sizeof(header) - sizeof(payload)) throw std::runtime_error( "lexer image truncated");
		if (header[3] != fnv1a( body, (std::size_t)payload)) throw std::runtime_error( "lexer image damaged (checksum)");
		ImageReader r( body, (std::size_t)payload);
		LexerProgram fresh;
		uint32_t bytechar;
		r.pod( fresh.m_flags); r.pod( bytechar); r.pod( fresh.m_fwdStates); r.pod( fresh.m_revStates);
		fresh.m_bytechar = bytechar != 0;
		uint32_t scalars[ NofImageScalars];
		r.pod( scalars);
		{
			SpLexerTables& W = fresh.m_tables.view;
			std::memset( &W, 0, sizeof(W));
			W.fwd_states = scalars[0]; W.fwd_stride = scalars[1]; W.fwd_init = scalars[2]; W.fwd_acc_base = scalars[3]; W.fwd_cut_lo = scalars[4];
			W.fwd_cut_n = scalars[5]; W.fwd_max_acc = scalars[6]; W.npat = scalars[7]; W.sym_mask = scalars[8]; W.has_symbols = scalars[9];
		}
		uint64_t n;
		r.pod( n);
		for (uint64_t i=0; i<n; ++i)
		{
			LexemDefinition d;
			r.str( d.expression); r.pod( d.id); r.pod( d.level); r.pod( d.posbind); r.pod( d.resultidx); r.pod( d.editdist); r.pod( d.symtab);
			fresh.m_defs.push_back( d);
		}
		r.pod( n);
		for (uint64_t i=0; i<n; ++i) {uint32_t k; uint8_t v; r.pod( k); r.pod( v); fresh.m_symtabOfLexem[ k] = v;}
		r.pod( n);
		for (uint64_t i=0; i<n; ++i) {uint32_t lexemid, symid; std::string name; r.pod( lexemid); r.str( name); r.pod( symid); fresh.m_symbols[ std::make_pair( lexemid, name)] = symid;}
		LexerTablesHost& T = fresh.m_tables;
		r.vec( T.fwd_next); r.vec( T.fwd_class); r.vec( T.fwd_acc_off); r.vec( T.fwd_acc_span); r.vec( T.fwd_acc_pat);
		r.vec( T.rev_next); r.vec( T.rev_classes); r.vec( T.cap_next); r.vec( T.cap_classes); r.vec( T.cap);
		r.vec( T.pat); r.vec( T.sym_slots); r.vec( T.sym_pool);
		if (!r.atEnd()) throw std::runtime_error( "lexer image has trailing bytes");
		// the arrays must be as large as the scalars say: a kernel indexes them without further checks
		const SpLexerTables& V = T.view;
		if (T.fwd_next.size() != (std::size_t)V.fwd_states * V.fwd_stride || T.fwd_class.size() != 256 || V.fwd_acc_base > V.fwd_states
		||  T.fwd_acc_off.size() != (std::size_t)(V.fwd_states - V.fwd_acc_base) + 1 || T.fwd_acc_span.size() + 1 != T.fwd_acc_off.size()
		||  T.fwd_acc_off.back() != T.fwd_acc_pat.size() || T.pat.size() != V.npat || fresh.m_defs.size() != V.npat
		||  (V.sym_mask && T.sym_slots.size() != (std::size_t)V.sym_mask + 1))
		{
			throw std::runtime_error( "lexer image inconsistent");
		}
		// ... and every index stored inside them must stay inside: a crafted or stale image (the checksum guards against
		// neither) would otherwise make the kernels read out of bounds, which ends the process with a sticky CUDA error
		{
			bool ok = V.fwd_init < V.fwd_states && V.fwd_stride >= 2 && (uint64_t)V.fwd_cut_lo + V.fwd_cut_n <= V.fwd_states
				&& (V.sym_mask == 0 || ((V.sym_mask + 1) & V.sym_mask) == 0);
			for (std::size_t i=0; ok && i<T.fwd_next.size(); ++i) ok = T.fwd_next[ i] < V.fwd_states;
			for (std::size_t i=0; ok && i<256; ++i) ok = (uint32_t)T.fwd_class[ i] + 1 < V.fwd_stride;
			for (std::size_t i=0; ok && i<T.fwd_acc_pat.size(); ++i) ok = T.fwd_acc_pat[ i] < V.npat;
			for (std::size_t i=0; ok && i+1<T.fwd_acc_off.size(); ++i)
			{
				ok = T.fwd_acc_off[ i] <= T.fwd_acc_off[ i+1] && T.fwd_acc_span[ i].off == T.fwd_acc_off[ i]
					&& T.fwd_acc_span[ i].n == T.fwd_acc_off[ i+1] - T.fwd_acc_off[ i] && T.fwd_acc_span[ i].n <= V.fwd_max_acc;
			}
			for (std::size_t i=0; ok && i<T.rev_next.size(); ++i) ok = (T.rev_next[ i] & SP_REV_ROWMASK) < T.rev_next.size() || (T.rev_next[ i] & SP_REV_STOP);
			for (std::size_t i=0; ok && i<T.cap_next.size(); ++i) ok = (T.cap_next[ i] & SP_REV_ROWMASK) < T.cap_next.size() || (T.cap_next[ i] & SP_REV_STOP);
			for (std::size_t i=0; ok && i<T.pat.size(); ++i)
			{
				const SpPatternInfo& pi = T.pat[ i];
				ok = pi.rev_stride >= 1 && (uint64_t)pi.rev_class + 256 <= T.rev_classes.size() && pi.posbind <= 3
					&& (!pi.resultidx || (uint64_t)pi.cap_first + 5 <= T.cap.size());
				for (int k=0; ok && k<4; ++k) ok = (uint64_t)pi.rev_init[ k] + pi.rev_stride <= T.rev_next.size();
				for (std::size_t b=0; ok && b<256; ++b) ok = (uint32_t)T.rev_classes[ pi.rev_class + b] < pi.rev_stride;
			}
			for (std::size_t i=0; ok && i<T.cap.size(); ++i)
			{
				const SpAnchoredDfa& c = T.cap[ i];
				ok = c.stride >= 1 && (uint64_t)c.classes + 256 <= T.cap_classes.size();
				for (int k=0; ok && k<4; ++k) ok = (uint64_t)c.init[ k] + c.stride <= T.cap_next.size();
				for (std::size_t b=0; ok && b<256; ++b) ok = (uint32_t)T.cap_classes[ c.classes + b] < c.stride;
			}
			for (std::size_t i=0; ok && i<T.sym_slots.size(); ++i) ok = !T.sym_slots[ i].hash || (uint64_t)T.sym_slots[ i].str + T.sym_slots[ i].len <= T.sym_pool.size();
			if (!ok) throw std::runtime_error( "lexer image inconsistent: an index inside the tables is out of range");
		}
		fresh.m_compiled = true;
		*this = fresh;
		m_tables.pointToHost();
	}

	// PatternLexerInstance::defineOption, src/patternLexer.cpp:1031-1066
	void defineOption( const std::string& name)
	{
		if (m_compiled) throw std::runtime_error( "called define option after calling 'compile'");
		std::string n( name);
		for (char& c: n) c = (char)toupper( (unsigned char)c);
		if (n == "CASELESS") m_flags |= RX_CASELESS;
		else if (n == "DOTALL") m_flags |= RX_DOTALL;
		else if (n == "MULTILINE") m_flags |= RX_MULTILINE;
		else if (n == "ALLOWEMPTY") m_flags |= RX_ALLOWEMPTY;
		else if (n == "UCP") m_flags |= RX_UCP;
		else if (n == "BYTECHAR") m_bytechar = true;
		else throw std::runtime_error( "unknown option '" + name + "'");
	}

	void defineLexem( uint32_t id, const std::string& expression_, unsigned int resultIndex, unsigned int level, unsigned int posbind)
	{
		if (m_compiled) throw std::runtime_error( "called define pattern after calling 'compile'");
		LexemDefinition def;
		def.expression = expression_;
		unsigned int editdist = stripEditDistance( def.expression);
		if (id > MaxPatternId) throw std::runtime_error( "pattern id out of range");
		if (level > 255) throw std::runtime_error( "level out of range");
		if (resultIndex > 255) throw std::runtime_error( "result index out of range");
		if (editdist > 255) throw std::runtime_error( "edit distance out of range");
		if (posbind > 3) throw std::runtime_error( "unknown position bind value");
		def.id = id; def.level = (uint8_t)level; def.posbind = (uint8_t)posbind;
		def.resultidx = (uint8_t)resultIndex; def.editdist = (uint8_t)editdist; def.symtab = 0;
		m_defs.push_back( def);
	}

	void defineSymbol( uint32_t symbolid, uint32_t lexemid, const std::string& name)
	{
		if (m_compiled) throw std::runtime_error( "called define symbol after calling 'compile'");
		if (lexemid > MaxPatternId) throw std::runtime_error( "pattern id out of range");
		if (symbolid > MaxPatternId) throw std::runtime_error( "symbol id out of range");
		if (name.size() > 0xFFFF) throw std::runtime_error( "symbol too long");
		uint8_t table = tableOf( lexemid, true);
		std::pair<uint32_t,std::string> key( table, name);
		if (m_symbols.find( key) != m_symbols.end()) throw std::runtime_error( "symbol defined twice: '" + name + "'");
		m_symbols[ key] = symbolid;
	}

	uint32_t getSymbol( uint32_t lexemid, const std::string& name) const
	{
		std::map<uint32_t,uint8_t>::const_iterator ti = m_symtabOfLexem.find( lexemid);
		if (ti == m_symtabOfLexem.end()) return 0;
		std::map<std::pair<uint32_t,std::string>,uint32_t>::const_iterator si = m_symbols.find( std::make_pair( (uint32_t)ti->second, name));
		return si == m_symbols.end() ? 0 : si->second;
	}

	// Throws SyntaxError / UnsupportedError / std::runtime_error; the message names the offending expression
	// like the reference does (src/patternLexer.cpp:1087-1113).
	void compile()
	{
		if (m_bytechar) throw UnsupportedError( "option BYTECHAR (one byte character map) is not supported");
		ClassGraphCache cache;
		ByteNfa fwd;
		std::vector<int> entries;
		std::vector<RxRef> trees;
		m_tables = LexerTablesHost();
		m_tables.pat.resize( m_defs.size());
		for (std::size_t di=0; di<m_defs.size(); ++di)
		{
			LexemDefinition& def = m_defs[di];
			std::map<uint32_t,uint8_t>::const_iterator ti = m_symtabOfLexem.find( def.id);
			def.symtab = (ti == m_symtabOfLexem.end()) ? 0 : ti->second;
			try
			{
				if (def.editdist) throw UnsupportedError( "approximate matching (~N) is not supported");
				RegexParser parser( def.expression, m_flags);
				RxRef tree = parser.parse();
				if ((m_flags & RX_UCP) && containsWordAssertion( tree)) throw UnsupportedError( "\\b and \\B with option UCP are not supported");
				trees.push_back( tree);
				NfaBuilder builder( fwd, cache, false);
				int entry = builder.addPattern( tree, (int)di);
				if (!(m_flags & RX_ALLOWEMPTY) && nfaNullable( fwd, entry)) throw SyntaxError( "pattern matches empty buffer; use option ALLOWEMPTY");
				entries.push_back( entry);
				buildReverse( di, tree, cache);
				SpPatternInfo& P = m_tables.pat[di];
				P.id = def.id; P.level = def.level; P.posbind = def.posbind; P.symtab = def.symtab; P.resultidx = def.resultidx;
				P.cap_first = 0;
				if (def.resultidx) buildCapture( di, cache);
			}
			catch (const UnsupportedError& e)
			{
				throw UnsupportedError( "failed to compile pattern \"" + def.expression + "\": " + e.what());
			}
			catch (const std::runtime_error& e)
			{
				throw SyntaxError( "failed to compile pattern \"" + def.expression + "\": " + e.what());
			}
		}
		if (m_defs.size() > 65535) throw UnsupportedError( "too many lexem definitions");
		buildForward( fwd, entries);
		buildSymbols();
		m_tables.view.npat = (uint32_t)m_defs.size();
		m_tables.pointToHost();
		m_compiled = true;
	}

private:
	// restates extractEditDistFromExpression, src/patternLexer.cpp:605-626
	static unsigned int stripEditDistance( std::string& expr)
	{
		long se = (long)expr.size() - 1;
		while (se >= 0 && (unsigned char)expr[se] <= 32) --se;
		long de = se;
		while (se >= 0 && expr[se] >= '0' && expr[se] <= '9') --se;
		if (de == se) return 0;
		std::string digits = expr.substr( se+1, de-se);
		while (se >= 0 && (unsigned char)expr[se] <= 32) --se;
		if (se < 0 || expr[se] != '~') return 0;
		while (se > 0 && (unsigned char)expr[se-1] <= 32) --se;
		expr.resize( se);
		return (unsigned int)std::atoi( digits.c_str());
	}
	uint8_t tableOf( uint32_t lexemid, bool create)
	{
		std::map<uint32_t,uint8_t>::const_iterator ti = m_symtabOfLexem.find( lexemid);
		if (ti != m_symtabOfLexem.end()) return ti->second;
		if (!create) return 0;
		if (m_symtabOfLexem.size() >= 255) throw std::runtime_error( "too many pattern symbol tables defined");
		uint8_t rt = (uint8_t)(m_symtabOfLexem.size() + 1);
		m_symtabOfLexem[ lexemid] = rt;
		return rt;
	}

	// appends an anchored automaton to `next`/`classes` in the row|flags entry format; returns init rows and stride
	static void appendAnchored( const Dfa& dfa, std::vector<uint32_t>& next, std::vector<uint8_t>& classes,
				    uint32_t init[4], uint32_t& stride, uint32_t& classOffset)
	{
		stride = dfa.stride();
		classOffset = (uint32_t)classes.size();
		classes.insert( classes.end(), dfa.byteclass, dfa.byteclass + 256);
		std::size_t base = next.size();
		if (base + (std::size_t)dfa.nstates * stride >= SP_REV_ROWMASK) throw UnsupportedError( "automata too large");
		std::vector<char> dying( dfa.nstates, 0);
		for (uint32_t s=0; s<dfa.nstates; ++s)
		{
			bool allDead = true;
			for (uint32_t c=0; c<stride && allDead; ++c) allDead = dfa.next[ (std::size_t)s*stride + c] == dfa.dead;
			dying[s] = allDead;
		}
		next.resize( base + (std::size_t)dfa.nstates * stride);
		for (uint32_t s=0; s<dfa.nstates; ++s)
		{
			for (uint32_t c=0; c<stride; ++c)
			{
				uint32_t t = dfa.next[ (std::size_t)s*stride + c];
				uint32_t e = (uint32_t)(base + (std::size_t)t * stride);
				if (!dfa.accepts[t].empty()) e |= SP_REV_ACCEPT;
				if (dying[t]) e |= SP_REV_STOP;
				next[ base + (std::size_t)s*stride + c] = e;
			}
		}
		for (int k=0; k<4; ++k) init[k] = (uint32_t)(base + (std::size_t)dfa.init[k] * stride);
	}

	void buildReverse( std::size_t di, const RxRef& tree, ClassGraphCache& cache)
	{
		ByteNfa nfa;
		NfaBuilder builder( nfa, cache, true);
		std::vector<int> entry( 1, builder.addPattern( tree, 0));
		DfaOptions opt; opt.unanchored = false; opt.allowEmpty = false;
		Dfa dfa = DfaBuilder( nfa, entry, opt).build();
		SpPatternInfo& P = m_tables.pat[di];
		appendAnchored( dfa, m_tables.rev_next, m_tables.rev_classes, P.rev_init, P.rev_stride, P.rev_class);
		m_revStates += dfa.nstates;
	}

	// The selected sub expression must be a top level group: expression = A (G) B. The helper automata are
	// compiled without the instance options, as the reference hands the bare expression to libtre
	// (REG_EXTENDED only, src/patternLexer.cpp:474); '.' includes newline there.
	void buildCapture( std::size_t di, ClassGraphCache& cache)
	{
		const LexemDefinition& def = m_defs[di];
		RegexParser parser( def.expression, RX_DOTALL);
		RxRef tree = parser.parse();
		if ((int)def.resultidx > parser.captureCount()) throw SyntaxError( "sub expression selection index out of range");
		std::vector<RxRef> parts;
		if (tree->kind == RxConcat) parts = tree->sub; else parts.push_back( tree);
		std::size_t gi = 0;
		for (; gi<parts.size(); ++gi) if (parts[gi]->kind == RxGroup && parts[gi]->capture == (int)def.resultidx) break;
		if (gi == parts.size()) throw UnsupportedError( "the selected sub expression must be a top level group of the expression");
		RxRef pieces[5] = {tree, Rx::concatOf( parts, 0, gi), Rx::concatOf( parts, gi, parts.size()),
				   Rx::concatOf( parts, gi, gi+1), Rx::concatOf( parts, gi+1, parts.size())};
		m_tables.pat[di].cap_first = (uint32_t)m_tables.cap.size();
		for (int k=0; k<5; ++k)
		{
			ByteNfa nfa;
			NfaBuilder builder( nfa, cache, false);
			std::vector<int> entry( 1, builder.addPattern( pieces[k], 0));
			DfaOptions opt; opt.unanchored = false; opt.allowEmpty = true;
			Dfa dfa = DfaBuilder( nfa, entry, opt).build();
			SpAnchoredDfa A; A.accept_on_entry = 0;
			appendAnchored( dfa, m_tables.cap_next, m_tables.cap_classes, A.init, A.stride, A.classes);
			m_tables.cap.push_back( A);
		}
	}

	void buildForward( const ByteNfa& nfa, const std::vector<int>& entries)
	{
		// States reached by a transition that no older thread survived are told apart (Dfa::cut): behind such a byte no
		// pending match reaches back in front of it, which is where the fused kernel closes a segment of events. If
		// that split makes the automaton too large, it is built without (the lexer then runs the island pipeline only).
		DfaOptions opt; opt.unanchored = true; opt.allowEmpty = false; opt.maxStates = 65000; opt.markCuts = true;
		Dfa dfa;
		try {dfa = DfaBuilder( nfa, entries, opt).build();}
		catch (const DfaTooLarge&)
		{
			opt.markCuts = false;
			try {dfa = DfaBuilder( nfa, entries, opt).build();}
			catch (const DfaTooLarge&)
			{
				// all definitions share one forward automaton with 16 bit state numbers (limits)
				throw UnsupportedError( "the " + std::to_string( m_defs.size()) + " lexem definitions together need more than "
							+ std::to_string( opt.maxStates) + " automaton states (counted repeats or '.*' in front of long tails multiply states); "
							"split the set over several lexer instances");
			}
		}
		// renumber: [plain][plain, cut][reporting, cut][reporting] so that the scan loop tests `state >= fwd_acc_base`
		// for matches and `state - fwd_cut_lo < fwd_cut_n` for cuts
		std::vector<uint32_t> newid( dfa.nstates);
		uint32_t nextid = 0, nplain = 0, cutlo = 0;
		for (int group=0; group<4; ++group)
		{
			if (group == 1) cutlo = nextid;
			if (group == 2) nplain = nextid;
			for (uint32_t s=0; s<dfa.nstates; ++s)
			{
				bool acc = !dfa.accepts[s].empty(), cut = dfa.cut[s] != 0;
				int g = acc ? (cut ? 2 : 3) : (cut ? 1 : 0);
				if (g == group) newid[s] = nextid++;
			}
			if (group == 2) {m_tables.view.fwd_cut_lo = cutlo; m_tables.view.fwd_cut_n = nextid - cutlo;}
		}
		uint32_t stride = dfa.stride();
		LexerTablesHost& T = m_tables;
		T.fwd_next.resize( (std::size_t)dfa.nstates * stride);
		T.fwd_acc_off.assign( dfa.nstates - nplain + 1, 0);
		for (uint32_t s=0; s<dfa.nstates; ++s)
		{
			for (uint32_t c=0; c<stride; ++c) T.fwd_next[ (std::size_t)newid[s]*stride + c] = (uint16_t)newid[ dfa.next[ (std::size_t)s*stride + c]];
		}
		std::vector<std::vector<uint32_t> > acc( dfa.nstates - nplain);
		for (uint32_t s=0; s<dfa.nstates; ++s) if (!dfa.accepts[s].empty()) acc[ newid[s]-nplain] = dfa.accepts[s];
		for (std::size_t a=0; a<acc.size(); ++a)
		{
			T.fwd_acc_off[a] = (uint32_t)T.fwd_acc_pat.size();
			for (uint32_t p: acc[a]) T.fwd_acc_pat.push_back( (uint16_t)p);	// ascending definition index = canonical event order
		}
		T.fwd_acc_off[ acc.size()] = (uint32_t)T.fwd_acc_pat.size();
		T.fwd_acc_span.resize( acc.size());
		for (std::size_t a=0; a<acc.size(); ++a) {T.fwd_acc_span[a].off = T.fwd_acc_off[a]; T.fwd_acc_span[a].n = T.fwd_acc_off[a+1] - T.fwd_acc_off[a];}
		T.view.fwd_max_acc = 0;
		for (std::size_t a=0; a<acc.size(); ++a) if (acc[a].size() > T.view.fwd_max_acc) T.view.fwd_max_acc = (uint32_t)acc[a].size();
		T.fwd_class.assign( dfa.byteclass, dfa.byteclass + 256);
		T.view.fwd_states = dfa.nstates;
		T.view.fwd_stride = stride;
		T.view.fwd_init = newid[ dfa.init[ KindEdge]];
		T.view.fwd_acc_base = nplain;
		m_fwdStates = dfa.nstates;
	}

	void buildSymbols()
	{
		LexerTablesHost& T = m_tables;
		T.view.has_symbols = m_symbols.empty() ? 0 : 1;
		T.view.sym_mask = 0;
		if (m_symbols.empty()) {T.sym_slots.assign( 1, SpSymbolSlot()); std::memset( &T.sym_slots[0], 0, sizeof(SpSymbolSlot)); T.sym_pool.assign( 1, 0); return;}
		std::size_t nslots = 16;
		while (nslots < m_symbols.size() * 2) nslots <<= 1;
		SpSymbolSlot empty; std::memset( &empty, 0, sizeof(empty));
		T.sym_slots.assign( nslots, empty);
		T.view.sym_mask = (uint32_t)nslots - 1;
		for (const auto& kv: m_symbols)
		{
			const std::string& name = kv.first.second;
			uint32_t h = symbolHash( (const uint8_t*)name.data(), (uint32_t)name.size(), kv.first.first);
			std::size_t slot = h & T.view.sym_mask;
			while (T.sym_slots[ slot].hash) slot = (slot + 1) & T.view.sym_mask;
			SpSymbolSlot& e = T.sym_slots[ slot];
			e.hash = h; e.symid = kv.second; e.str = (uint32_t)T.sym_pool.size(); e.len = (uint16_t)name.size(); e.symtab = (uint16_t)kv.first.first;
			T.sym_pool.insert( T.sym_pool.end(), name.begin(), name.end());
		}
		if (T.sym_pool.empty()) T.sym_pool.push_back( 0);
	}
	struct ImageWriter
	{
		std::vector<uint8_t> bytes;
		template <typename T> void pod( const T& v) {const uint8_t* p = (const uint8_t*)&v; bytes.insert( bytes.end(), p, p + sizeof(T));}
		void str( const std::string& v) {pod( (uint64_t)v.size()); bytes.insert( bytes.end(), v.begin(), v.end());}
		template <typename T> void vec( const std::vector<T>& v)
		{
			pod( (uint64_t)v.size());
			const uint8_t* p = (const uint8_t*)v.data();
			bytes.insert( bytes.end(), p, p + v.size() * sizeof(T));
		}
	};
	struct ImageReader
	{
		const uint8_t* p; std::size_t left;
		ImageReader( const uint8_t* p_, std::size_t n) :p(p_),left(n){}
		void need( std::size_t n) {if (n > left) throw std::runtime_error( "lexer image truncated");}
		template <typename T> void pod( T& v) {need( sizeof(T)); std::memcpy( &v, p, sizeof(T)); p += sizeof(T); left -= sizeof(T);}
		void str( std::string& v) {uint64_t n; pod( n); need( (std::size_t)n); v.assign( (const char*)p, (std::size_t)n); p += n; left -= (std::size_t)n;}
		template <typename T> void vec( std::vector<T>& v)
		{
			uint64_t n; pod( n);
			if (n > left / sizeof(T)) throw std::runtime_error( "lexer image truncated");
			v.resize( (std::size_t)n);
			if (n) std::memcpy( (void*)v.data(), p, (std::size_t)n * sizeof(T));
			p += n * sizeof(T); left -= (std::size_t)n * sizeof(T);
		}
		bool atEnd() const {return left == 0;}
	};
	static uint32_t fnv1a( const uint8_t* s, std::size_t n)
	{
		uint32_t h = 2166136261u;
		for (std::size_t i=0; i<n; ++i) {h ^= s[i]; h *= 16777619u;}
		return h;
	}
	static uint32_t layoutHash()
	{
		const uint32_t sizes[] = {(uint32_t)sizeof(SpLexerTables), (uint32_t)sizeof(SpPatternInfo), (uint32_t)sizeof(SpAnchoredDfa),
					(uint32_t)sizeof(SpSymbolSlot), (uint32_t)sizeof(SpAccSpan), SP_REV_ACCEPT, SP_REV_STOP};
		return fnv1a( (const uint8_t*)sizes, sizeof(sizes));
	}
	static uint32_t symbolHash( const uint8_t* s, uint32_t len, uint32_t symtab)
	{
		// must equal sp_symbol_hash (lexer_device.cuh)
		uint32_t h = 2166136261u ^ (symtab * 0x9E3779B1u);
		for (uint32_t i=0; i<len; ++i) {h ^= s[i]; h *= 16777619u;}
		return h ? h : 1;
	}

	std::vector<LexemDefinition> m_defs;
	std::map<uint32_t,uint8_t> m_symtabOfLexem;
	std::map<std::pair<uint32_t,std::string>,uint32_t> m_symbols;
	unsigned int m_flags;
	bool m_bytechar;
	bool m_compiled;
	LexerTablesHost m_tables;
	uint32_t m_fwdStates, m_revStates;
};

} // namespace sp
#endif
