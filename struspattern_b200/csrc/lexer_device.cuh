// Per-thread building blocks of the lexer kernels (lexer_kernels.cu). Written as host+device inline
// functions so that tests/hostsim can run the very same table walks on the CPU against the oracle while no
// GPU is around; the product (libstrus_pattern_b200.so) only ever calls them from CUDA kernels.
//
// Reference behaviour restated here (file:line under the strusPattern tree):
//   scan contract of hs_scan with HS_FLAG_SOM_LEFTMOST          src/patternLexer.cpp:875-879, 403
//   match_event_handler (supersede logic, symbols, size check)  src/patternLexer.cpp:717-826
//   sub expression selection                                    src/patternLexer.cpp:414-426, 488-507
//   ordinal position assignment                                 src/patternLexer.cpp:893-945
#ifndef SP_LEXER_DEVICE_CUH_INCLUDED
#define SP_LEXER_DEVICE_CUH_INCLUDED
#include "lexer_tables.h"
#include "strus_pattern_b200.h"

#ifdef __CUDACC__
#define SP_HD __device__ __forceinline__
#define SP_LDG(p) __ldg(p)
#else
#define SP_HD inline
#define SP_LDG(p) (*(p))
#endif

#define SP_PAT_DROPPED   0xFFFFFFFFu	/* event removed by the sub expression selection */
#define SP_PAT_SIZEERROR 0xFFFFFFFEu	/* matched term of 65535 bytes or more: error of the document */

typedef struct alignas(16) SpRawEvent
{
	uint32_t pat;		/* 0-based definition index (reference: 1-based hyperscan id, src/patternLexer.cpp:390) */
	uint32_t from;		/* byte offsets relative to the document */
	uint32_t to;
	uint32_t aux;		/* document index until the start is resolved, then symbol id (0 = none) */
} SpRawEvent;

typedef struct alignas(16) SpMatchEvent	/* = MatchEvent, src/patternLexer.cpp:665-679 */
{
	uint32_t id;
	uint32_t origpos;
	uint32_t origsize;
	uint32_t meta;		/* level | posbind << 8 */
} SpMatchEvent;

// 16 byte records move as one vector access on the device (the arrays are 16 byte aligned)
SP_HD SpRawEvent sp_load_event( const SpRawEvent* p)
{
#ifdef __CUDACC__
	const uint4 v = *reinterpret_cast<const uint4*>( p);
	SpRawEvent e; e.pat = v.x; e.from = v.y; e.to = v.z; e.aux = v.w;
	return e;
#else
	return *p;
#endif
}
SP_HD void sp_store_event( SpRawEvent* p, const SpRawEvent& e)
{
#ifdef __CUDACC__
	*reinterpret_cast<uint4*>( p) = make_uint4( e.pat, e.from, e.to, e.aux);
#else
	*p = e;
#endif
}

SP_HD uint32_t sp_kind_of_byte( uint32_t b)
{
	if (b == '\n') return 2;
	return ((b - '0') < 10u || ((b | 0x20) - 'a') < 26u || b == '_') ? 1 : 0;
}

// index of the non empty document that contains the byte at global offset `pos` (pos < total)
SP_HD uint32_t sp_find_doc( const uint64_t* doc_off, uint32_t ndocs, uint64_t pos)
{
	uint32_t lo = 0, hi = ndocs;	// invariant: doc_off[lo] <= pos < doc_off[hi]
	while (hi - lo > 1)
	{
		uint32_t mid = lo + ((hi - lo) >> 1);
		if (SP_LDG( doc_off + mid) <= pos) lo = mid; else hi = mid;
	}
	return lo;
}

SP_HD SpAccSpan sp_load_span( const SpAccSpan* p)
{
#ifdef __CUDACC__
	const uint2 v = __ldg( reinterpret_cast<const uint2*>( p));
	SpAccSpan rt; rt.off = v.x; rt.n = v.y;
	return rt;
#else
	return *p;
#endif
}

struct SpCountSink
{
	const SpLexerTables* T;
	uint32_t count;
	SP_HD void accept( uint32_t state, uint32_t, uint32_t)
	{
		count += SP_LDG( &T->fwd_acc_span[ state - T->fwd_acc_base].n);
	}
	SP_HD void docBegin( uint32_t) {}
	SP_HD void cut() {}
};

#define SP_HEAD_NOCUT 0xFFFFFFFFu	/* chunk without a cut: all its events belong to a segment that began in front of it */

struct SpEmitSink
{
	const SpLexerTables* T;
	SpRawEvent* events;
	uint64_t cursor;
	uint64_t* doc_ev_begin;
	uint32_t pendingCut, seenCut, head;	/* a cut lies between the next event and the one in front of it; events in front of the first cut */
	SP_HD void accept( uint32_t state, uint32_t to, uint32_t doc)
	{
		const SpAccSpan span = sp_load_span( T->fwd_acc_span + (state - T->fwd_acc_base));
		if (!seenCut) head += span.n;
		for (uint32_t pi = span.off, pe = span.off + span.n; pi < pe; ++pi)
		{
			// until the start is resolved `from` says whether the event is the first one behind a cut
			SpRawEvent ev; ev.pat = SP_LDG( T->fwd_acc_pat + pi); ev.from = pendingCut; ev.to = to; ev.aux = doc;
			pendingCut = 0;
			sp_store_event( events + cursor++, ev);
		}
	}
	SP_HD void docBegin( uint32_t doc) {doc_ev_begin[ doc] = cursor;}
	SP_HD void cut() {pendingCut = 1; seenCut = 1;}
};

// Forward scan of the global byte range [begin,end) of the concatenated documents starting in `state`
// (ignored when `begin` is the first byte of a document). Returns the state behind the last byte.
template <class Sink>
SP_HD uint32_t sp_scan_range( const SpLexerTables& T, const uint8_t* text, const uint64_t* doc_off, uint32_t ndocs,
				uint64_t begin, uint64_t end, uint32_t state, Sink& sink)
{
	const uint32_t stride = T.fwd_stride, accbase = T.fwd_acc_base;
	uint32_t d = sp_find_doc( doc_off, ndocs, begin);
	uint64_t dstart = SP_LDG( doc_off + d), dend = SP_LDG( doc_off + d + 1);
	if (dstart == begin) state = T.fwd_init;
	uint64_t pos = begin;
	while (pos < end)
	{
		uint64_t stop = dend < end ? dend : end;
		while (pos < stop)
		{
			if ((((uintptr_t)(text + pos)) & 15) == 0 && pos + 16 <= stop)
			{
				uint4 v = *reinterpret_cast<const uint4*>( text + pos);
				uint32_t w[4] = {v.x, v.y, v.z, v.w};
				#pragma unroll
				for (int wi=0; wi<4; ++wi)
				{
					#pragma unroll
					for (int bi=0; bi<4; ++bi)
					{
						uint32_t b = (w[wi] >> (8*bi)) & 0xFF;
						state = SP_LDG( T.fwd_next + state * stride + T.fwd_class[ b]);
						if (state >= accbase) sink.accept( state, (uint32_t)(pos - dstart), d);
						if ((state - T.fwd_cut_lo) < T.fwd_cut_n) sink.cut();
						++pos;
					}
				}
			}
			else
			{
				uint32_t b = text[ pos];	// generic load: `text` may be a shared memory tile addressed like the text (k_lex_count<true>)
				state = SP_LDG( T.fwd_next + state * stride + T.fwd_class[ b]);
				if (state >= accbase) sink.accept( state, (uint32_t)(pos - dstart), d);
						if ((state - T.fwd_cut_lo) < T.fwd_cut_n) sink.cut();
				++pos;
			}
		}
		if (pos == dend)
		{
			// end of the document: the virtual end-of-text symbol reports the matches that end here
			state = SP_LDG( T.fwd_next + state * stride + (stride-1));
			if (state >= accbase) sink.accept( state, (uint32_t)(dend - dstart), d);
			sink.cut();	// a document border closes every segment
			state = T.fwd_init;
			for (++d; d < ndocs; ++d)
			{
				sink.docBegin( d);
				if (SP_LDG( doc_off + d + 1) > pos) break;	// next non empty document
			}
			if (d >= ndocs) break;
			dstart = pos; dend = SP_LDG( doc_off + d + 1);
		}
	}
	return state;
}

// Leftmost start of a match of definition `pat` that ends at `to` (document relative); doc = text of the
// document, doclen its size. SP_NO_POS if the reverse automaton finds none (table inconsistency).
SP_HD uint32_t sp_leftmost_start( const SpLexerTables& T, const SpPatternInfo& P, const uint8_t* doc, uint32_t doclen, uint32_t to)
{
	uint32_t pk = (to < doclen) ? sp_kind_of_byte( SP_LDG( doc + to)) : 3;
	uint32_t row = P.rev_init[ pk];
	const uint8_t* cls = T.rev_classes + P.rev_class;
	uint32_t from = SP_NO_POS;
	for (uint32_t i=to; i>0; --i)
	{
		uint32_t e = SP_LDG( T.rev_next + row + SP_LDG( cls + SP_LDG( doc + i - 1)));
		if (e & SP_REV_ACCEPT) from = i;
		if (e & SP_REV_STOP) return from;
		row = e & SP_REV_ROWMASK;
	}
	uint32_t e = SP_LDG( T.rev_next + row + (P.rev_stride-1));
	if (e & SP_REV_ACCEPT) from = 0;
	return from;
}

// The same with a memo of the last walk of the definition (exact: the reverse automaton is deterministic). A run of
// matches of one definition that end at consecutive offsets ("h", "he", "hel", ...) would walk back over the whole
// match for every end offset; but as soon as the walk from `to` is, behind the byte in front of the memorized end offset,
// in the row the memorized walk was in at that place, the rest of the two walks is the same and its outcome is known.
typedef struct SpStartMemo
{
	uint32_t pat;		/* definition (SP_NO_POS = empty) */
	uint32_t to;		/* end offset the memorized walk started from (same document) */
	uint32_t row1;		/* its row behind the first byte (SP_NO_POS: it stopped there) */
	uint32_t fromAfter;	/* its result counting only starts reported behind the first byte (SP_NO_POS: none) */
} SpStartMemo;

SP_HD uint32_t sp_leftmost_start_memo( const SpLexerTables& T, uint32_t pat, const SpPatternInfo& P, const uint8_t* doc, uint32_t doclen, uint32_t to, SpStartMemo& m)
{
	const uint32_t pk = (to < doclen) ? sp_kind_of_byte( SP_LDG( doc + to)) : 3;
	uint32_t row = P.rev_init[ pk];
	const uint8_t* cls = T.rev_classes + P.rev_class;
	const bool usable = m.pat == pat && m.to < to;
	const uint32_t mto = m.to, mrow1 = m.row1, mfrom = m.fromAfter;
	uint32_t from = SP_NO_POS, fromAfter = SP_NO_POS, row1 = SP_NO_POS;
	uint32_t i = to;
	for (; i>0; --i)
	{
		const uint32_t e = SP_LDG( T.rev_next + row + SP_LDG( cls + SP_LDG( doc + i - 1)));
		if (e & SP_REV_ACCEPT) {from = i; if (i != to) fromAfter = i;}
		if (e & SP_REV_STOP) break;
		row = e & SP_REV_ROWMASK;
		if (i == to) row1 = row;
		if (usable && i == mto && row == mrow1)
		{
			if (mfrom != SP_NO_POS) from = fromAfter = mfrom;
			break;
		}
	}
	if (i == 0)
	{
		const uint32_t e = SP_LDG( T.rev_next + row + (P.rev_stride-1));
		if (e & SP_REV_ACCEPT) {from = 0; if (to) fromAfter = 0;}
	}
	m.pat = pat; m.to = to; m.row1 = row1; m.fromAfter = fromAfter;
	return from;
}

// ---- capture selection (libtre re-match of the reference restated with anchored automata)
// true if the automaton matches exactly doc[a,b) with the real neighbours visible to assertions
SP_HD bool sp_cap_span( const SpLexerTables& T, const SpAnchoredDfa& A, const uint8_t* doc, uint32_t doclen, uint32_t a, uint32_t b)
{
	uint32_t pk = a ? sp_kind_of_byte( SP_LDG( doc + a - 1)) : 3;
	uint32_t row = A.init[ pk];
	const uint8_t* cls = T.cap_classes + A.classes;
	for (uint32_t i=a; i<b; ++i)
	{
		uint32_t e = SP_LDG( T.cap_next + row + SP_LDG( cls + SP_LDG( doc + i)));
		if (e & SP_REV_STOP) return false;
		row = e & SP_REV_ROWMASK;
	}
	uint32_t e = (b < doclen) ? SP_LDG( T.cap_next + row + SP_LDG( cls + SP_LDG( doc + b))) : SP_LDG( T.cap_next + row + (A.stride-1));
	return (e & SP_REV_ACCEPT) != 0;
}

// PatternTable::matchSubExpression (src/patternLexer.cpp:414-426) + SubExpressionDef::match (:488-507):
// POSIX leftmost-longest overall match from `from`, then the selected group as early and as long as possible.
SP_HD bool sp_select_subexpression( const SpLexerTables& T, const SpPatternInfo& P, const uint8_t* doc, uint32_t doclen, uint32_t& from, uint32_t& to)
{
	const SpAnchoredDfa* C = T.cap + P.cap_first;	// whole, A, GB, G, B
	uint32_t best = SP_NO_POS;
	{
		uint32_t pk = from ? sp_kind_of_byte( SP_LDG( doc + from - 1)) : 3;
		uint32_t row = C[0].init[ pk];
		const uint8_t* cls = T.cap_classes + C[0].classes;
		uint32_t i = from;
		for (;; ++i)
		{
			if (i == doclen)
			{
				if (SP_LDG( T.cap_next + row + (C[0].stride-1)) & SP_REV_ACCEPT) best = i;
				break;
			}
			uint32_t b = SP_LDG( doc + i);
			uint32_t e = SP_LDG( T.cap_next + row + SP_LDG( cls + b));
			if (e & SP_REV_ACCEPT) best = i;
			if (b == 0 || (e & SP_REV_STOP)) break;		// the subject of the re-match is a C string
			row = e & SP_REV_ROWMASK;
		}
	}
	if (best == SP_NO_POS) return false;
	uint32_t gs = SP_NO_POS, ge = SP_NO_POS;
	for (uint32_t s=from; s<=best; ++s)
	{
		if (sp_cap_span( T, C[1], doc, doclen, from, s) && sp_cap_span( T, C[2], doc, doclen, s, best)) {gs = s; break;}
	}
	if (gs == SP_NO_POS) return false;
	for (uint32_t e=best+1; e>gs; --e)
	{
		if (sp_cap_span( T, C[3], doc, doclen, gs, e-1) && sp_cap_span( T, C[4], doc, doclen, e-1, best)) {ge = e-1; break;}
	}
	if (ge == SP_NO_POS) return false;
	if (ge > to) return false;	// mt.rm_eo > to - from (:503)
	from = gs; to = ge;
	return true;
}

SP_HD uint32_t sp_symbol_hash( const uint8_t* s, uint32_t len, uint32_t symtab)
{
	uint32_t h = 2166136261u ^ (symtab * 0x9E3779B1u);
	for (uint32_t i=0; i<len; ++i) {h ^= SP_LDG( s + i); h *= 16777619u;}
	return h ? h : 1;
}

// SymbolTable lookup of the matched bytes (src/patternLexer.cpp:312-317, 740-744); 0 = no symbol
SP_HD uint32_t sp_symbol_lookup( const SpLexerTables& T, uint32_t symtab, const uint8_t* s, uint32_t len)
{
	if (!T.sym_mask || len > 0xFFFF) return 0;
	uint32_t h = sp_symbol_hash( s, len, symtab);
	for (uint32_t slot = h & T.sym_mask;; slot = (slot + 1) & T.sym_mask)
	{
		SpSymbolSlot e = T.sym_slots[ slot];
		if (e.hash == 0) return 0;
		if (e.hash == h && e.len == len && e.symtab == symtab)
		{
			uint32_t i = 0;
			while (i < len && SP_LDG( T.sym_pool + e.str + i) == SP_LDG( s + i)) ++i;
			if (i == len) return e.symid;
		}
	}
}

// Completes one raw event whose leftmost start is known: size check, sub expression selection, symbol.
SP_HD void sp_finish_event_from( const SpLexerTables& T, const SpPatternInfo& P, SpRawEvent& ev, uint32_t from, const uint8_t* doc, uint32_t doclen)
{
	uint32_t to = ev.to;
	ev.aux = 0;
	if (from == SP_NO_POS || to - from >= 65535) {ev.pat = SP_PAT_SIZEERROR; ev.from = 0; return;}	// src/patternLexer.cpp:727-730
	if (P.resultidx && !sp_select_subexpression( T, P, doc, doclen, from, to)) {ev.pat = SP_PAT_DROPPED; ev.from = from; return;}
	ev.from = from; ev.to = to;
	if (P.symtab) ev.aux = sp_symbol_lookup( T, P.symtab, doc + from, to - from);
}

// Completes one raw event: leftmost start, size check, sub expression selection, symbol.
SP_HD void sp_finish_event( const SpLexerTables& T, SpRawEvent& ev, const uint8_t* doc, uint32_t doclen)
{
	const SpPatternInfo P = T.pat[ ev.pat];
	sp_finish_event_from( T, P, ev, sp_leftmost_start( T, P, doc, doclen, ev.to), doc, doclen);
}

SP_HD SpMatchEvent sp_make_match_event( uint32_t id, const SpPatternInfo& P, uint32_t from, uint32_t to)
{
	SpMatchEvent m; m.id = id; m.origpos = from; m.origsize = to - from; m.meta = (uint32_t)P.level | ((uint32_t)P.posbind << 8);
	return m;
}

// match_event_handler (src/patternLexer.cpp:746-822) for one completed event on the array ar[0..n).
template <class EventArray>	/* SpMatchEvent* or an accessor with operator[] (lexer_fused.cuh) */
SP_HD void sp_supersede( const SpLexerTables& T, const SpRawEvent& rev, EventArray ar, uint32_t& n)
{
	const SpPatternInfo P = T.pat[ rev.pat];
	SpMatchEvent ev = sp_make_match_event( P.id, P, rev.from, rev.to);
	bool withSymbol = rev.aux != 0 && rev.aux != P.id;
	SpMatchEvent symev = ev; symev.id = rev.aux;
	if (n == 0)
	{
		ar[ n++] = ev;
		if (withSymbol) ar[ n++] = symev;
		return;
	}
	const uint32_t evLevel = ev.meta & 0xFF, evEnd = ev.origpos + ev.origsize;
	if (!withSymbol)
	{
		// the usual case decided with one read: the new element starts behind the last one, so nothing is deleted (the first
		// loop below stops at once) and it is appended at the end (so does the third); the second loop looks at the last one only
		const SpMatchEvent last = ar[ n-1];
		if (last.origpos < ev.origpos)
		{
			if (last.origpos + last.origsize < evEnd) {ar[ n++] = ev; return;}
			if ((last.meta & 0xFF) > evLevel) return;
		}
	}
	// elements replaced by the new one (same lexem at the same place) or covered by it with a lower level
	uint32_t deleted = 0;
	for (uint32_t k=n; k>0; --k)
	{
		SpMatchEvent old = ar[ k-1];
		if (old.origpos < ev.origpos) break;
		uint32_t oldLevel = old.meta & 0xFF;
		if ((old.id == ev.id && old.origpos == ev.origpos && oldLevel == evLevel)
		||  (oldLevel < evLevel && old.origpos + old.origsize <= evEnd))
		{
			for (uint32_t m=k; m<n; ++m) {const SpMatchEvent moved = ar[ m]; ar[ m-1] = moved;}
			--n; ++deleted;
		}
	}
	if (!deleted)
	{
		// the new element is dropped if an element of higher level at the tail covers it
		for (uint32_t k=n; k>0; --k)
		{
			SpMatchEvent old = ar[ k-1];
			if (old.origpos + old.origsize < evEnd) break;
			if ((old.meta & 0xFF) > evLevel && old.origpos <= ev.origpos) return;
		}
	}
	if (!withSymbol)
	{
		uint32_t k = n;
		for (; k>0; --k)
		{
			const SpMatchEvent later = ar[ k-1];
			if (later.origpos <= ev.origpos) break;
			ar[ k] = later;
		}
		ar[ k] = ev;
		++n;
	}
	else
	{
		// two elements are inserted; as in the reference (:809-820) the first later-starting element moves
		// by two places and every further one by one only, which loses an element and leaves a zeroed one
		SpMatchEvent zero; zero.id = 0; zero.origpos = 0; zero.origsize = 0; zero.meta = 0;
		ar[ n] = zero; ar[ n+1] = zero;
		uint32_t dst = n + 1, k = n;
		for (; k>0; --k)
		{
			const SpMatchEvent later = ar[ k-1];
			if (later.origpos <= ev.origpos) break;
			ar[ dst] = later; dst = k-1;
		}
		ar[ k] = ev;
		ar[ k+1] = symev;
		n += 2;
	}
}

// Ordinal position assignment (src/patternLexer.cpp:893-945). out == 0: only count.
SP_HD uint32_t sp_assign_ordpos( const SpMatchEvent* ar, uint32_t n, sp_lexem* out)
{
	uint32_t nout = 0, mi = 0, ordpos = 0, origpos = 0, lastbind = SP_BIND_CONTENT;
	for (; mi < n; ++mi)
	{
		SpMatchEvent e = ar[ mi];
		uint32_t bind = (e.meta >> 8) & 0xFF;
		lastbind = bind;
		if (bind == SP_BIND_UNIQUE || bind == SP_BIND_CONTENT)
		{
			ordpos = 1; origpos = e.origpos;
			if (out) {sp_lexem lx; lx.id = e.id; lx.ordpos = 1; lx.origpos = e.origpos; lx.origsize = e.origsize; out[ nout] = lx;}
			++nout; ++mi;
			break;
		}
		else if (bind == SP_BIND_SUCCESSOR)
		{
			if (out) {sp_lexem lx; lx.id = e.id; lx.ordpos = 1; lx.origpos = e.origpos; lx.origsize = e.origsize; out[ nout] = lx;}
			++nout;
		}
	}
	if (ordpos == 0) return 0;	// no content element: empty result (:919-922)
	for (; mi < n; ++mi)
	{
		SpMatchEvent e = ar[ mi];
		uint32_t bind = (e.meta >> 8) & 0xFF;
		uint32_t pos = 0; bool emit = true;
		if (bind == SP_BIND_UNIQUE && lastbind == SP_BIND_UNIQUE) emit = false;
		else if (bind == SP_BIND_UNIQUE || bind == SP_BIND_CONTENT)
		{
			if (e.origpos > origpos) {origpos = e.origpos; ++ordpos;}
			pos = ordpos;
		}
		else if (bind == SP_BIND_SUCCESSOR) pos = ordpos + 1;
		else pos = ordpos;
		if (emit)
		{
			if (out) {sp_lexem lx; lx.id = e.id; lx.ordpos = pos; lx.origpos = e.origpos; lx.origsize = e.origsize; out[ nout] = lx;}
			++nout;
		}
		lastbind = bind;
	}
	return nout;
}

// ------------------------------------------------------------------------------------------------
// Kernel bodies: one call = the work of one CUDA thread (lexer_kernels.cu) or one loop iteration (tests/hostsim).
// Pipeline of one batch (DESIGN.md "lexer pipeline"):
//   count (speculative start states) -> verify/fix-up rounds -> prefix sum -> emit -> finish events
//   -> resolve per document -> prefix sum -> write lexems
typedef struct SpScanParams
{
	SpLexerTables T;
	const uint8_t* text;		/* concatenated documents */
	const uint64_t* doc_off;	/* [ndocs+1] */
	uint32_t ndocs;
	uint64_t total;			/* = doc_off[ndocs] */
	uint32_t chunk;			/* bytes per scan thread */
	uint32_t warmup;		/* bytes scanned ahead of a chunk to guess its start state */
	uint64_t nchunks;
	uint32_t* start_state;		/* [nchunks] */
	uint32_t* end_state;		/* [nchunks] */
	uint32_t* ev_count;		/* [nchunks] */
	uint8_t* known;			/* [nchunks] chunk starts at a document start */
	uint8_t* redo;			/* [nchunks] chunk has to be scanned again from start_state */
	uint32_t* nbad;			/* number of chunks flagged by the last verification */
	const uint64_t* ev_off;		/* [nchunks+1] exclusive prefix sum of ev_count */
	SpRawEvent* events;
	uint32_t* ev_doc;		/* [nev] document of every event (written by the finish step) */
	uint64_t* doc_ev_begin;		/* [ndocs+1] first event of every document */
	SpMatchEvent* ar;		/* match event arrays, document d at doc_ev_begin[d]*ar_mult */
	uint32_t ar_mult;		/* 2 if symbols are defined (an event may insert two elements) else 1 */
	uint32_t* doc_kept;		/* [ndocs] elements of the match event array */
	uint32_t* doc_nlex;		/* [ndocs] lexems written */
	const uint64_t* lex_off;	/* [ndocs+1] exclusive prefix sum of doc_nlex */
	sp_lexem* lexems;
	uint32_t* status;		/* [ndocs] */
	uint32_t* head;			/* [nchunks] events of the chunk in front of its first cut, SP_HEAD_NOCUT if it has none */
} SpScanParams;

struct SpNullSink
{
	SP_HD void accept( uint32_t, uint32_t, uint32_t) {}
	SP_HD void docBegin( uint32_t) {}
	SP_HD void cut() {}
};

// scan_text: where the bytes [begin, end) of the chunk are read from (P.text, or a staged copy addressed like P.text)
SP_HD void spk_count_chunk( const SpScanParams& P, uint64_t j, int round, const uint8_t* scan_text)
{
	uint64_t begin = j * P.chunk, end = begin + P.chunk < P.total ? begin + P.chunk : P.total;
	uint32_t state;
	if (round == 0)
	{
		uint32_t d = sp_find_doc( P.doc_off, P.ndocs, begin);
		uint64_t dstart = SP_LDG( P.doc_off + d);
		state = P.T.fwd_init;
		P.known[ j] = (dstart == begin);
		P.redo[ j] = 0;
		if (dstart != begin)
		{
			uint64_t wb = (begin - dstart > P.warmup) ? begin - P.warmup : dstart;
			SpNullSink none;
			state = sp_scan_range( P.T, P.text, P.doc_off, P.ndocs, wb, begin, wb == dstart ? P.T.fwd_init : P.T.fwd_init, none);
		}
		P.start_state[ j] = state;
	}
	else
	{
		if (!P.redo[ j]) return;
		P.redo[ j] = 0;
		state = P.start_state[ j];
	}
	SpCountSink sink; sink.T = &P.T; sink.count = 0;
	P.end_state[ j] = sp_scan_range( P.T, scan_text, P.doc_off, P.ndocs, begin, end, state, sink);
	P.ev_count[ j] = sink.count;
}

// returns 1 if the chunk has to be scanned again
SP_HD int spk_verify_chunk( const SpScanParams& P, uint64_t j)
{
	if (j == 0 || P.known[ j]) return 0;
	uint32_t truth = P.end_state[ j-1];
	if (P.start_state[ j] == truth) return 0;
	P.start_state[ j] = truth;
	P.redo[ j] = 1;
	return 1;
}

SP_HD void spk_emit_chunk( const SpScanParams& P, uint64_t j, const uint8_t* scan_text)
{
	uint64_t begin = j * P.chunk, end = begin + P.chunk < P.total ? begin + P.chunk : P.total;
	SpEmitSink sink; sink.T = &P.T; sink.events = P.events; sink.cursor = P.ev_off[ j]; sink.doc_ev_begin = P.doc_ev_begin;
	sink.head = 0;
	sink.pendingCut = sink.seenCut = P.known[ j] ? 1 : 0;	// a chunk that starts with a document starts with a cut
	sp_scan_range( P.T, scan_text, P.doc_off, P.ndocs, begin, end, P.start_state[ j], sink);
	if (P.head) P.head[ j] = sink.seenCut ? sink.head : SP_HEAD_NOCUT;
}

SP_HD void spk_finish_event( const SpScanParams& P, uint64_t i)
{
	SpRawEvent ev = sp_load_event( P.events + i);
	if (P.ev_doc) P.ev_doc[ i] = P.head ? (ev.aux | (ev.from << 31)) : ev.aux;	// chunk resolution: bit 31 = first event behind a cut
	uint64_t dstart = SP_LDG( P.doc_off + ev.aux);
	uint32_t doclen = (uint32_t)(SP_LDG( P.doc_off + ev.aux + 1) - dstart);
	sp_finish_event( P.T, ev, P.text + dstart, doclen);
	sp_store_event( P.events + i, ev);
}

SP_HD void spk_resolve_doc( const SpScanParams& P, uint32_t d)
{
	uint64_t eb = P.doc_ev_begin[ d], ee = P.doc_ev_begin[ d+1];
	SpMatchEvent* ar = P.ar + eb * P.ar_mult;
	uint32_t n = 0, status = 0;
	if (P.doc_off[ d+1] - P.doc_off[ d] >= 0xFFFFFFFFull) status = (uint32_t)-SP_ERR_RUNTIME;	// src/patternLexer.cpp:866-869
	for (uint64_t i=eb; i<ee && !status; ++i)
	{
		SpRawEvent ev = P.events[ i];
		if (ev.pat == SP_PAT_DROPPED) continue;
		if (ev.pat == SP_PAT_SIZEERROR) {status = (uint32_t)-SP_ERR_RUNTIME; break;}
		sp_supersede( P.T, ev, ar, n);
	}
	if (status) n = 0;
	P.status[ d] = status;
	P.doc_kept[ d] = n;
	P.doc_nlex[ d] = sp_assign_ordpos( ar, n, 0);
}

SP_HD void spk_write_doc( const SpScanParams& P, uint32_t d)
{
	const SpMatchEvent* ar = P.ar + P.doc_ev_begin[ d] * P.ar_mult;
	sp_assign_ordpos( ar, P.doc_kept[ d], P.lexems + P.lex_off[ d]);
}

// ------------------------------------------------------------------------------------------------
// Parallel formulation of the resolution (DESIGN.md "islands"): the supersede logic only ever touches array
// elements that overlap the new event, so the event sequence of a document can be cut wherever every earlier
// event ends at or before the start of every later one; each such "island" is resolved independently and the
// results are concatenated. Ordinal positions are then three scans over the kept elements.

// Span of an event as the cut criterion sees it: [lo, r). Events of length 0 (empty selected sub expression)
// are widened by one byte on both sides so that they never sit on a cut.
SP_HD uint64_t sp_event_lo_key( const SpRawEvent& ev, uint32_t doc)
{
	if (ev.pat >= SP_PAT_SIZEERROR) return ~0ull;
	uint32_t lo = (ev.to == ev.from && ev.from) ? ev.from - 1 : ev.from;
	return ((uint64_t)doc << 32) | lo;
}
SP_HD uint64_t sp_event_r_key( const SpRawEvent& ev, uint32_t doc)
{
	if (ev.pat >= SP_PAT_SIZEERROR) return 0;
	uint32_t r = ev.to > ev.from ? ev.to : ev.from + 1;
	return ((uint64_t)doc << 32) | r;
}

#define SP_ISLAND_LOCAL 16	/* islands whose elements fit are resolved in thread local memory */
#define SP_META_BIND( meta) (((meta) >> 8) & 0xFF)
#define SP_META_INC 0x10000u	/* dense element: ordinal position advances here */
#define SP_DENSE_BIND( sizeword) (((sizeword) >> 16) & 0xFF)
#define SP_SIZE_EMIT 0x80000000u	/* dense element, origsize word: element is part of the output */

// Resolves the events [eb,ee) of one island into `ar`; returns the number of kept elements. A size error
// marks the document.
SP_HD uint32_t sp_resolve_island( const SpLexerTables& T, const SpRawEvent* events, uint64_t eb, uint64_t ee, SpMatchEvent* ar, uint32_t* doc_status)
{
	uint32_t n = 0;
	for (uint64_t i=eb; i<ee; ++i)
	{
		SpRawEvent ev = events[ i];
		if (ev.pat == SP_PAT_DROPPED) continue;
		if (ev.pat == SP_PAT_SIZEERROR) {*doc_status = (uint32_t)-SP_ERR_RUNTIME; continue;}
		sp_supersede( T, ev, ar, n);
	}
	return n;
}

// scan 1 (prefix max): key of an element that can advance the ordinal position, else 0.
// prev_* describe the element in front (same document) for the BindUnique rule (src/patternLexer.cpp:926-929).
SP_HD uint64_t sp_ord_key( uint32_t bind, uint32_t origpos, uint32_t doc, bool hasPrev, uint32_t prevBind)
{
	if (bind == SP_BIND_CONTENT || (bind == SP_BIND_UNIQUE && !(hasPrev && prevBind == SP_BIND_UNIQUE)))
	{
		return ((uint64_t)doc << 32) | ((uint64_t)origpos + 1);
	}
	return 0;
}

// after scan 2 (segmented count of advances, inclusive): final ordinal position and whether the element is emitted
SP_HD bool sp_ord_decide( uint32_t bind, bool counts, uint32_t cnt, uint32_t& ordpos)
{
	switch (bind)
	{
		case SP_BIND_CONTENT: case SP_BIND_UNIQUE: ordpos = cnt; return counts;
		case SP_BIND_SUCCESSOR: ordpos = cnt + 1; return true;
		default: ordpos = cnt; return cnt > 0;	// BindPredecessor: dropped in front of the first content element
	}
}


// ------------------------------------------------------------------------------------------------
// Slot formulation of the second half of the pipeline (DESIGN.md "lexer pipeline", steps 7-8): every island is
// resolved ONCE and its kept elements are written into the island's own slot range [eb*M, ee*M) (M = ar_mult;
// an island never keeps more than M elements per event), unused slots are marked empty. The decisions of the
// ordinal position passes that only need the island are taken right there:
//   COUNTS  the element can advance the position (BindContent, or BindUnique not behind a BindUnique; :926-934)
//   INC     it does advance it: its start lies behind every counting element in front of it. Elements of earlier
//           islands of the document start strictly in front of this island's first event, so the running
//           maximum is seeded with that start and no scan over the document is needed.
// One sum scan over the slots (low word: INC, high word: emitted) then yields ordinal positions and output indices.
typedef struct alignas(16) SpSlot
{
	uint32_t id;
	uint32_t origpos;
	uint32_t info;		/* origsize (16 bit) | bind << 16 | SP_SLOT_* flags */
	uint32_t doc;
} SpSlot;

#define SP_SLOT_VALID  0x01000000u
#define SP_SLOT_COUNTS 0x02000000u
#define SP_SLOT_INC    0x04000000u
#define SP_SLOT_BIND( info) (((info) >> 16) & 0xFF)
#define SP_NO_SLOT 0xFFFFFFFFu

#ifdef __CUDACC__
#define SP_ATOMIC_MIN( ptr, val) atomicMin( (ptr), (val))
#else
#define SP_ATOMIC_MIN( ptr, val) do {if ((val) < *(ptr)) *(ptr) = (val);} while (0)
#endif

// COUNTS/INC flags of the elements of one island, one element after the other. prevBind: bind of the element in
// front of the island in the same document (SP_BIND_CONTENT if there is none); minStart: smallest event start of the island.
typedef struct SpFlagState
{
	uint32_t prevBind, mx, ncounting;
} SpFlagState;

SP_HD void sp_flags_begin( SpFlagState& st, uint32_t prevBind, uint32_t minStart)
{
	st.prevBind = prevBind; st.mx = minStart; st.ncounting = 0;	// keys are origpos+1; earlier islands hold keys <= minStart
}
SP_HD uint32_t sp_flags_next( SpFlagState& st, uint32_t bind, uint32_t origpos)
{
	uint32_t f = SP_SLOT_VALID;
	if (bind == SP_BIND_CONTENT || (bind == SP_BIND_UNIQUE && st.prevBind != SP_BIND_UNIQUE))
	{
		f |= SP_SLOT_COUNTS;
		st.ncounting += 1;
		if (origpos + 1 > st.mx) {st.mx = origpos + 1; f |= SP_SLOT_INC;}
	}
	st.prevBind = bind;
	return f;
}

typedef struct SpSlotParams
{
	SpLexerTables T;
	const SpRawEvent* events;
	const uint32_t* ev_doc;
	uint64_t nev;
	uint32_t ar_mult;
	SpSlot* slots;			/* [nev*ar_mult] */
	const uint32_t* isl_start;	/* [nisl+1] */
	uint64_t nisl;
	uint32_t* doc_status;		/* [ndocs] */
	uint8_t* doc_has_content;	/* [ndocs] */
	uint32_t* doc_first_content;	/* [ndocs] slot of the first counting element (SP_NO_SLOT = none) */
	uint32_t* fix_list;		/* islands whose first element is BindUnique: flags depend on the element in front */
	uint32_t* fix_count;
	const uint64_t* doc_ev_begin;	/* [ndocs+1] */
	uint32_t ndocs;
	uint64_t* doc_base;		/* [ndocs+1] scan value in front of the first slot of every document */
	sp_lexem* lexems;
	uint64_t* lex_off;		/* [ndocs+1] */
} SpSlotParams;

// Resolves island i into its slots. `ar` is a work array of (ee-eb)*M elements: thread local memory, or the slot
// range of the island itself (same record size; element k is read before slot k is written).
// Returns 1 if the island goes on the fix list.
SP_HD int sp_island_to_slots( const SpSlotParams& P, uint64_t i, SpMatchEvent* ar)
{
	const uint64_t eb = P.isl_start[ i], ee = P.isl_start[ i+1];
	const uint32_t doc = P.ev_doc[ eb], M = P.ar_mult;
	uint32_t n = 0, minStart = SP_NO_POS;
	for (uint64_t e=eb; e<ee; ++e)
	{
		SpRawEvent ev = P.events[ e];
		if (ev.pat == SP_PAT_DROPPED) continue;
		if (ev.pat == SP_PAT_SIZEERROR) {P.doc_status[ doc] = (uint32_t)-SP_ERR_RUNTIME; continue;}
		if (ev.from < minStart) minStart = ev.from;
		sp_supersede( P.T, ev, ar, n);
	}
	if (minStart == SP_NO_POS) minStart = 0;
	SpSlot* out = P.slots + eb * M;
	const uint32_t cap = (uint32_t)(ee - eb) * M;
	SpFlagState st; sp_flags_begin( st, SP_BIND_CONTENT, minStart);
	uint32_t first = SP_NO_SLOT, firstBind = SP_BIND_CONTENT;
	for (uint32_t k=0; k<cap; ++k)
	{
		SpSlot s; s.id = 0; s.origpos = 0; s.info = 0; s.doc = doc;
		if (k < n)
		{
			SpMatchEvent e = ar[ k];
			uint32_t bind = SP_META_BIND( e.meta);
			if (k == 0) firstBind = bind;
			uint32_t f = sp_flags_next( st, bind, e.origpos);
			if ((f & SP_SLOT_COUNTS) && first == SP_NO_SLOT) first = k;
			s.id = e.id; s.origpos = e.origpos; s.info = (e.origsize & 0xFFFFu) | (bind << 16) | f;
		}
		out[ k] = s;
	}
	if (first != SP_NO_SLOT)
	{
		// most islands lose against an earlier one: look before the atomic (the value only ever decreases)
		uint32_t cand = (uint32_t)(eb * M) + first;
		if (!P.doc_has_content[ doc]) P.doc_has_content[ doc] = 1;
		if (cand < P.doc_first_content[ doc]) SP_ATOMIC_MIN( P.doc_first_content + doc, cand);
	}
	// a leading BindUnique element is dropped if the element in front of the island is BindUnique too
	bool firstOfDoc = (eb == 0 || P.ev_doc[ eb-1] != doc);
	return (n && !firstOfDoc && firstBind == SP_BIND_UNIQUE) ? 1 : 0;
}

// Second look at an island that starts with a BindUnique element: bind of the last valid slot in front of it.
SP_HD void sp_island_fix_unique( const SpSlotParams& P, uint64_t i)
{
	const uint64_t eb = P.isl_start[ i], ee = P.isl_start[ i+1];
	const uint32_t M = P.ar_mult, doc = P.ev_doc[ eb];
	uint64_t s = eb * M;
	uint32_t prevBind = SP_BIND_CONTENT;
	while (s > 0)
	{
		SpSlot prev = P.slots[ s-1];
		if (prev.doc != doc) break;
		if (prev.info & SP_SLOT_VALID) {prevBind = SP_SLOT_BIND( prev.info); break;}
		--s;
	}
	if (prevBind != SP_BIND_UNIQUE) return;
	SpSlot* slots = P.slots + eb * M;
	const uint32_t cap = (uint32_t)(ee - eb) * M;
	uint32_t minStart = SP_NO_POS;
	for (uint64_t e=eb; e<ee; ++e) {SpRawEvent ev = P.events[ e]; if (ev.pat < SP_PAT_SIZEERROR && ev.from < minStart) minStart = ev.from;}
	if (minStart == SP_NO_POS) minStart = 0;
	SpFlagState st; sp_flags_begin( st, prevBind, minStart);
	for (uint32_t k=0; k<cap && (slots[ k].info & SP_SLOT_VALID); ++k)
	{
		uint32_t info = slots[ k].info;
		slots[ k].info = (info & 0x00FFFFFFu) | sp_flags_next( st, SP_SLOT_BIND( info), slots[ k].origpos);
	}
}

// Scan value of slot j: low word 1 if the position advances here, high word 1 if the element is part of the output.
SP_HD uint64_t sp_slot_value( const SpSlotParams& P, uint64_t j, const SpSlot& s)
{
	if (!(s.info & SP_SLOT_VALID)) return 0;
	uint64_t v = (s.info & SP_SLOT_INC) ? 1 : 0;
	if (SP_LDG( P.doc_status + s.doc) || !SP_LDG( P.doc_has_content + s.doc)) return v;
	bool emit;
	switch (SP_SLOT_BIND( s.info))
	{
		case SP_BIND_CONTENT: emit = true; break;
		case SP_BIND_UNIQUE: emit = (s.info & SP_SLOT_COUNTS) != 0; break;
		case SP_BIND_SUCCESSOR: emit = true; break;
		default: emit = j > SP_LDG( P.doc_first_content + s.doc); break;	// BindPredecessor: dropped in front of the first content element
	}
	return emit ? (v | (1ull << 32)) : v;
}

// Writes the lexem of slot j given the exclusive scan value in front of it.
SP_HD void sp_slot_emit( const SpSlotParams& P, uint64_t j, const SpSlot& s, uint64_t excl, uint64_t value)
{
	if (!(value >> 32)) return;
	uint32_t cnt = (uint32_t)(excl + value) - (uint32_t)SP_LDG( P.doc_base + s.doc);	// positions advanced in the document up to here
	sp_lexem lx; lx.id = s.id; lx.origpos = s.origpos; lx.origsize = s.info & 0xFFFFu;
	lx.ordpos = (SP_SLOT_BIND( s.info) == SP_BIND_SUCCESSOR) ? cnt + 1 : cnt;
	P.lexems[ excl >> 32] = lx;
}

#endif
