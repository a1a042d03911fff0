// Fused second half of the lexer (DESIGN.md "fused segment kernel"): the thread that scans a text chunk also
// completes, resolves and flags the events it finds, so that raw events never travel through HBM.
//
// Segments. The forward automaton tells apart the states that are reached by a transition which no older
// thread survived (SpLexerTables::fwd_cut_lo/fwd_cut_n, automaton.hpp Dfa::cut). Behind such a byte at position
// p every match that is still pending started at p or later, and every match reported so far ended at p or
// before: no event spans p. The supersede logic of the reference (src/patternLexer.cpp:754-822) only ever
// touches elements that overlap the new event, so the event sequence can be cut at p and the pieces
// ("segments", unions of the islands of lexer_islands.cuh) are resolved independently, in event order.
//
// Ownership. The thread of chunk j owns the segments that BEGIN in its chunk: it skips (only counts) the events in
// front of the first cut of its chunk, and scans on behind the end of its chunk until the segment that is open
// there closes. Kept elements go densely into the slot range that starts at the index of the thread's first
// owned event (times ar_mult): ranges of different threads are disjoint and large enough by construction.
//
// Per chunk the thread leaves the number of kept elements and the pair (positions advanced, lexems emitted);
// one exclusive scan over the chunks then gives every chunk its ordinal position / output base, and
// k_fused_write turns the kept elements into sp_lexem records (src/patternLexer.cpp:893-945).
//
// Not handled here (need_legacy is raised and the batch runs through the island pipeline instead): matched
// terms of 65535 bytes or more, empty selected sub expressions, segments reaching further than SP_FUSED_MAXREL bytes
// behind the begin of their chunk (fused scan only), a BindUnique element as first element of a thread's
// range, BindSuccessor elements in a document without content element.
#ifndef SP_LEXER_FUSED_CUH_INCLUDED
#define SP_LEXER_FUSED_CUH_INCLUDED
#include "lexer_device.cuh"

#define SP_FUSED_BUF 24			/* pending entries per thread */
#define SP_FUSED_AR 8			/* elements of an open segment kept in the thread's fast work memory */
#define SP_FUSED_MEMO 4			/* memorized reverse walks per thread (power of two) */
#define SP_FUSED_MAXREL 65000u		/* a thread never scans further than this behind the begin of its chunk */
#define SP_FUSED_CUTMARK 0x80000000u	/* pending entry: segment ends here; else definition | relative end << 15 */
#define SP_FUSED_MAXPAT 32767u
#define SP_EVDOC_CUT 0x80000000u

typedef struct SpFusedParams
{
	SpLexerTables T;
	const uint8_t* text;
	const uint64_t* doc_off;	/* [ndocs+1] */
	uint32_t ndocs;
	uint64_t total;
	uint32_t chunk;
	uint64_t nchunks;
	const uint32_t* start_state;	/* [nchunks] verified start states (count/verify rounds) */
	const uint64_t* ev_off;		/* [nchunks+1] index of the first raw event of every chunk */
	uint32_t ar_mult;
	SpSlot* slots;			/* [nev*ar_mult] kept elements, dense per chunk from region_begin */
	/* per chunk */
	uint64_t* region_begin;		/* first slot of the chunk's range */
	uint32_t* kept;			/* elements written */
	uint64_t* agg;			/* positions advanced | lexems emitted << 32 */
	uint32_t* lead_pred;		/* BindPredecessor elements in front of the first advance whose document began earlier;
					   after the fix step: those that are NOT emitted (0 if they are) */
	uint32_t* inc_tail;		/* advances since the last document begin in the range (or in the whole range) */
	uint8_t* has_docstart;		/* range begins at or passes a document begin */
	const uint64_t* prefix;		/* [nchunks+1] exclusive scan of agg */
	/* per document */
	uint32_t* doc_chunk;		/* chunk whose thread passed the begin of the document */
	uint64_t* doc_local;		/* that thread's (advanced | emitted << 32) at this point */
	uint8_t* doc_has_content;
	uint8_t* doc_succ_early;	/* a BindSuccessor element was seen in front of the first advance some thread saw */
	uint64_t* doc_base;		/* [ndocs+1] scan value in front of the first element of the document */
	uint64_t* lex_off;		/* [ndocs+1] */
	uint32_t* need_legacy;
	sp_lexem* lexems;
	/* chunk resolution over completed events (spk_resolve_chunk) */
	const SpRawEvent* events;	/* [nev] completed by the finish step */
	const uint32_t* ev_doc;		/* [nev] document | SP_EVDOC_CUT if the event is the first one behind a cut */
	const uint32_t* head;		/* [nchunks] */
	uint64_t nev;
} SpFusedParams;

// Per thread work memory lives outside the thread state so that the kernel can keep it in shared memory, element i of
// thread t at [i*stride + t]: all lanes of a warp hit different banks whatever their indices are.
// Elements SP_FUSED_AR and up of an open segment (rare: long chains of overlapping matches) live in the slots they will
// end up in: the segment's kept elements go to the slots from *out on, element k is read before slot k is written, and
// the thread's slot range has room for ar_mult elements per event.
struct SpStridedEvents
{
	SpMatchEvent* base; uint32_t stride;
	SpMatchEvent* spill;	/* = slots + out of the thread, kept up to date by sp_fused_set_out (a pointer to the thread's `out`
				   would force the whole thread state into local memory) */
	// element access by value through a proxy: a reference would have to point into either of two address spaces and
	// make every access a generic one; this way the common case stays a plain shared memory access
	struct Element
	{
		const SpStridedEvents* a; uint32_t i;
		SP_HD operator SpMatchEvent() const {return i < SP_FUSED_AR ? a->base[ i * a->stride] : a->spill[ i];}
		SP_HD const Element& operator=( const SpMatchEvent& v) const
		{
			if (i < SP_FUSED_AR) a->base[ i * a->stride] = v; else a->spill[ i] = v;
			return *this;
		}
	};
	SP_HD Element operator[]( uint32_t i) const {Element e; e.a = this; e.i = i; return e;}
};

typedef struct SpFusedThread
{
	uint32_t* buf; uint32_t bstride;	/* pending entries */
	SpStridedEvents ar;			/* elements of the open segment */
	SpStartMemo* memo;			/* SP_FUSED_MEMO last reverse walks by definition & (SP_FUSED_MEMO-1), stride bstride */
	uint32_t nbuf, n, minStart;
	uint32_t prevBind, prevUnknown;
	uint32_t docinc, docSeenBegin;
	uint32_t inc, emit, lead, nkept;
	uint64_t out;
	uint32_t doc, doclen, relbase, contentDoc;
	const uint8_t* docptr;
	uint32_t legacy;
} SpFusedThread;

struct SpFusedThread;
SP_HD void sp_fused_set_out( const struct SpFusedParams& P, struct SpFusedThread& st, uint64_t out);

SP_HD void sp_store_slot( SpSlot* dst, const SpSlot& s)
{
#ifdef __CUDACC__
	*reinterpret_cast<uint4*>( dst) = make_uint4( s.id, s.origpos, s.info, s.doc);	// slot ranges start at multiples of 16 bytes
#else
	*dst = s;
#endif
}

SP_HD void sp_fused_set_out( const SpFusedParams& P, SpFusedThread& st, uint64_t out)
{
	st.out = out; st.ar.spill = reinterpret_cast<SpMatchEvent*>( P.slots + out);
}

// Closes the open segment: COUNTS/INC flags (src/patternLexer.cpp:926-934), slot records, the thread's counters.
SP_HD void sp_fused_finalize( const SpFusedParams& P, SpFusedThread& st)
{
	if (st.n)
	{
		SpFlagState fs; sp_flags_begin( fs, st.prevUnknown ? SP_BIND_CONTENT : st.prevBind, st.minStart == SP_NO_POS ? 0 : st.minStart);
		for (uint32_t k=0; k<st.n; ++k)
		{
			const SpMatchEvent e = st.ar[ k];
			const uint32_t bind = SP_META_BIND( e.meta);
			if (k == 0 && st.prevUnknown && bind == SP_BIND_UNIQUE) st.legacy = 1;	// depends on the element in front of the range
			const uint32_t f = sp_flags_next( fs, bind, e.origpos);
			SpSlot s; s.id = e.id; s.origpos = e.origpos; s.info = (e.origsize & 0xFFFFu) | (bind << 16) | f; s.doc = st.doc;
			sp_store_slot( P.slots + st.out + k, s);	// st.out stays until the loop is through: the accessor's spill range starts there
			if (f & SP_SLOT_INC) {st.inc += 1; st.docinc += 1;}
			if ((f & SP_SLOT_COUNTS) && st.contentDoc != st.doc) {P.doc_has_content[ st.doc] = 1; st.contentDoc = st.doc;}
			switch (bind)
			{
				case SP_BIND_CONTENT: st.emit += 1; break;
				case SP_BIND_UNIQUE: if (f & SP_SLOT_COUNTS) st.emit += 1; break;
				case SP_BIND_SUCCESSOR: st.emit += 1; if (!st.docinc) P.doc_succ_early[ st.doc] = 1; break;
				default:
					// BindPredecessor: dropped in front of the first content element of the document (:893-922)
					if (st.docinc) st.emit += 1;
					else if (!st.docSeenBegin) {st.emit += 1; st.lead += 1;}	// decided by the fix step
					break;
			}
		}
		st.nkept += st.n; sp_fused_set_out( P, st, st.out + st.n);
		st.prevBind = fs.prevBind; st.prevUnknown = 0;
	}
	st.n = 0; st.minStart = SP_NO_POS;
}

// Resolves the queued entries in order: events are completed (leftmost start with the memo of the definition's last
// walk, sub expression, symbol) and go through the supersede logic, cut marks close the segment.
SP_HD void sp_fused_drain( const SpFusedParams& P, SpFusedThread& st)
{
	const uint32_t M = P.ar_mult;
	for (uint32_t i=0; i<st.nbuf; ++i)
	{
		const uint32_t e = st.buf[ i * st.bstride];
		if (e & SP_FUSED_CUTMARK) {sp_fused_finalize( P, st); continue;}
		SpRawEvent ev; ev.pat = e & SP_FUSED_MAXPAT; ev.from = SP_NO_POS; ev.to = (e >> 15) + st.relbase; ev.aux = 0;
		const SpPatternInfo I = P.T.pat[ ev.pat];
		SpStartMemo* mp = st.memo + (ev.pat & (SP_FUSED_MEMO-1)) * st.bstride;
		SpStartMemo m = *mp;
		const uint32_t from = sp_leftmost_start_memo( P.T, ev.pat, I, st.docptr, st.doclen, ev.to, m);
		*mp = m;
		sp_finish_event_from( P.T, I, ev, from, st.docptr, st.doclen);
		if (ev.pat == SP_PAT_DROPPED) continue;
		if (ev.pat == SP_PAT_SIZEERROR || ev.from == ev.to) {st.legacy = 1; continue;}
		if (ev.from < st.minStart) st.minStart = ev.from;
		sp_supersede( P.T, ev, st.ar, st.n);
	}
	st.nbuf = 0;
}

// Scan state of the thread while it walks the text.
typedef struct SpFusedWalk
{
	uint32_t state, own, overhang, done, finished, docend;
	uint64_t evidx, begin, end, pos;
	uint64_t region_begin;		/* first slot of the thread's range (valid once own is set) */
	uint32_t d;			/* current document */
	uint64_t dstart, dend;
	uint32_t roomMax;		/* a byte may be taken while nbuf <= roomMax */
} SpFusedWalk;

// The state just entered: queues its events (or counts them while the thread does not own the text yet) and handles a
// cut. The caller guarantees room for fwd_max_acc + 1 entries.
SP_HD void sp_fused_arrive( const SpFusedParams& P, SpFusedThread& st, SpFusedWalk& w, bool cut)
{
	const SpLexerTables& T = P.T;
	if (w.own)
	{
		if (w.state >= T.fwd_acc_base)
		{
			const uint32_t a = w.state - T.fwd_acc_base, rel = (uint32_t)(w.pos - w.begin) << 15;
			uint32_t pi = SP_LDG( T.fwd_acc_off + a), pe = SP_LDG( T.fwd_acc_off + a + 1);
			for (; pi < pe; ++pi) st.buf[ (st.nbuf++) * st.bstride] = (uint32_t)SP_LDG( T.fwd_acc_pat + pi) | rel;
		}
		if (cut)
		{
			st.buf[ (st.nbuf++) * st.bstride] = SP_FUSED_CUTMARK;
			if (w.overhang) w.done = 1;
		}
	}
	else
	{
		if (w.state >= T.fwd_acc_base)
		{
			const uint32_t a = w.state - T.fwd_acc_base;
			w.evidx += SP_LDG( T.fwd_acc_off + a + 1) - SP_LDG( T.fwd_acc_off + a);
		}
		if (cut)
		{
			// first cut of the chunk: everything reported up to here belongs to the thread in front
			w.own = 1;
			w.region_begin = w.evidx * P.ar_mult; sp_fused_set_out( P, st, w.region_begin);
			st.prevUnknown = 1; st.docSeenBegin = 0;
		}
	}
}

// One byte of text.
SP_HD void sp_fused_step( const SpFusedParams& P, SpFusedThread& st, SpFusedWalk& w, uint32_t b)
{
	const SpLexerTables& T = P.T;
	w.state = SP_LDG( T.fwd_next + w.state * T.fwd_stride + T.fwd_class[ b]);
	sp_fused_arrive( P, st, w, (w.state - T.fwd_cut_lo) < T.fwd_cut_n);
	++w.pos;
}

SP_HD void sp_fused_set_doc( const SpFusedParams& P, SpFusedThread& st, SpFusedWalk& w, uint32_t d)
{
	w.d = d; w.dstart = SP_LDG( P.doc_off + d); w.dend = SP_LDG( P.doc_off + d + 1);
	st.doc = d; st.docptr = P.text + w.dstart; st.doclen = (uint32_t)(w.dend - w.dstart); st.relbase = (uint32_t)(w.begin - w.dstart);
	for (uint32_t k=0; k<SP_FUSED_MEMO; ++k) st.memo[ k * st.bstride].pat = SP_NO_POS;	// offsets are document relative
}

// Start of the work of the thread of chunk j. buf: SP_FUSED_BUF entries, ar: SP_FUSED_AR elements (both strided).
SP_HD void sp_fused_init( const SpFusedParams& P, uint64_t j, SpFusedThread& st, SpFusedWalk& w, uint32_t* buf, uint32_t bstride, SpStridedEvents ar, SpStartMemo* memo)
{
	st.buf = buf; st.bstride = bstride; st.ar = ar; st.memo = memo;
	st.nbuf = 0; st.n = 0; st.minStart = SP_NO_POS;
	st.prevBind = SP_BIND_CONTENT; st.prevUnknown = 0; st.docinc = 0; st.docSeenBegin = 0;
	st.inc = st.emit = st.lead = st.nkept = 0; st.legacy = 0; st.contentDoc = 0xFFFFFFFFu;
	w.begin = j * P.chunk; w.end = w.begin + P.chunk < P.total ? w.begin + P.chunk : P.total; w.pos = w.begin;
	sp_fused_set_doc( P, st, w, sp_find_doc( P.doc_off, P.ndocs, w.begin));
	w.state = P.start_state[ j]; w.own = 0; w.overhang = 0; w.done = 0; w.finished = 0; w.docend = 0; w.evidx = P.ev_off[ j];
	w.region_begin = w.evidx * P.ar_mult; sp_fused_set_out( P, st, w.region_begin);
	if (w.dstart == w.begin) {w.state = P.T.fwd_init; w.own = 1; st.docSeenBegin = 1;}
	w.roomMax = SP_FUSED_BUF - 1 - P.T.fwd_max_acc;
}

// Walk phase: takes bytes while the queue has room, never across the end of the document and, unless the segment
// that is open at the end of the chunk has to be finished, never across the end of the chunk.
SP_HD void sp_fused_walk( const SpFusedParams& P, SpFusedThread& st, SpFusedWalk& w)
{
	const SpLexerTables& T = P.T;
	const uint64_t hard = (w.overhang || w.dend < w.end) ? w.dend : w.end;
	w.docend = 0;
	while (w.pos < hard && !w.done && st.nbuf <= w.roomMax)
	{
		if (!w.overhang && w.pos + 8 <= hard && (((uintptr_t)(P.text + w.pos)) & 7) == 0)
		{
			const uint2 v = *reinterpret_cast<const uint2*>( P.text + w.pos);
			#pragma unroll
			for (int bi=0; bi<8; ++bi)
			{
				if (st.nbuf > w.roomMax) break;
				sp_fused_step( P, st, w, ((bi < 4 ? v.x : v.y) >> (8*(bi&3))) & 0xFF);
			}
		}
		else sp_fused_step( P, st, w, SP_LDG( P.text + w.pos));
	}
	if (w.pos == w.dend && !w.done && st.nbuf <= w.roomMax)
	{
		// end of the document: the virtual end-of-text symbol reports the matches that end here and closes the segment
		w.state = SP_LDG( T.fwd_next + w.state * T.fwd_stride + (T.fwd_stride-1));
		sp_fused_arrive( P, st, w, w.own != 0);
		w.done = 0;
		w.docend = 1;
	}
}

// What follows the drain of a walk phase: end conditions, switch to the next document, overhang.
SP_HD void sp_fused_after( const SpFusedParams& P, uint64_t j, SpFusedThread& st, SpFusedWalk& w)
{
	if (w.done) {w.finished = 1; return;}
	if (w.pos - w.begin >= SP_FUSED_MAXREL) {st.legacy = 1; w.finished = 1; return;}
	if (w.docend)
	{
		if (w.overhang) {w.finished = 1; return;}
		// the next document begins here: its first segment belongs to this thread if it lies in this chunk
		w.state = P.T.fwd_init;
		if (!w.own) {w.own = 1; w.region_begin = w.evidx * P.ar_mult; sp_fused_set_out( P, st, w.region_begin);}
		st.prevBind = SP_BIND_CONTENT; st.prevUnknown = 0; st.docinc = 0; st.docSeenBegin = 1;
		uint32_t d = w.d;
		for (++d; d < P.ndocs; ++d)
		{
			P.doc_chunk[ d] = (uint32_t)j;
			P.doc_local[ d] = (uint64_t)st.inc | ((uint64_t)st.emit << 32);
			if (SP_LDG( P.doc_off + d + 1) > w.pos) break;	// next non empty document
		}
		if (d >= P.ndocs) {w.finished = 1; return;}
		sp_fused_set_doc( P, st, w, d);
		if (w.pos >= w.end) w.finished = 1;
	}
	else if (w.pos >= w.end && w.pos != w.dend)
	{
		if (!w.own) w.finished = 1;	// no cut in the whole chunk: everything belongs to a thread in front
		else w.overhang = 1;
	}
}

SP_HD void sp_fused_store( const SpFusedParams& P, uint64_t j, const SpFusedThread& st, const SpFusedWalk& w)
{
	P.region_begin[ j] = w.region_begin;
	P.kept[ j] = st.nkept;
	P.agg[ j] = (uint64_t)st.inc | ((uint64_t)st.emit << 32);
	P.lead_pred[ j] = st.lead;
	P.inc_tail[ j] = st.docinc;
	P.has_docstart[ j] = (uint8_t)st.docSeenBegin;
	if (st.legacy) *P.need_legacy = 1;
}

// Work of the thread of chunk j. buf: SP_FUSED_BUF entries, ar: SP_FUSED_AR elements, memo: SP_FUSED_MEMO records, all
// strided by bstride (1 = plain arrays; the kernel keeps them in shared memory with the block size as stride).
SP_HD void spk_fused_chunk( const SpFusedParams& P, uint64_t j, uint32_t* buf, uint32_t bstride, SpStridedEvents ar, SpStartMemo* memo)
{
	SpFusedThread st; SpFusedWalk w;
	sp_fused_init( P, j, st, w, buf, bstride, ar, memo);
	while (!w.finished)
	{
		sp_fused_walk( P, st, w);
		sp_fused_drain( P, st);
		sp_fused_after( P, j, st, w);
	}
	sp_fused_store( P, j, st, w);
}

// ------------------------------------------------------------------------------------------------
// Chunk resolution over completed events: the second half of the fused formulation alone. The emit step has marked the
// events behind a cut (bit 31 of ev_doc) and counted the events in front of the first cut of every chunk (head); the finish
// step has completed the events. The thread of chunk j resolves the events from its first cut up to the first cut of
// the next chunk that has one, exactly like the fused thread does, and leaves the same per chunk / per document records.
typedef struct SpChunkRange
{
	uint64_t a, b;		/* events of the thread */
	uint32_t curdoc;	/* document of the event in front of the current one */
} SpChunkRange;

SP_HD void sp_resolve_begin( const SpFusedParams& P, uint64_t j, SpFusedThread& st, SpChunkRange& R, SpStridedEvents ar)
{
	st.buf = 0; st.bstride = 1; st.ar = ar; st.memo = 0;
	st.nbuf = 0; st.n = 0; st.minStart = SP_NO_POS;
	st.prevBind = SP_BIND_CONTENT; st.prevUnknown = 1; st.docinc = 0; st.docSeenBegin = 0;
	st.inc = st.emit = st.lead = st.nkept = 0; st.legacy = 0; st.contentDoc = 0xFFFFFFFFu;
	st.docptr = 0; st.doclen = 0; st.relbase = 0;
	const uint32_t h = P.head[ j];
	R.a = P.ev_off[ j+1]; R.b = R.a;
	if (h != SP_HEAD_NOCUT)
	{
		R.a = P.ev_off[ j] + h;
		uint64_t k = j + 1;
		while (k < P.nchunks && P.head[ k] == SP_HEAD_NOCUT) ++k;
		R.b = (k < P.nchunks) ? P.ev_off[ k] + P.head[ k] : P.nev;
	}
	sp_fused_set_out( P, st, R.a * P.ar_mult);
	R.curdoc = R.a ? (P.ev_doc[ R.a-1] & ~SP_EVDOC_CUT) : 0xFFFFFFFFu;
	st.doc = R.curdoc;
}

// One completed event of the thread's range; dflag = its ev_doc word.
SP_HD void sp_resolve_event( const SpFusedParams& P, uint64_t j, SpFusedThread& st, SpChunkRange& R, const SpRawEvent& ev, uint32_t dflag)
{
	const uint32_t d = dflag & ~SP_EVDOC_CUT;
	if (d != R.curdoc)
	{
		// first event of its document: the documents up to it begin here as far as the output is concerned
		sp_fused_finalize( P, st);
		for (uint32_t dd=R.curdoc+1; dd<=d; ++dd)
		{
			P.doc_chunk[ dd] = (uint32_t)j;
			P.doc_local[ dd] = (uint64_t)st.inc | ((uint64_t)st.emit << 32);
		}
		R.curdoc = d; st.doc = d;
		st.prevBind = SP_BIND_CONTENT; st.prevUnknown = 0; st.docinc = 0; st.docSeenBegin = 1;
	}
	else if (dflag & SP_EVDOC_CUT) sp_fused_finalize( P, st);
	if (ev.pat == SP_PAT_DROPPED) return;
	if (ev.pat == SP_PAT_SIZEERROR || ev.from == ev.to) {st.legacy = 1; return;}
	if (ev.from < st.minStart) st.minStart = ev.from;
	sp_supersede( P.T, ev, st.ar, st.n);
}

// The events [0,left) of the thread's stream as `src` shows them (src.ev(k), src.df(k)); streamEnds: the stream ends
// behind them. (A separate short way for segments of one or two events was measured and dropped: two code paths in
// one warp cost more than the loops of the general one, 35 ms against 23 ms per GiB.)
template <class Source>
SP_HD void sp_resolve_view( const SpFusedParams& P, uint64_t j, SpFusedThread& st, SpChunkRange& R, const Source& src, uint32_t left, bool streamEnds)
{
	(void)streamEnds;
	for (uint32_t k=0; k<left; ++k) sp_resolve_event( P, j, st, R, src.ev( k), src.df( k));
}

struct SpEventStream	/* completed events in HBM from index `first` on */
{
	const SpRawEvent* events; const uint32_t* ev_doc; uint64_t first;
	SP_HD SpRawEvent ev( uint32_t k) const {return events[ first + k];}
	SP_HD uint32_t df( uint32_t k) const {return ev_doc[ first + k];}
};

SP_HD void sp_resolve_end( const SpFusedParams& P, uint64_t j, SpFusedThread& st, const SpChunkRange& R)
{
	sp_fused_finalize( P, st);
	P.region_begin[ j] = R.a * P.ar_mult;
	P.kept[ j] = st.nkept;
	P.agg[ j] = (uint64_t)st.inc | ((uint64_t)st.emit << 32);
	P.lead_pred[ j] = st.lead;
	P.inc_tail[ j] = st.docinc;
	P.has_docstart[ j] = (uint8_t)st.docSeenBegin;
	if (st.legacy) *P.need_legacy = 1;
}

// Sequential form (tests/hostsim); k_lex_resolve_chunks feeds the same functions from shared memory tiles that the
// warp fills with coalesced loads.
SP_HD void spk_resolve_chunk( const SpFusedParams& P, uint64_t j, SpStridedEvents ar)
{
	SpFusedThread st; SpChunkRange R;
	sp_resolve_begin( P, j, st, R, ar);
	// in views of at most 8 events like the kernel's tiles, so that the borders between views are exercised as well
	for (uint64_t i=R.a; i<R.b; i+=8)
	{
		SpEventStream src; src.events = P.events; src.ev_doc = P.ev_doc; src.first = i;
		const uint32_t left = (uint32_t)(R.b - i < 8 ? R.b - i : 8);
		sp_resolve_view( P, j, st, R, src, left, i + left == R.b);
	}
	sp_resolve_end( P, j, st, R);
}

// Fix step: BindPredecessor elements at the begin of a range are emitted iff the document has advanced its position
// before; look at the ranges in front until one answers.
SP_HD void spk_fused_fix_chunk( const SpFusedParams& P, uint64_t j)
{
	const uint32_t lead = P.lead_pred[ j];
	if (!lead) return;
	for (uint64_t k=j; k>0;)
	{
		--k;
		if (P.inc_tail[ k]) {P.lead_pred[ j] = 0; return;}
		if (P.has_docstart[ k]) break;
	}
	P.agg[ j] -= (uint64_t)lead << 32;
}

// Scan value in front of the first element of document d (d == ndocs: the total), output offsets, late checks.
SP_HD void spk_fused_doc( const SpFusedParams& P, uint32_t d, uint64_t total)
{
	uint64_t base = total;
	if (d < P.ndocs)
	{
		const uint32_t c = P.doc_chunk[ d];
		if (c != 0xFFFFFFFFu) base = P.prefix[ c] + P.doc_local[ d] - ((uint64_t)P.lead_pred[ c] << 32);	// else: behind the last event
		if (P.doc_succ_early[ d] && !P.doc_has_content[ d]) *P.need_legacy = 1;
	}
	P.doc_base[ d] = base;
	P.lex_off[ d] = base >> 32;
}

// Emit decision and ordinal position of one kept element; cnt = positions advanced in its document up to and including it.
SP_HD bool sp_fused_lexem( const SpSlot& s, uint32_t cnt, sp_lexem& lx)
{
	bool emit;
	const uint32_t bind = SP_SLOT_BIND( s.info);
	switch (bind)
	{
		case SP_BIND_CONTENT: emit = true; break;
		case SP_BIND_UNIQUE: emit = (s.info & SP_SLOT_COUNTS) != 0; break;
		case SP_BIND_SUCCESSOR: emit = true; break;
		default: emit = cnt > 0; break;
	}
	lx.id = s.id; lx.origpos = s.origpos; lx.origsize = s.info & 0xFFFFu;
	lx.ordpos = (bind == SP_BIND_SUCCESSOR) ? cnt + 1 : cnt;
	return emit;
}

// Sequential form of the write step for one chunk (tests/hostsim; the kernel does the same with one warp per chunk).
SP_HD void spk_fused_write_chunk( const SpFusedParams& P, uint64_t j)
{
	const uint64_t rb = P.region_begin[ j], pre = P.prefix[ j];
	const uint32_t n = P.kept[ j];
	uint32_t incp = (uint32_t)pre;
	uint64_t emitp = pre >> 32;
	for (uint32_t k=0; k<n; ++k)
	{
		const SpSlot s = P.slots[ rb + k];
		if (s.info & SP_SLOT_INC) ++incp;
		sp_lexem lx;
		if (sp_fused_lexem( s, incp - (uint32_t)P.doc_base[ s.doc], lx)) P.lexems[ emitp++] = lx;
	}
}

#endif
