// CUDA kernels of the window evaluator (matcher_eval.cuh) for sm_100a.
//
// A "work item" is one rule instance: (key token, site, entry of the site's program list). The items of a batch are
// numbered in document order (token, then the token event before its follow events, then list order); that order
// is also the order of the results inside a document.
//
//   k_eval_prepare      per token: event slot and number of work items; irregular tokens flag their document.
//   (prefix sum: first work item of every token)
//   k_eval_count        one warp per run of 256 tokens. The tokens of the run (+ 64 behind it) are staged in shared
//                       memory as (id, ordinal position). Every lane owns a contiguous block of the run's work items and
//                       follows one rule at a time through its window from the compiled list entry (sp_eval_lite_step:
//                       32 bytes per rule, no program / trigger definition loads); a lane whose rule has ended takes its
//                       next item at once, so the lanes stay busy whatever the lengths of lists and windows are. Output:
//                       results / items per token, one bit per work item that has a result, the statistics counters per
//                       document; the few items the entry cannot decide are queued.
//   k_eval_count_full   the queued items, one per thread, by the full fold (sp_eval_fold).
//   (prefix sums over the per token counts, k_eval_docs, prefix sums over the per document counts)
//   k_eval_write        one warp per run: walks the result bits of the run, 32 results at a time: full fold with the
//                       taken events recorded, result record and items written at their final place.
//   k_eval_docs         per document: counts, status, list of flagged documents (they run on the state machine of
//                       matcher_kernels.cuh and are copied into place by k_eval_place).
// No arena and no inter-thread dependency; atomics only for the per document counters and the result bits.
#ifndef SP_MATCHER_EVAL_KERNELS_CUH_INCLUDED
#define SP_MATCHER_EVAL_KERNELS_CUH_INCLUDED
#include "matcher_eval.cuh"
#include "matcher_kernels.cuh"

namespace sp {

enum {EvalWarpsPerBlock = 4, EvalThreads = EvalWarpsPerBlock * 32, EvalRun = 256, EvalLook = 64, EvalPerLane = EvalRun / 32, EvalSteps = 4, EvalRefill = 8};

#define SP_EVALQ_STATS_DONE (1ull << 63)	/* queued item whose counters are already added: only result, items and bit are missing */

struct EvalParams
{
	SpMatcherTables T;
	const sp_lexem* tokens;
	const uint32_t* origsegs;		// may be 0
	const uint64_t* tok_off;		// [ndocs+1]
	uint32_t ndocs;
	uint64_t ntokens;
	uint32_t* tok_slot; uint32_t* tok_work; uint32_t* tok_doc;	// [ntokens] k_eval_prepare: event slot, work items, document
	const uint64_t* tok_work_off;					// [ntokens+1] first work item of a token
	uint32_t* tok_nres; uint32_t* tok_nitems;			// [ntokens] count pass output
	const uint64_t* tok_res_off; const uint64_t* tok_item_off;	// [ntokens+1] prefix sums of the above
	const uint64_t* doc_res_off; const uint64_t* doc_item_off;	// [ndocs+1] final places of the documents
	uint32_t* resbits;			// one bit per work item: the rule has a result
	unsigned long long* queue; uint64_t queue_cap; unsigned long long* queue_count;
	sp_result* results; sp_item* items;
	sp_matcher_stats* stats;
	uint32_t* flags;			// [ndocs] 1 = run the document on the state machine
};

// document that owns token i: last d with tok_off[d] <= i
__device__ __forceinline__ uint32_t evalFindDoc( const uint64_t* __restrict__ tok_off, uint32_t ndocs, uint64_t i)
{
	uint32_t lo = 0, hi = ndocs;	// invariant: tok_off[lo] <= i < tok_off[hi]
	while (hi - lo > 1)
	{
		uint32_t mid = lo + (hi - lo) / 2;
		if (__ldg( &tok_off[ mid]) <= i) lo = mid; else hi = mid;
	}
	return lo;
}

__device__ __forceinline__ void evalAddStats( sp_matcher_stats* st, uint64_t installed, uint32_t altkey, uint64_t signals, uint64_t open)
{
	if (installed) atomicAdd( (unsigned long long*)&st->nof_programs_installed, (unsigned long long)installed);
	if (altkey) atomicAdd( (unsigned long long*)&st->nof_altkey_programs_installed, (unsigned long long)altkey);
	if (signals) atomicAdd( (unsigned long long*)&st->nof_signals_fired, (unsigned long long)signals);
	if (open) atomicAdd( (unsigned long long*)&st->nof_open_patterns, (unsigned long long)open);
}

// number of work items of a token with event slot `slotidx`
__device__ __forceinline__ uint32_t evalTokenWork( const SpMatcherTables& T, uint32_t slotidx)
{
	return slotidx == SP_EVAL_NOSLOT ? 0u : __ldg( &T.eval_slots[ slotidx].xcount);
}

static __global__ void __launch_bounds__(256) k_eval_prepare( EvalParams P)
{
	__shared__ uint32_t sdoc;
	const uint64_t b0 = (uint64_t)blockIdx.x * 256;
	if (threadIdx.x == 0) sdoc = evalFindDoc( P.tok_off, P.ndocs, b0);
	__syncthreads();
	const uint64_t i = b0 + threadIdx.x;
	if (i >= P.ntokens) return;
	uint32_t d = sdoc;
	while (__ldg( &P.tok_off[ d+1]) <= i) ++d;
	const uint64_t tb = __ldg( &P.tok_off[ d]);
	uint32_t slotidx = SP_EVAL_NOSLOT, work = 0;
	if (i >= tb)	// tokens in front of the first document belong to nobody
	{
		const sp_lexem tok = P.tokens[ i];
		const uint32_t prev = i > tb ? __ldg( &P.tokens[ i-1].ordpos) : 0;
		if (sp_eval_token_irregular( tok, P.origsegs ? P.origsegs[ i] : 0, i == tb, prev)) P.flags[ d] = 1;
		slotidx = sp_find_event_index( P.T, tok.id);
		work = evalTokenWork( P.T, slotidx);
	}
	P.tok_slot[ i] = slotidx; P.tok_work[ i] = work; P.tok_doc[ i] = d;
}

// Position inside the work items of a token: site and entry.
struct EvalCursor
{
	uint32_t c;			// 0 = token event, else follow event number c-1
	uint32_t k;			// entry inside the list
	uint32_t lbegin, lcount;	// program list of the site
	uint32_t fb, fc;		// static follow events of the token
};

// cursor at work item `rem` of a token with event slot `slotidx` (rem < work items of the token): one load of the flattened list
__device__ __forceinline__ void evalCursorSeek( const SpMatcherTables& T, uint32_t slotidx, uint32_t rem, EvalCursor& cur)
{
	const uint4 es = __ldg( reinterpret_cast<const uint4*>( T.eval_slots + slotidx));	// follow_begin, follow_count, xbegin, xcount
	const uint2 it = __ldg( reinterpret_cast<const uint2*>( T.eval_xlist + es.z + rem));
	cur.fb = es.x; cur.fc = es.y;
	cur.c = it.y; cur.k = 0; cur.lbegin = it.x; cur.lcount = 1;
}

__device__ __forceinline__ void evalQueuePush( const EvalParams& P, uint64_t item)
{
	const unsigned long long q = atomicAdd( P.queue_count, 1ull);
	if (q < P.queue_cap) P.queue[ q] = item;
}

static __global__ void __launch_bounds__(EvalThreads) k_eval_count_step( EvalParams P)
{
	__shared__ uint2 s_tok[ EvalWarpsPerBlock][ EvalRun + EvalLook];		// id, ordinal position
	__shared__ uint32_t s_work[ EvalWarpsPerBlock][ EvalRun + 1];			// first work item of a token, relative to the run
	__shared__ uint16_t s_doc[ EvalWarpsPerBlock][ EvalRun];			// document of a token, relative to the first document of the run
	__shared__ uint32_t s_nres[ EvalWarpsPerBlock][ EvalRun];
	__shared__ uint32_t s_nitems[ EvalWarpsPerBlock][ EvalRun];
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint2* stok = s_tok[ warp]; uint32_t* swork = s_work[ warp];
	uint16_t* sdoc = s_doc[ warp]; uint32_t* snres = s_nres[ warp]; uint32_t* snitems = s_nitems[ warp];

	const uint64_t run = (uint64_t)blockIdx.x * EvalWarpsPerBlock + warp;
	const uint64_t r0 = run * EvalRun;
	if (r0 >= P.ntokens) return;
	const uint32_t nrun = (uint32_t)min( (uint64_t)EvalRun, P.ntokens - r0);
	const uint32_t nstaged = (uint32_t)min( (uint64_t)(EvalRun + EvalLook), P.ntokens - r0);
	// ---- stage the run: every lane owns EvalPerLane consecutive tokens
	uint32_t dbase = 0;
	if (lane == 0) dbase = evalFindDoc( P.tok_off, P.ndocs, r0);
	dbase = __shfl_sync( SP_FULL, dbase, 0);
	for (uint32_t j=lane; j<nstaged; j+=32) stok[ j] = __ldg( reinterpret_cast<const uint2*>( P.tokens + r0 + j));
	uint32_t mywork = 0;
	{
		uint32_t d = dbase;
		for (uint32_t q=0; q<EvalPerLane; ++q)
		{
			const uint32_t j = lane * EvalPerLane + q;
			uint32_t w = 0;
			if (j < nrun)
			{
				const uint64_t i = r0 + j;
				while (__ldg( &P.tok_off[ d+1]) <= i) ++d;
				sdoc[ j] = (uint16_t)min( d - dbase, 0xFFFFu);
				w = P.tok_work[ i];
			}
			snres[ j] = 0; snitems[ j] = 0;
			swork[ j] = w;		// turned into the prefix sum below
			mywork += w;
		}
	}
	uint32_t incl = mywork;
	#pragma unroll
	for (int ofs=1; ofs<32; ofs<<=1)
	{
		uint32_t up = __shfl_up_sync( SP_FULL, incl, ofs);
		if (lane >= (unsigned int)ofs) incl += up;
	}
	const uint32_t W = __shfl_sync( SP_FULL, incl, 31);
	{
		uint32_t run_sum = incl - mywork;
		for (uint32_t q=0; q<EvalPerLane; ++q)
		{
			const uint32_t j = lane * EvalPerLane + q;
			const uint32_t w = swork[ j];
			swork[ j] = run_sum;
			run_sum += w;
		}
		if (lane == 31) swork[ EvalRun] = W;
	}
	__syncwarp();
	if (W == 0)
	{
		for (uint32_t j=lane; j<nrun; j+=32) {P.tok_nres[ r0 + j] = 0; P.tok_nitems[ r0 + j] = 0;}
		return;
	}
	const uint64_t wbase = __ldg( &P.tok_work_off[ r0]);	// number of the run's first work item

	// ---- every lane follows the rules of its block of work items
	const uint32_t B = (W + 31) / 32;
	uint32_t w = min( lane * B, W);
	const uint32_t wend = min( w + B, W);
	uint32_t t = 0, tslot = SP_EVAL_NOSLOT;	// token of the run the cursor is in, its event slot
	EvalCursor cur; cur.c = cur.k = cur.lbegin = cur.lcount = cur.fb = cur.fc = 0;
	if (w < wend)
	{
		uint32_t lo = 0, hi = EvalRun;		// last token with swork[t] <= w
		while (hi - lo > 1) {uint32_t mid = (lo + hi) / 2; if (swork[ mid] <= w) lo = mid; else hi = mid;}
		t = lo;
		tslot = P.tok_slot[ r0 + t];
		evalCursorSeek( P.T, tslot, w - swork[ t], cur);
	}
	// counters of the document the lane is in
	uint32_t accdoc = (w < wend) ? dbase + sdoc[ t] : 0xFFFFFFFFu, acc_installed = 0;
	uint64_t acc_signals = 0, acc_open = 0;
	// the rule the lane follows
	bool have = false;
	SpEvalEntry en; SpEvalLite L;
	uint32_t expiry = 0, wpos = 0, wte = 0;
	en.ev[0] = en.ev[1] = en.ev[2] = en.ev[3] = 0xFFFFFFFFu; en.codes = en.value0 = en.meta = en.meta2 = 0;
	L.value = L.count = L.endpos = L.flags = L.nitems = L.signals = L.alive_tokens = 0;

	// The lanes step their rules in lock step (EvalSteps window tokens per round, one uniform instruction stream); ended
	// rules are accounted and replaced at the begin of a round. A lane waits at most EvalSteps-1 steps for its next rule.
	auto advance = [&]()
	{
		++w;
		if (w < wend)
		{
			while (swork[ t+1] <= w) ++t;		// w < wend: there is a token that owns item w
			evalCursorSeek( P.T, P.tok_slot[ r0 + t], w - swork[ t], cur);
		}
	};
	auto account = [&]()
	{
		if (L.flags & SP_LITE_FULL) {evalQueuePush( P, wbase + w); return;}
		acc_installed += 1; acc_signals += L.signals;
		acc_open += (uint64_t)L.alive_tokens * SP_EVAL_META_NTRIG( en.meta);
		if (L.flags & SP_LITE_RESULT)
		{
			uint32_t extra = 0;
			if ((en.meta & SP_EVAL_META_ITEMSFULL)
			    && !sp_eval_lite_follow_items( P.T, P.tokens, r0 + t, __ldg( &P.tok_off[ accdoc + 1]), en, P.T.eval_follows[ cur.fb + cur.c - 1], extra))
			{
				evalQueuePush( P, (wbase + w) | SP_EVALQ_STATS_DONE);
			}
			else
			{
				atomicAdd( &snres[ t], 1u);
				if (L.flags & SP_LITE_DREFEMIT) atomicAdd( &snitems[ t], L.nitems + extra);
				const uint64_t bit = wbase + w;
				atomicOr( P.resbits + (bit >> 5), 1u << (bit & 31));
			}
		}
	};
	while (__any_sync( SP_FULL, have || w < wend))
	{
		// rules are replaced when enough lanes wait for one (the replacement code runs for the waiting lanes only)
		const unsigned int waiting = __ballot_sync( SP_FULL, !(L.flags & SP_LITE_ALIVE) && (have || w < wend));
		const unsigned int running = __ballot_sync( SP_FULL, (L.flags & SP_LITE_ALIVE) != 0);
		if (!(L.flags & SP_LITE_ALIVE) && (__popc( waiting) >= EvalRefill || !running))
		{
			if (have) {account(); have = false; advance();}
			while (w < wend)
			{
				// take the item the cursor stands on
				const uint4* ep = reinterpret_cast<const uint4*>( P.T.eval_entries + cur.lbegin + cur.k);
				const uint4 a = __ldg( ep), b = __ldg( ep + 1);
				en.ev[0] = a.x; en.ev[1] = a.y; en.ev[2] = a.z; en.ev[3] = a.w; en.codes = b.x; en.value0 = b.y; en.meta = b.z; en.meta2 = b.w;
				const uint32_t d = dbase + sdoc[ t];
				if (d != accdoc)
				{
					evalAddStats( P.stats + accdoc, acc_installed, 0, acc_signals, acc_open);
					accdoc = d; acc_installed = 0; acc_signals = 0; acc_open = 0;
				}
				if (en.meta & SP_EVAL_META_GENERIC) {evalQueuePush( P, wbase + w); advance(); continue;}
				const uint32_t spos = stok[ t].y;
				sp_eval_lite_init( en, spos, L);
				if (!(L.flags & SP_LITE_ALIVE)) {account(); advance(); continue;}	// ended by its installation
				expiry = spos + SP_EVAL_META_RANGE( en.meta);
				wpos = t + 1; wte = (uint32_t)min( __ldg( &P.tok_off[ d + 1]) - r0, (uint64_t)0xFFFFFFFFu);
				have = true;
				break;
			}
		}
		#pragma unroll
		for (int step=0; step<EvalSteps; ++step)
		{
			if (L.flags & SP_LITE_ALIVE)
			{
				if (wpos < wte)
				{
					uint2 tk;
					if (wpos < EvalRun + EvalLook) tk = stok[ wpos];
					else tk = __ldg( reinterpret_cast<const uint2*>( P.tokens + r0 + wpos));
					sp_eval_lite_step( en, expiry, tk.x, tk.y, L);
					++wpos;
				}
				else L.flags &= ~SP_LITE_ALIVE;
			}
		}
	}
	if (accdoc != 0xFFFFFFFFu) evalAddStats( P.stats + accdoc, acc_installed, 0, acc_signals, acc_open);
	__syncwarp();
	for (uint32_t j=lane; j<nrun; j+=32) {P.tok_nres[ r0 + j] = snres[ j]; P.tok_nitems[ r0 + j] = snitems[ j];}
}

// ---- count pass as a hash join (default) ------------------------------------------------------------------------------
// A rule instance that never gets a signal after its installation (the fate of more than nine rules out of ten) is known as
// such as soon as none of the events of its live triggers occurs among the tokens of its window. All rules installed by
// one token share that window, so the warp builds, once per key token, a small hash table in shared memory: token id ->
// distance (in ordinal positions) of its first occurrence behind the key token, for the tokens up to EvalJoinRange
// positions away, plus the number of window tokens per distance and the list of the distinct ids.
// The quiet rules are never touched: what they add to the counters is summed up per event slot at compile time
// (matcher_tables.h: SpEvalAgg) and needs the window token counts only. The rules that CAN get a signal are found from the
// window's side: every distinct id X of the window is looked up with the key token's event slot S in the pair index
// (SpEvalPair: the entries of S that have a live trigger on X). Those entries are loaded and probed against the table
// (an entry reached from several ids is taken by the first of its events that is in the window); the ones that get a
// signal inside their range are queued in shared memory and followed token by token, 32 at a time (sp_eval_lite_step),
// and their outcome replaces their quiet share of the counters. Entries outside the aggregate (generic ones, a result at
// the installation, a range beyond the table) and the tokens whose window does not fit the table are handled one by one.
enum {EvalTile = 8, EvalTileLook = 56, EvalJoinSlots = 64, EvalJoinCap = 64, EvalJoinRange = 15, EvalStepQueue = 64, EvalJoinIds = 32};
#define SP_JOIN_EMPTY 0xFFFFFFFFFFFFFFFFull
#define SP_JOINQ_DELTA 0x80000000u	/* queued work item of the aggregate: its outcome is accounted as the difference to its quiet share */

struct EvalLaneAcc		// statistics of the document a lane is in (sums modulo 2^64: quiet shares are taken back by subtraction)
{
	uint32_t doc;
	uint64_t installed, signals, open;
};

// (the rare and the bulky parts of the join kernel are kept out of line: the kernel is bound by instruction fetch, its
// warps are in different phases most of the time)
__device__ __noinline__ void evalFlushStats( sp_matcher_stats* st, uint64_t installed, uint64_t signals, uint64_t open)
{
	evalAddStats( st, installed, 0, signals, open);
}

__device__ __forceinline__ void evalAccDoc( const EvalParams& P, EvalLaneAcc& acc, uint32_t d)
{
	if (d == acc.doc) return;
	if (acc.doc != 0xFFFFFFFFu) evalFlushStats( P.stats + acc.doc, acc.installed, acc.signals, acc.open);
	acc.doc = d; acc.installed = 0; acc.signals = 0; acc.open = 0;
}

// result count / item count / result bit of an ended rule that has a result, or a place in the queue of the full fold
__device__ __noinline__ void evalAccountResult( const EvalParams& P, const SpEvalEntry& en, uint32_t nitems, bool drefemit, uint64_t item, uint64_t tok, uint64_t te,
						     const SpEvalFollow* F, uint32_t* snres, uint32_t* snitems)
{
	uint32_t extra = 0;
	if ((en.meta & SP_EVAL_META_ITEMSFULL) && !sp_eval_lite_follow_items( P.T, P.tokens, tok, te, en, *F, extra))
	{
		evalQueuePush( P, item | SP_EVALQ_STATS_DONE);
		return;
	}
	atomicAdd( snres, 1u);
	if (drefemit) atomicAdd( snitems, nitems + extra);
	atomicOr( P.resbits + (item >> 5), 1u << (item & 31));
}

// A batch of rules, one per lane, followed token by token in lock step until the last one has ended (one copy of the
// state machine of sp_eval_lite_step in the kernel).
__device__ __noinline__ void evalStepBatch( const SpEvalEntry* enp, SpEvalLite* Lp, uint32_t expiry, uint64_t k, uint64_t te, uint64_t r0, const uint2* stok, const sp_lexem* tokens)
{
	const SpEvalEntry en = *enp;
	SpEvalLite L = *Lp;
	// token numbers relative to the first token of the tile: 32 bit arithmetic in the loop
	uint32_t kr = (uint32_t)(k - r0);
	const uint32_t ter = (uint32_t)min( te - r0, (uint64_t)0xFFFFFFFFu);
	const uint2* tile_tokens = reinterpret_cast<const uint2*>( tokens + r0);	// (id, ordpos) are the first half of a 16 byte record
	while (__any_sync( SP_FULL, (L.flags & SP_LITE_ALIVE) != 0))
	{
		#pragma unroll 1
		for (int step=0; step<4; ++step)
		{
			if (L.flags & SP_LITE_ALIVE)
			{
				if (kr < ter)
				{
					const uint2 tk = (kr < EvalTile + EvalTileLook) ? stok[ kr] : __ldg( tile_tokens + 2 * (uint64_t)kr);
					sp_eval_lite_step( en, expiry, tk.x, tk.y, L);
					++kr;
				}
				else L.flags &= ~SP_LITE_ALIVE;
			}
		}
	}
	*Lp = L;
}

// what an ended rule leaves behind: counters, result count / item count / result bit, or a place in the queue of the full fold
__device__ __forceinline__ void evalAccount( const EvalParams& P, const SpEvalEntry& en, const SpEvalLite& L, uint64_t item, uint64_t tok, uint64_t te,
						const SpEvalFollow* F, uint32_t* snres, uint32_t* snitems, EvalLaneAcc& acc)
{
	if (L.flags & SP_LITE_FULL) {evalQueuePush( P, item); return;}
	acc.installed += 1; acc.signals += L.signals;
	acc.open += (uint64_t)L.alive_tokens * SP_EVAL_META_NTRIG( en.meta);
	if (L.flags & SP_LITE_RESULT) evalAccountResult( P, en, L.nitems, (L.flags & SP_LITE_DREFEMIT) != 0, item, tok, te, F, snres, snitems);
}

// the same for an entry of the aggregate (alive behind its installation, no result there) whose quiet share - one program
// installed, the signals of the installation, its live triggers over the `quiet_alive` tokens of its window - is already
// in the counters
__device__ __forceinline__ void evalAccountDelta( const EvalParams& P, const SpEvalEntry& en, const SpEvalLite& L, uint32_t quiet_alive, uint64_t item, uint64_t tok, uint64_t te,
						     const SpEvalFollow* F, uint32_t* snres, uint32_t* snitems, EvalLaneAcc& acc)
{
	const uint64_t ntrig = SP_EVAL_META_NTRIG( en.meta), quiet_signals = SP_EVAL_META_NKEY( en.meta);
	if (L.flags & SP_LITE_FULL)
	{
		// the full fold adds all counters of the rule
		evalQueuePush( P, item);
		acc.installed -= 1; acc.signals -= quiet_signals; acc.open -= ntrig * quiet_alive;
		return;
	}
	acc.signals += (uint64_t)L.signals - quiet_signals;
	acc.open += ntrig * L.alive_tokens - ntrig * quiet_alive;
	if (L.flags & SP_LITE_RESULT) evalAccountResult( P, en, L.nitems, (L.flags & SP_LITE_DREFEMIT) != 0, item, tok, te, F, snres, snitems);
}

__device__ __forceinline__ uint32_t evalJoinHash( uint32_t id) {return (id * 0x9E3779B1u) >> 26;}	// 6 bits: EvalJoinSlots

// number of leading values of the ascending sequence held by the lanes (inclusive prefix sums) that are <= x
__device__ __forceinline__ uint32_t evalLanesUpTo( uint32_t incl, uint32_t x)
{
	uint32_t lo = 0;
	#pragma unroll
	for (int step=16; step>0; step>>=1)
	{
		const uint32_t v = __shfl_sync( SP_FULL, incl, (lo + step - 1) & 31);
		if (v <= x) lo += step;
	}
	return lo;
}

template <int BLOCKS_PER_SM>
static __global__ void __launch_bounds__(EvalThreads, BLOCKS_PER_SM) k_eval_count_join( const __grid_constant__ EvalParams P)
{
	__shared__ uint32_t s_key[ EvalWarpsPerBlock][ EvalTile][ EvalJoinSlots];		// token id, 0xFFFFFFFF = empty
	__shared__ uint8_t s_dist[ EvalWarpsPerBlock][ EvalTile][ EvalJoinSlots];		// distance of its first occurrence
	__shared__ uint8_t s_cnt[ EvalWarpsPerBlock][ EvalTile][ EvalJoinRange + 1];		// window tokens up to a distance
	__shared__ uint8_t s_wslot[ EvalWarpsPerBlock][ EvalTile][ EvalJoinIds];		// table slots of the distinct ids of a window
	__shared__ uint32_t s_wn[ EvalWarpsPerBlock][ EvalTile];				// their number (0: no window lookups for this token)
	__shared__ uint32_t s_slot[ EvalWarpsPerBlock][ EvalTile];				// event slot of a token
	__shared__ uint32_t s_queue[ EvalWarpsPerBlock][ EvalStepQueue];			// items of the tile to follow token by token
	__shared__ uint2 s_tok[ EvalWarpsPerBlock][ EvalTile + EvalTileLook];
	__shared__ uint32_t s_nres[ EvalWarpsPerBlock][ EvalTile], s_nitems[ EvalWarpsPerBlock][ EvalTile];
	__shared__ uint4 s_info[ EvalWarpsPerBlock][ EvalTile][ 2];	// per token: (first item, work list, program list, direct items), (follow begin, document, position, follow count)
	__shared__ unsigned long long s_te[ EvalWarpsPerBlock][ EvalTile];
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint4 (*sinfo)[ 2] = s_info[ warp];
	unsigned long long* ste = s_te[ warp];
	uint32_t (*skey)[ EvalJoinSlots] = s_key[ warp];
	uint8_t (*sdist)[ EvalJoinSlots] = s_dist[ warp];
	uint8_t (*scnt)[ EvalJoinRange + 1] = s_cnt[ warp];
	uint8_t (*swslot)[ EvalJoinIds] = s_wslot[ warp];
	uint32_t* swn = s_wn[ warp]; uint32_t* sslot = s_slot[ warp];
	uint32_t* squeue = s_queue[ warp]; uint32_t* snres = s_nres[ warp]; uint32_t* snitems = s_nitems[ warp];
	uint2* stok = s_tok[ warp];

	const uint64_t tile = (uint64_t)blockIdx.x * EvalWarpsPerBlock + warp;
	const uint64_t r0 = tile * EvalTile;
	if (r0 >= P.ntokens) return;
	const uint32_t ntile = (uint32_t)min( (uint64_t)EvalTile, P.ntokens - r0);
	// ---- the tokens of the tile: lane j < EvalTile holds token j
	uint32_t my_work = 0, my_slot = SP_EVAL_NOSLOT, my_doc = 0, my_pos = 0;
	uint32_t my_lbegin = 0, my_pbegin = 0, my_lcount = 0, my_fb = 0, my_fc = 0;	// work item list, program list of the token event, static follow events
	uint64_t my_te = 0;
	if (lane < ntile)
	{
		my_work = P.tok_work[ r0 + lane];
		my_slot = P.tok_slot[ r0 + lane];
		if (my_work)
		{
			my_lbegin = __ldg( &P.T.eval_slots[ my_slot].xbegin);
			my_pbegin = __ldg( &P.T.event_slots[ my_slot].prog_begin); my_lcount = __ldg( &P.T.event_slots[ my_slot].prog_count);
			my_fb = __ldg( &P.T.eval_slots[ my_slot].follow_begin); my_fc = __ldg( &P.T.eval_slots[ my_slot].follow_count);
		}
	}
	const unsigned int workmask = __ballot_sync( SP_FULL, my_work != 0);
	if (lane < EvalTile) {snres[ lane] = 0; snitems[ lane] = 0; swn[ lane] = 0; sslot[ lane] = my_slot;}
	if (!workmask)
	{
		if (lane < ntile) {P.tok_nres[ r0 + lane] = 0; P.tok_nitems[ r0 + lane] = 0;}
		return;
	}
	if (lane < ntile)
	{
		my_doc = P.tok_doc[ r0 + lane]; my_te = __ldg( &P.tok_off[ my_doc + 1]);
	}
	// the tokens of the tile and behind it, (id, ordinal position): the windows of the rules that are followed step by step
	for (uint32_t j=lane; j<EvalTile+EvalTileLook; j+=32)
	{
		stok[ j] = (r0 + j < P.ntokens) ? __ldg( reinterpret_cast<const uint2*>( P.tokens + r0 + j)) : make_uint2( 0, 0);
	}
	__syncwarp();
	if (lane < ntile) my_pos = stok[ lane].y;
	// prefix sums of the work items over the tile (lanes 0..EvalTile)
	uint32_t pre = my_work;
	#pragma unroll
	for (int ofs=1; ofs<EvalTile; ofs<<=1)
	{
		uint32_t up = __shfl_up_sync( SP_FULL, pre, ofs);
		if (lane >= (unsigned int)ofs) pre += up;
	}
	pre -= my_work;		// exclusive
	const uint64_t wbase = __ldg( &P.tok_work_off[ r0]);
	// ---- hash tables of the windows of the tokens that install something
	const uint32_t joinrange = P.T.eval_join_range;		// <= EvalJoinRange
	unsigned int bigwindow = 0;	// tokens whose window does not fit: all their rules are followed token by token
	for (unsigned int m=workmask; m; m &= m - 1)
	{
		const int t = __ffs( m) - 1;
		const uint64_t i = r0 + t, te = __shfl_sync( SP_FULL, my_te, t);
		const uint32_t spos = __shfl_sync( SP_FULL, my_pos, t);
		for (uint32_t h=lane; h<EvalJoinSlots; h+=32) skey[ t][ h] = 0xFFFFFFFFu;
		uint32_t cnt = 0;	// window tokens of the rounds done
		uint32_t wn = 0;	// distinct ids of the rounds done
		if (lane <= EvalJoinRange) scnt[ t][ lane] = 0;
		__syncwarp();
		bool more = true;
		for (uint32_t round=0; more; ++round)
		{
			const uint64_t k = i + 1 + round * 32 + lane;
			uint2 tk = make_uint2( 0, 0);
			bool in = false;
			if (k < te)
			{
				tk = (k - r0 < EvalTile + EvalTileLook) ? stok[ k - r0] : __ldg( reinterpret_cast<const uint2*>( P.tokens + k));
				in = tk.y <= spos + joinrange && tk.y >= spos;
			}
			const unsigned int inmask = __ballot_sync( SP_FULL, in);
			// positions ascend with the lanes and with the rounds: the first occurrence of an id is its lowest lane in the
			// first round that has it; that lane alone inserts the id, an id that is already there keeps its distance
			const unsigned int same = __match_any_sync( SP_FULL, in ? tk.x : 0xFFFFFFFFu - lane);
			bool inserted = false; uint32_t hslot = 0;
			if (in && (unsigned int)(__ffs( same) - 1) == lane)
			{
				for (uint32_t h = evalJoinHash( tk.x);; h = (h + 1) & (EvalJoinSlots - 1))
				{
					const uint32_t old = atomicCAS( &skey[ t][ h], 0xFFFFFFFFu, tk.x);
					if (old == 0xFFFFFFFFu) {sdist[ t][ h] = (uint8_t)(tk.y - spos); inserted = true; hslot = h; break;}
					if (old == tk.x) break;
				}
			}
			{
				const unsigned int insmask = __ballot_sync( SP_FULL, inserted);
				const uint32_t p = wn + __popc( insmask & laneLt( lane));
				if (inserted && p < EvalJoinIds) swslot[ t][ p] = (uint8_t)hslot;
				wn += __popc( insmask);
			}
			__syncwarp();
			// window tokens up to a distance: the distances ascend with the lanes, so lane L knows the count for every distance
			// from its own up to (not including) its neighbour's: base + L + 1
			{
				const uint32_t dist = tk.y - spos;
				const uint32_t nextdist = __shfl_down_sync( SP_FULL, in ? dist : 0xFFFFu, 1);
				const uint32_t upto = (lane == 31 || !((inmask >> (lane + 1)) & 1u)) ? joinrange + 1 : nextdist;
				if (in) for (uint32_t r=dist; r<upto; ++r) scnt[ t][ r] = (uint8_t)min( cnt + lane + 1, 255u);
				cnt += __popc( inmask);
			}
			more = inmask == SP_FULL;		// positions ascend: the window ends inside this round otherwise
			if (more && (round + 1) * 32 >= EvalJoinCap) {bigwindow |= 1u << t; more = false;}
		}
		if (wn > EvalJoinIds) bigwindow |= 1u << t;
		if (lane == 0) swn[ t] = ((bigwindow >> t) & 1u) ? 0u : wn;
	}
	__syncwarp();

	EvalLaneAcc acc; acc.doc = 0xFFFFFFFFu; acc.installed = 0; acc.signals = 0; acc.open = 0;
	// ---- the quiet share of the aggregated entries, and the entries to handle one by one (all of them for a window that
	// does not fit the table)
	uint32_t my_ic = 0, my_ib = 0;
	if (lane < ntile && my_work)
	{
		const SpEvalAgg* ag = P.T.eval_agg + my_slot;
		const uint4 head = __ldg( reinterpret_cast<const uint4*>( ag));	// n, signals, ind_begin, ind_count
		if ((bigwindow >> lane) & 1u) my_ic = my_work;
		else
		{
			my_ic = head.w; my_ib = head.z;
			uint64_t open = 0;
			#pragma unroll
			for (int q=0; q<SP_EVAL_AGG_RANGES/4; ++q)
			{
				const uint4 o = __ldg( reinterpret_cast<const uint4*>( ag->open) + q);
				open += (uint64_t)o.x * scnt[ lane][ 4*q] + (uint64_t)o.y * scnt[ lane][ 4*q+1] + (uint64_t)o.z * scnt[ lane][ 4*q+2] + (uint64_t)o.w * scnt[ lane][ 4*q+3];
			}
			acc.doc = my_doc; acc.installed = head.x; acc.signals = head.y; acc.open = open;
		}
	}
	uint32_t nqueue = 0;		// uniform
	// first items of the tokens, in registers of every lane
	uint32_t pfx[ EvalTile];
	#pragma unroll
	for (int j=0; j<EvalTile; ++j) pfx[ j] = __shfl_sync( SP_FULL, pre, j);
	if (lane < EvalTile)
	{
		sinfo[ lane][ 0] = make_uint4( pre, my_lbegin, my_pbegin, my_lcount);
		sinfo[ lane][ 1] = make_uint4( my_fb, my_doc, my_pos, my_fc);
		ste[ lane] = my_te;
	}
	__syncwarp();
	auto decode = [&]( uint32_t wl, uint32_t& t, EvalCursor& cur)
	{
		// the token of item wl: the last one whose first item is not behind wl (tokens without items share their first item
		// with the token behind them, and wl is an item of the tile)
		uint32_t n = 0;
		#pragma unroll
		for (int j=1; j<EvalTile; ++j) n += (wl >= pfx[ j]) ? 1u : 0u;
		t = n;
		const uint4 a = sinfo[ t][ 0], b = sinfo[ t][ 1];
		const uint32_t rem = wl - a.x;
		cur.fb = b.x; cur.fc = b.w;
		cur.c = 0; cur.k = 0; cur.lbegin = a.z + rem; cur.lcount = 1;
		if (rem >= a.w)	// an item of a follow event of the token
		{
			const uint2 it = __ldg( reinterpret_cast<const uint2*>( P.T.eval_xlist + a.y + rem));
			cur.c = it.y; cur.lbegin = it.x;
		}
	};
	auto load_entry = [&]( const EvalCursor& cur, SpEvalEntry& en)
	{
		const uint4* ep = reinterpret_cast<const uint4*>( P.T.eval_entries + cur.lbegin + cur.k);
		const uint4 a = __ldg( ep), b = __ldg( ep + 1);
		en.ev[0] = a.x; en.ev[1] = a.y; en.ev[2] = a.z; en.ev[3] = a.w; en.codes = b.x; en.value0 = b.y; en.meta = b.z; en.meta2 = b.w;
	};
	auto follow_steps = [&]( uint32_t nbatch)
	{
		// the first nbatch queued items, one per lane: followed token by token in lock step
		const bool valid = lane < nbatch;
		const uint32_t qv = valid ? squeue[ lane] : 0;
		const uint32_t wl = qv & ~SP_JOINQ_DELTA;
		uint32_t t; EvalCursor cur;
		decode( wl, t, cur);
		const uint32_t d = sinfo[ t][ 1].y, spos = sinfo[ t][ 1].z;
		const uint64_t te = ste[ t];
		SpEvalEntry en; SpEvalLite L;
		en.ev[0] = en.ev[1] = en.ev[2] = en.ev[3] = 0xFFFFFFFFu; en.codes = en.value0 = en.meta = en.meta2 = 0;
		L.value = L.count = L.endpos = L.flags = L.nitems = L.signals = L.alive_tokens = 0;
		uint32_t expiry = 0;
		uint64_t k = r0 + t + 1;
		if (valid)
		{
			load_entry( cur, en);
			sp_eval_lite_init( en, spos, L);
			expiry = spos + SP_EVAL_META_RANGE( en.meta);
		}
		evalStepBatch( &en, &L, expiry, k, te, r0, stok, P.tokens);
		if (valid)
		{
			evalAccDoc( P, acc, d);
			const SpEvalFollow* F = cur.c ? P.T.eval_follows + cur.fb + cur.c - 1 : 0;
			if (qv & SP_JOINQ_DELTA) evalAccountDelta( P, en, L, scnt[ t][ SP_EVAL_META_RANGE( en.meta)], wbase + wl, r0 + t, te, F, snres + t, snitems + t, acc);
			else evalAccount( P, en, L, wbase + wl, r0 + t, te, F, snres + t, snitems + t, acc);
		}
		__syncwarp();
		// the rest of the queue moves to the front
		const uint32_t rest = nqueue - nbatch;
		uint32_t mv[ EvalStepQueue / 32];
		#pragma unroll
		for (int q=0; q<EvalStepQueue/32; ++q) mv[ q] = (q * 32 + lane < rest) ? squeue[ nbatch + q * 32 + lane] : 0;
		__syncwarp();
		#pragma unroll
		for (int q=0; q<EvalStepQueue/32; ++q) if (q * 32 + lane < rest) squeue[ q * 32 + lane] = mv[ q];
		nqueue = rest;
		__syncwarp();
	};
	auto enqueue = [&]( bool tostep, uint32_t value)
	{
		// at most 31 items wait in the queue here, so 32 more fit
		const unsigned int sm = __ballot_sync( SP_FULL, tostep);
		if (tostep) squeue[ nqueue + __popc( sm & laneLt( lane))] = value;
		nqueue += __popc( sm);
		__syncwarp();
		while (nqueue >= 32) follow_steps( 32);
	};
	// ---- entries handled one by one, 32 per round
	{
		uint32_t ipre = my_ic;
		#pragma unroll
		for (int ofs=1; ofs<EvalTile; ofs<<=1)
		{
			uint32_t up = __shfl_up_sync( SP_FULL, ipre, ofs);
			if (lane >= (unsigned int)ofs) ipre += up;
		}
		const uint32_t It = __shfl_sync( SP_FULL, ipre, EvalTile - 1);
		ipre -= my_ic;		// exclusive
		for (uint32_t base=0; base<It; base+=32)
		{
			const uint32_t il = base + lane;
			const bool valid = il < It;
			// the token of the entry: the last one whose first entry is not behind il
			uint32_t t = 0;
			#pragma unroll
			for (int j=1; j<EvalTile; ++j) t += (il >= __shfl_sync( SP_FULL, ipre, j)) ? 1u : 0u;
			const uint32_t tfirst = __shfl_sync( SP_FULL, ipre, t), tib = __shfl_sync( SP_FULL, my_ib, t);
			bool tostep = false, generic = false;
			uint32_t wl = 0;
			if (valid)
			{
				const bool big = (bigwindow >> t) & 1u;
				const uint32_t rem = big ? il - tfirst : __ldg( P.T.eval_indlist + tib + (il - tfirst));
				wl = sinfo[ t][ 0].x + rem;
				uint32_t t2; EvalCursor cur;
				decode( wl, t2, cur);
				SpEvalEntry en;
				load_entry( cur, en);
				if (en.meta & SP_EVAL_META_GENERIC) generic = true;
				else
				{
					const uint32_t range = SP_EVAL_META_RANGE( en.meta);
					bool quiet = true;
					if (!(en.meta & SP_EVAL_META_DEAD))
					{
						if (range > joinrange || big) quiet = false;
						else
						{
							bool hit = false;
							#pragma unroll
							for (int e=0; e<SP_EVAL_NEV; ++e)
							{
								const uint32_t ev = en.ev[ e];
								uint32_t h = evalJoinHash( ev);
								uint32_t key = skey[ t][ h];
								while (key != 0xFFFFFFFFu && key != ev) {h = (h + 1) & (EvalJoinSlots - 1); key = skey[ t][ h];}
								hit |= (key == ev) & (ev != 0xFFFFFFFFu) & (sdist[ t][ h] <= range);
							}
							quiet = !hit;
						}
					}
					if (quiet)
					{
						SpEvalLite L;
						sp_eval_lite_init( en, sinfo[ t][ 1].z, L);
						if (L.flags & SP_LITE_ALIVE) L.alive_tokens = scnt[ t][ range];
						evalAccDoc( P, acc, sinfo[ t][ 1].y);
						evalAccount( P, en, L, wbase + wl, r0 + t, ste[ t], cur.c ? P.T.eval_follows + cur.fb + cur.c - 1 : 0, snres + t, snitems + t, acc);
					}
					else tostep = true;
				}
			}
			// the entries for the full fold: one addition to the queue counter per round
			{
				const unsigned int gm = __ballot_sync( SP_FULL, generic);
				if (gm)
				{
					const int leader = __ffs( gm) - 1;
					unsigned long long qb = 0;
					if ((int)lane == leader) qb = atomicAdd( P.queue_count, (unsigned long long)__popc( gm));
					qb = __shfl_sync( SP_FULL, qb, leader);
					const unsigned long long q = qb + __popc( gm & laneLt( lane));
					if (generic && q < P.queue_cap) P.queue[ q] = wbase + wl;
				}
			}
			enqueue( tostep, wl);
		}
	}
	// ---- the aggregated entries that can get a signal: pair index lookups for the distinct ids of the windows, 32 per round
	{
		uint32_t my_wn = lane < EvalTile ? swn[ lane] : 0;
		uint32_t dpre = my_wn;
		#pragma unroll
		for (int ofs=1; ofs<EvalTile; ofs<<=1)
		{
			uint32_t up = __shfl_up_sync( SP_FULL, dpre, ofs);
			if (lane >= (unsigned int)ofs) dpre += up;
		}
		const uint32_t Dt = __shfl_sync( SP_FULL, dpre, EvalTile - 1);
		dpre -= my_wn;		// exclusive
		const uint32_t pairmask = P.T.eval_pair_mask;
		for (uint32_t base=0; base<Dt; base+=32)
		{
			const uint32_t il = base + lane;
			uint32_t t = 0;
			#pragma unroll
			for (int j=1; j<EvalTile; ++j) t += (il >= __shfl_sync( SP_FULL, dpre, j)) ? 1u : 0u;
			const uint32_t tfirst = __shfl_sync( SP_FULL, dpre, t);
			uint32_t cnt = 0, begin = 0, X = 0xFFFFFFFFu;
			if (il < Dt)
			{
				X = skey[ t][ swslot[ t][ il - tfirst]];
				const uint32_t slot = sslot[ t];
				for (uint32_t i = sp_eval_pair_hash( slot, X) & pairmask;; i = (i + 1) & pairmask)
				{
					const uint4 p = __ldg( reinterpret_cast<const uint4*>( P.T.eval_pairs + i));
					if (p.x == 0xFFFFFFFFu) break;
					if (p.x == slot && p.y == X) {begin = p.z; cnt = p.w; break;}
				}
			}
			// the candidates of the round, 32 at a time: lane c takes candidate number c of the concatenated lists
			uint32_t incl = cnt;
			#pragma unroll
			for (int ofs=1; ofs<32; ofs<<=1)
			{
				uint32_t up = __shfl_up_sync( SP_FULL, incl, ofs);
				if (lane >= (unsigned int)ofs) incl += up;
			}
			const uint32_t Ct = __shfl_sync( SP_FULL, incl, 31);
			for (uint32_t cb=0; cb<Ct; cb+=32)
			{
				const uint32_t ci = cb + lane;
				const uint32_t src = evalLanesUpTo( incl, ci);
				const uint32_t ct = __shfl_sync( SP_FULL, t, src), cbegin = __shfl_sync( SP_FULL, begin, src);
				const uint32_t cfirst = __shfl_sync( SP_FULL, incl - cnt, src), cX = __shfl_sync( SP_FULL, X, src);
				bool tostep = false;
				uint32_t wl = 0;
				if (ci < Ct)
				{
					wl = sinfo[ ct][ 0].x + __ldg( P.T.eval_pairlist + cbegin + (ci - cfirst));
					uint32_t t2; EvalCursor cur;
					decode( wl, t2, cur);
					SpEvalEntry en;
					load_entry( cur, en);
					const uint32_t range = SP_EVAL_META_RANGE( en.meta);
					// the first event of the entry that is in the window owns the entry (it is reached from every one of them)
					bool hit = false, owned = false; uint32_t owner = 0xFFFFFFFFu;
					#pragma unroll
					for (int e=0; e<SP_EVAL_NEV; ++e)
					{
						const uint32_t ev = en.ev[ e];
						uint32_t h = evalJoinHash( ev);
						uint32_t key = skey[ ct][ h];
						while (key != 0xFFFFFFFFu && key != ev) {h = (h + 1) & (EvalJoinSlots - 1); key = skey[ ct][ h];}
						const bool present = (key == ev) & (ev != 0xFFFFFFFFu);
						if (present && !owned) {owned = true; owner = ev;}
						hit |= present & (sdist[ ct][ h] <= range);
					}
					tostep = hit && owner == cX;
				}
				enqueue( tostep, wl | SP_JOINQ_DELTA);
			}
		}
	}
	while (nqueue) follow_steps( min( nqueue, 32u));
	// the counters of the lanes: one set of atomics per warp when all lanes are in the same document (the usual case)
	{
		const uint32_t d0 = __shfl_sync( SP_FULL, my_doc, __ffs( workmask) - 1);
		if (__all_sync( SP_FULL, acc.doc == d0 || acc.doc == 0xFFFFFFFFu))
		{
			uint64_t inst = acc.doc == d0 ? acc.installed : 0, sig = acc.doc == d0 ? acc.signals : 0, open = acc.doc == d0 ? acc.open : 0;
			#pragma unroll
			for (int ofs=16; ofs>0; ofs>>=1)
			{
				inst += __shfl_xor_sync( SP_FULL, inst, ofs); sig += __shfl_xor_sync( SP_FULL, sig, ofs); open += __shfl_xor_sync( SP_FULL, open, ofs);
			}
			if (lane == 0) evalAddStats( P.stats + d0, inst, 0, sig, open);
			acc.doc = 0xFFFFFFFFu;
		}
	}
	evalAccDoc( P, acc, 0xFFFFFFFFu);
	__syncwarp();
	if (lane < ntile) {P.tok_nres[ r0 + lane] = snres[ lane]; P.tok_nitems[ r0 + lane] = snitems[ lane];}
}

// work item number -> token (binary search over the first items of the tokens), site and list entry
__device__ __forceinline__ void evalDecodeItem( const EvalParams& P, uint64_t item, uint64_t& tok, uint32_t& slotidx, EvalCursor& cur)
{
	uint64_t lo = 0, hi = P.ntokens;	// last token with tok_work_off[t] <= item
	while (hi - lo > 1)
	{
		uint64_t mid = lo + (hi - lo) / 2;
		if (__ldg( &P.tok_work_off[ mid]) <= item) lo = mid; else hi = mid;
	}
	tok = lo;
	slotidx = P.tok_slot[ lo];
	evalCursorSeek( P.T, slotidx, (uint32_t)(item - __ldg( &P.tok_work_off[ lo])), cur);
}

__device__ __forceinline__ void evalItemSite( const EvalParams& P, uint64_t tok, uint32_t slotidx, const EvalCursor& cur, SpEvalSite& site)
{
	SpEvalSite base;
	sp_eval_site_token( P.T, P.tokens[ tok], P.origsegs ? P.origsegs[ tok] : 0, tok, slotidx, base);
	if (cur.c) sp_eval_site_follow( base, cur.c - 1, P.T.eval_follows[ cur.fb + cur.c - 1], site);
	else site = base;
}

// The queued items of the count pass by the full fold, one per thread.
static __global__ void __launch_bounds__(128) k_eval_count_full( EvalParams P)
{
	const uint64_t nq = min( (uint64_t)*P.queue_count, P.queue_cap);
	for (uint64_t q = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; q < nq; q += (uint64_t)gridDim.x * blockDim.x)
	{
		const uint64_t entry = P.queue[ q];
		const uint64_t item = entry & ~SP_EVALQ_STATS_DONE;
		uint64_t tok; uint32_t slotidx; EvalCursor cur;
		evalDecodeItem( P, item, tok, slotidx, cur);
		const uint32_t d = P.tok_doc[ tok];
		SpEvalSite site;
		evalItemSite( P, tok, slotidx, cur, site);
		SpEvalRule R;
		sp_eval_fold<false,0>( P.T, P.tokens, P.origsegs, __ldg( &P.tok_off[ d]), __ldg( &P.tok_off[ d+1]), site, P.T.program_triggers[ cur.lbegin + cur.k], false, R, 0);
		if (R.flag) P.flags[ d] = 1;
		if (!(entry & SP_EVALQ_STATS_DONE)) evalAddStats( P.stats + d, R.installed, R.altkey, R.signals, R.open);
		if (R.has_result)
		{
			atomicAdd( P.tok_nres + tok, 1u);
			const uint32_t nit = sp_eval_result_items( R);
			if (nit) atomicAdd( P.tok_nitems + tok, nit);
			atomicOr( P.resbits + (item >> 5), 1u << (item & 31));
		}
	}
}

// Write pass: the result bits of a run of tokens, 32 results at a time.
template <int BLOCKS_PER_SM>
static __global__ void __launch_bounds__(EvalThreads, BLOCKS_PER_SM) k_eval_write( EvalParams P)
{
	__shared__ uint32_t s_work[ EvalWarpsPerBlock][ EvalRun + 1];
	__shared__ uint32_t s_word[ EvalWarpsPerBlock][ 32];
	__shared__ uint32_t s_pre[ EvalWarpsPerBlock][ 33];
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint32_t* swork = s_work[ warp]; uint32_t* sword = s_word[ warp]; uint32_t* spre = s_pre[ warp];
	const uint64_t run = (uint64_t)blockIdx.x * EvalWarpsPerBlock + warp;
	const uint64_t r0 = run * EvalRun;
	if (r0 >= P.ntokens) return;
	const uint32_t nrun = (uint32_t)min( (uint64_t)EvalRun, P.ntokens - r0);
	const uint64_t res0 = __ldg( &P.tok_res_off[ r0]);
	const uint32_t nres_run = (uint32_t)(__ldg( &P.tok_res_off[ r0 + nrun]) - res0);
	if (nres_run == 0) return;
	const uint64_t wbase = __ldg( &P.tok_work_off[ r0]);
	const uint32_t W = (uint32_t)(__ldg( &P.tok_work_off[ r0 + nrun]) - wbase);
	for (uint32_t j=lane; j<=EvalRun; j+=32) swork[ j] = (uint32_t)(__ldg( &P.tok_work_off[ r0 + min( j, nrun)]) - wbase);
	__syncwarp();
	uint64_t res_done = 0;				// results of the run in front of the current group of words
	uint64_t carry_tok = ~0ull; uint32_t carry_items = 0;	// items of the results of the last token seen so far
	const uint64_t firstword = wbase >> 5, lastword = (wbase + W - 1) >> 5;
	for (uint64_t wordbase=firstword; wordbase<=lastword; wordbase+=32)
	{
		// 32 words of the bit map, restricted to the items of the run
		const uint64_t wi = wordbase + lane;
		uint32_t word = 0;
		if (wi <= lastword)
		{
			word = P.resbits[ wi];
			if (wi == firstword) word &= 0xFFFFFFFFu << (wbase & 31);
			if (wi == lastword && ((wbase + W) & 31)) word &= 0xFFFFFFFFu >> (32 - ((wbase + W) & 31));
		}
		uint32_t cnt = __popc( word), incl = cnt;
		#pragma unroll
		for (int ofs=1; ofs<32; ofs<<=1)
		{
			uint32_t up = __shfl_up_sync( SP_FULL, incl, ofs);
			if (lane >= (unsigned int)ofs) incl += up;
		}
		const uint32_t total = __shfl_sync( SP_FULL, incl, 31);
		if (!total) continue;
		__syncwarp();
		sword[ lane] = word; spre[ lane] = incl - cnt;
		if (lane == 31) spre[ 32] = total;
		__syncwarp();
		for (uint32_t j0=0; j0<total; j0+=32)
		{
			const uint32_t j = j0 + lane;
			const bool valid = j < total;
			SpEvalRule R; sp_eval_rule_init( R);
			SpEvalTake takes[ SP_EVAL_MAX_TAKES];
			SpEvalTrace tr; tr.flags = 0; tr.ntakes = 0; tr.nitems = 0; tr.first_rel = tr.emit_rel = tr.emit_end = 0;
			SpEvalSite site;
			uint64_t tok = ~0ull - 1 - lane, tb = 0, te = 0;
			uint32_t d = 0, ptidx = 0, nit = 0;
			bool live = false, lite = false;
			const SpEvalFollow* F = 0;
			if (valid)
			{
				uint32_t lo = 0, hi = 32;	// word that holds result j: last word with spre <= j
				while (hi - lo > 1) {uint32_t mid = (lo + hi) / 2; if (spre[ mid] <= j) lo = mid; else hi = mid;}
				const uint32_t bit = __fns( sword[ lo], 0, (int)(j - spre[ lo]) + 1);
				const uint32_t wl = (uint32_t)(((wordbase + lo) << 5) + bit - wbase);		// item of the run
				uint32_t tlo = 0, thi = EvalRun;
				while (thi - tlo > 1) {uint32_t mid = (tlo + thi) / 2; if (swork[ mid] <= wl) tlo = mid; else thi = mid;}
				tok = r0 + tlo;
				const uint32_t slotidx = P.tok_slot[ tok];
				EvalCursor cur;
				evalCursorSeek( P.T, slotidx, wl - swork[ tlo], cur);
				d = P.tok_doc[ tok];
				tb = __ldg( &P.tok_off[ d]); te = __ldg( &P.tok_off[ d+1]);
				live = !P.flags[ d];	// the results of flagged documents come from the state machine
				if (live)
				{
					evalItemSite( P, tok, slotidx, cur, site);
					ptidx = cur.lbegin + cur.k;
					if (cur.c) F = P.T.eval_follows + cur.fb + cur.c - 1;
				}
			}
			// the rule once more, with the taken events recorded: from its compiled list entry (all lanes run the same short
			// loop) or, where the entry is not enough, by the full fold
			if (live)
			{
				const SpEvalEntry en = P.T.eval_entries[ ptidx];
				lite = sp_eval_lite_trace( en, P.T.eval_vars[ ptidx], P.tokens, tb, te, site, tr);
				uint32_t extra = 0;
				if (lite && F && !sp_eval_lite_follow_items( P.T, P.tokens, tok, te, en, *F, extra)) lite = false;
				if (lite) nit = (tr.flags & SP_LITE_DREFEMIT) ? tr.nitems + extra : 0;
			}
			if (live && !lite)
			{
				sp_eval_fold<true,0>( P.T, P.tokens, P.origsegs, tb, te, site, P.T.program_triggers[ ptidx], false, R, takes);
				nit = sp_eval_result_items( R);
			}
			// items of the earlier results of the same token (the results of a token are neighbours)
			uint32_t seg = nit;
			#pragma unroll
			for (int ofs=1; ofs<32; ofs<<=1)
			{
				const uint32_t up = __shfl_up_sync( SP_FULL, seg, ofs);
				const uint64_t uptok = __shfl_up_sync( SP_FULL, tok, ofs);
				if (lane >= (unsigned int)ofs && uptok == tok) seg += up;
			}
			const uint32_t before = seg - nit + (tok == carry_tok ? carry_items : 0u);
			if (live && (lite || R.has_result))
			{
				const uint64_t res_in_doc = (res0 + res_done + j) - __ldg( &P.tok_res_off[ tb]);
				const uint64_t item_in_doc = (__ldg( &P.tok_item_off[ tok]) - __ldg( &P.tok_item_off[ tb])) + before;
				sp_item* iout = P.items + __ldg( &P.doc_item_off[ d]) + item_in_doc;
				sp_result& rout = P.results[ __ldg( &P.doc_res_off[ d]) + res_in_doc];
				const SpProgram prog = P.T.programs[ P.T.program_triggers[ ptidx].program];
				if (lite)
				{
					bool ok = true;
					if (nit) sp_eval_lite_write_items( P.T, P.tokens, P.origsegs, tb, te, site, P.T.eval_vars[ ptidx], tr, F, iout, ok);
					sp_eval_lite_result_record( prog, P.tokens, P.origsegs, tb, tr, (uint32_t)item_in_doc, nit, rout);
				}
				else
				{
					if (nit) sp_eval_write_items( P.T, P.tokens, P.origsegs, tb, te, takes, R.ntakes, iout);
					sp_eval_result_record( prog, R, (uint32_t)item_in_doc, nit, rout);
				}
			}
			const int last = (int)min( total - j0, 32u) - 1;
			const uint64_t ltok = __shfl_sync( SP_FULL, tok, last);
			const uint32_t lseg = __shfl_sync( SP_FULL, seg, last);
			carry_items = lseg + (ltok == carry_tok ? carry_items : 0u);
			carry_tok = ltok;
		}
		res_done += total;
	}
}

// Per document: counts from the token prefix sums, status, number of events; flagged documents are listed.
// totals[0] results, [1] items of the documents that stay on the evaluator; nflagged = length of the list.
static __global__ void k_eval_docs( EvalParams P, uint32_t* doc_res_count, uint32_t* doc_item_count, uint32_t* status,
				    uint32_t* list, uint32_t* nflagged, unsigned long long* totals)
{
	const uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
	if (d >= P.ndocs) return;
	const uint64_t tb = P.tok_off[ d], te = P.tok_off[ d+1];
	status[ d] = 0;
	if (P.flags[ d])
	{
		doc_res_count[ d] = 0; doc_item_count[ d] = 0;
		list[ atomicAdd( nflagged, 1u)] = d;
		return;
	}
	const uint64_t nr = P.tok_res_off[ te] - P.tok_res_off[ tb], ni = P.tok_item_off[ te] - P.tok_item_off[ tb];
	doc_res_count[ d] = (uint32_t)nr; doc_item_count[ d] = (uint32_t)ni;
	P.stats[ d].nof_events = te - tb;
	if (nr) atomicAdd( totals + 0, (unsigned long long)nr);
	if (ni) atomicAdd( totals + 1, (unsigned long long)ni);
}

// Results of the documents that ran on the state machine: from its pools to their place (one warp per listed document).
static __global__ void __launch_bounds__(256) k_eval_place( const uint32_t* list, uint32_t nlist, MatchParams M, const uint64_t* res_off, const uint64_t* item_off,
							   sp_result* results, sp_item* items)
{
	const unsigned int lane = threadIdx.x & 31;
	const uint32_t w = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (w >= nlist) return;
	const uint32_t d = list[ w];
	{
		const uint4* src = reinterpret_cast<const uint4*>( M.item_pool + M.doc_item_begin[ d]);
		uint4* dst = reinterpret_cast<uint4*>( items + item_off[ d]);
		for (uint64_t i=lane, n=2ull*M.doc_item_count[ d]; i<n; i+=32) dst[ i] = src[ i];
	}
	{
		const uint2* src = reinterpret_cast<const uint2*>( M.res_pool + M.doc_res_begin[ d]);
		uint2* dst = reinterpret_cast<uint2*>( results + res_off[ d]);
		for (uint64_t i=lane, n=5ull*M.doc_res_count[ d]; i<n; i+=32) dst[ i] = src[ i];
	}
}

} // namespace sp
#endif
