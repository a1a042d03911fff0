// CUDA kernels of the level-2 matcher (sm_100a): one warp per document, persistent grid.
//
//   k_match_documents  every warp owns one arena (matcher_device.cuh: SpArena) in HBM and one SpDocState in shared
//                      memory and pulls documents from an atomic work counter. Per input token: lane 0 advances
//                      the position (rule expiry) and queues the event; for every event of the follow list all 32
//                      lanes scan the event's trigger bucket (coalesced 128 byte rows + __ballot_sync, the analogue
//                      of getTriggers_SSE4, src/ruleMatcherAutomaton.cpp:179-226), lane 0 fires the hits in array
//                      order and installs the programs keyed by the event. At the end of the document the results
//                      and their items are written to pool space claimed with one atomicAdd per document.
//   k_match_collect    lists the documents whose arena overflowed (they run again with larger arenas).
//   k_match_compact    one warp per document: copies results and items from the pool into document order.
#ifndef SP_MATCHER_KERNELS_CUH_INCLUDED
#define SP_MATCHER_KERNELS_CUH_INCLUDED
#include "matcher_device.cuh"
#include "matcher_warp.cuh"

namespace sp {

enum {MatchWarpsPerBlock = 4, MatchThreads = MatchWarpsPerBlock * 32};

struct MatchParams
{
	SpMatcherTables T;
	const sp_lexem* tokens;
	const uint32_t* origsegs;		// may be 0
	const uint64_t* tok_off;		// [ndocs+1]
	uint32_t ndocs;
	const uint32_t* doc_list;		// 0: all documents, else the documents to run
	uint32_t nlist;
	uint8_t* arena_base;			// nwarps arenas of layout.bytes each
	SpArenaLayout layout;
	uint32_t* next_doc;			// work counter
	// output pools and per document descriptors
	sp_result* res_pool; sp_item* item_pool;
	uint64_t res_cap, item_cap;
	unsigned long long* cursors;		// [0] results, [1] items claimed so far
	uint64_t* doc_res_begin; uint32_t* doc_res_count;
	uint64_t* doc_item_begin; uint32_t* doc_item_count;
	sp_matcher_stats* stats;
	uint32_t* status;
};

// uniform view of the document status (lanes may look at the word at different times)
__device__ __forceinline__ bool warpFailed( const SpDocState& S) {return __any_sync( SP_FULL, S.status != 0) != 0;}

// Two builds of the kernel: the per-document state machine is a chain of dependent loads, so throughput comes from
// warps in flight. MINBLOCKS = 8 caps the kernel at 64 registers (32 warps per SM, spills go to L1): 75 instead of 55 M
// tokens/s when there are documents for all of them; with fewer documents than 16 warps per SM the build with
// MINBLOCKS = 4 (128 registers, 16 warps per SM) finishes each chain sooner (about 40 against 37.8 M tokens/s at 1 000 documents).
template <int MINBLOCKS>
static __global__ void __launch_bounds__(MatchThreads, MINBLOCKS) k_match_documents( MatchParams P)
{
	__shared__ SpDocState states[ MatchWarpsPerBlock];
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	SpDocState& S = states[ warp];
	SpArena A;
	sp_arena_bind( P.layout, P.arena_base + (size_t)(blockIdx.x * MatchWarpsPerBlock + warp) * P.layout.bytes, A);
	if (lane == 0) S.stamp = 0;
	const uint32_t ntodo = P.doc_list ? P.nlist : P.ndocs;
	for (;;)
	{
		uint32_t slot = 0;
		if (lane == 0) slot = atomicAdd( P.next_doc, 1u);
		slot = __shfl_sync( SP_FULL, slot, 0);
		if (slot >= ntodo) break;
		const uint32_t d = P.doc_list ? P.doc_list[ slot] : slot;
		const uint64_t tb = P.tok_off[ d], te = P.tok_off[ d+1];
		if (lane == 0) sp_doc_begin( S);
		__syncwarp();
		uint64_t nevents = 0;
		bool failed = false;
		for (uint64_t ti=tb; ti<te && !failed; ++ti)
		{
			// putInput: checks, rule expiry by the whole warp, first entry of the follow list
			const sp_lexem tok = P.tokens[ ti];
			const uint32_t seg = P.origsegs ? P.origsegs[ ti] : 0;
			if (lane == 0) sp_input_check( S, tok, seg);
			__syncwarp();
			if ((failed = warpFailed( S))) break;
			warpSetCurrentPos( A, S, lane, tok.ordpos);
			if (lane == 0 && !S.status) sp_input_queue( A, S, tok, seg);
			__syncwarp();
			if ((failed = warpFailed( S))) break;
			// doTransition: the follow list grows while it is worked off
			for (uint32_t fi=0;; ++fi)
			{
				const uint32_t nfollow = S.nfollow;
				__syncwarp();
				if (fi >= nfollow) break;
				const SpFollow follow = A.follow[ fi];
				if (follow.event)
				{
					const uint32_t b = sp_evhash( follow.event), n = S.bucket_size[ b];
					__syncwarp();
					warpFireHits( A, S, lane, follow.event, b, n, follow.data);
					if ((failed = warpFailed( S))) break;
				}
				SpEventSlot es; es.prog_count = 0; es.stopword = 0;
				const bool known = sp_find_event( P.T, follow.event, es);
				if (known && es.prog_count)
				{
					warpInstall( P.T, A, S, lane, follow.event, es, follow.data);
					if ((failed = warpFailed( S))) break;
				}
				const uint32_t ndispose = S.ndispose;
				__syncwarp();
				if (ndispose)
				{
					warpDeactivate( A, S, lane, A.dispose, ndispose, false, false);
					if (lane == 0) S.ndispose = 0;
				}
				if (lane == 0)
				{
					if (known && es.stopword)
					{
						SpStopLog& log = A.stoplog[ es.stopword-1];
						log.data = follow.data; log.timestmp = ++S.timestmp; log.stamp = S.stamp;
					}
					else if (follow.data.subdataref) sp_dataref_release( A, S, follow.data.subdataref);
				}
				__syncwarp();
				if ((failed = warpFailed( S))) break;
			}
			if (!failed) ++nevents;
		}
		__syncwarp();
		// fetchResults: one lane per result counts its items, the document claims pool space once, the lanes write
		if (lane == 0 && !S.status && P.T.exclusive) sp_mark_covered( P.T, A, S);
		__syncwarp();
		const uint32_t npending = warpFailed( S) ? 0 : S.nresults;
		uint32_t nres = 0, nitems = 0;
		for (uint32_t base=0; base<npending; base+=32)
		{
			const uint32_t i = base + lane;
			uint32_t keep = 0, cnt = 0;
			if (i < npending && !A.results[ i].eliminated)
			{
				keep = 1;
				if (A.results[ i].dataref) cnt = sp_gather_items( A, S, A.results[ i].dataref, 0);
			}
			uint32_t ki = keep, ci = cnt;
			#pragma unroll
			for (int ofs=1; ofs<32; ofs<<=1)
			{
				uint32_t uk = __shfl_up_sync( SP_FULL, ki, ofs), uc = __shfl_up_sync( SP_FULL, ci, ofs);
				if (lane >= (unsigned int)ofs) {ki += uk; ci += uc;}
			}
			if (keep) {A.results[ i].out_index = nres + ki - 1; A.results[ i].item_begin = nitems + ci - cnt;}
			nres += __shfl_sync( SP_FULL, ki, 31);
			nitems += __shfl_sync( SP_FULL, ci, 31);
		}
		__syncwarp();
		if (warpFailed( S)) {nres = 0; nitems = 0;}
		unsigned long long rb = 0, ib = 0;
		if (lane == 0)
		{
			if (nres) rb = atomicAdd( P.cursors + 0, (unsigned long long)nres);
			if (nitems) ib = atomicAdd( P.cursors + 1, (unsigned long long)nitems);
			if (rb + nres > P.res_cap || ib + nitems > P.item_cap) S.status = SP_MST_OVF_OUTPUT;
		}
		rb = __shfl_sync( SP_FULL, rb, 0); ib = __shfl_sync( SP_FULL, ib, 0);
		__syncwarp();
		if (warpFailed( S)) {nres = 0; nitems = 0;}
		if (nres)
		{
			for (uint32_t i=lane; i<npending; i+=32)
			{
				const SpPendingResult res = A.results[ i];
				if (res.eliminated) continue;
				uint32_t cnt = res.dataref ? sp_gather_items( A, S, res.dataref, P.item_pool + ib + res.item_begin) : 0;
				sp_result_record( res, res.item_begin, cnt, P.res_pool[ rb + res.out_index]);
			}
		}
		if (lane == 0)
		{
			P.doc_res_begin[ d] = rb; P.doc_res_count[ d] = nres;
			P.doc_item_begin[ d] = ib; P.doc_item_count[ d] = nitems;
			sp_matcher_stats st;
			sp_doc_stats( S, S.status ? 0 : nevents, st);
			if (S.status) st.nof_open_patterns = 0;
			P.stats[ d] = st;
			P.status[ d] = S.status;
		}
		__syncwarp();
	}
}

// documents that need another run: list[0..*count) ; kinds[0] |= arena overflow, kinds[1] |= output pool overflow
static __global__ void k_match_collect( const uint32_t* status, uint32_t ndocs, uint32_t* list, uint32_t* count, uint32_t* kinds)
{
	uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
	if (d >= ndocs) return;
	uint32_t st = status[ d];
	if (st < SP_MST_OVERFLOW) return;
	if (st == SP_MST_OVF_OUTPUT) atomicOr( kinds + 1, 1u);
	else if (st == SP_MST_OVF_DEPTH) atomicOr( kinds + 2, 1u);
	else {atomicOr( kinds + 0, 1u << (st - SP_MST_OVERFLOW)); list[ atomicAdd( count, 1u)] = d;}	// bit = which pool of the arena
}

static __global__ void __launch_bounds__(256) k_match_compact( MatchParams P, const uint64_t* res_off, const uint64_t* item_off, sp_result* results, sp_item* items)
{
	const unsigned int lane = threadIdx.x & 31;
	const uint32_t d = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (d >= P.ndocs) return;
	{
		const uint4* src = reinterpret_cast<const uint4*>( P.item_pool + P.doc_item_begin[ d]);	// sp_item = 2 x 16 bytes
		uint4* dst = reinterpret_cast<uint4*>( items + item_off[ d]);
		for (uint64_t i=lane, n=2ull*P.doc_item_count[ d]; i<n; i+=32) dst[ i] = src[ i];
	}
	{
		const uint2* src = reinterpret_cast<const uint2*>( P.res_pool + P.doc_res_begin[ d]);		// sp_result = 5 x 8 bytes
		uint2* dst = reinterpret_cast<uint2*>( results + res_off[ d]);
		for (uint64_t i=lane, n=5ull*P.doc_res_count[ d]; i<n; i+=32) dst[ i] = src[ i];
	}
}

// ---- compact wire format (include/strus_pattern_b200.h: sp_result_packed, sp_item_packed): elementwise, *notpackable is
// raised by any record that does not fit
static __global__ void __launch_bounds__(256) k_pack_results( const sp_result* __restrict__ in, uint64_t n, sp_result_packed* __restrict__ out, uint32_t* notpackable)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const sp_result r = in[ i];
	const uint32_t ordlen = r.ordend - r.ordpos, origlen = r.endpos - r.origpos;
	if (r.origseg | r.endseg | (r.name >> 18) | (r.format >> 14) | (ordlen >> 8) | (origlen >> 16) | (r.item_count >> 8)) *notpackable = 1;
	uint4 o;
	o.x = r.name | (r.format << 18); o.y = r.ordpos; o.z = r.origpos; o.w = ordlen | (r.item_count << 8) | (origlen << 16);
	*reinterpret_cast<uint4*>( out + i) = o;
}
static __global__ void __launch_bounds__(256) k_pack_items( const sp_item* __restrict__ in, uint64_t n, sp_item_packed* __restrict__ out, uint32_t* notpackable)
{
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= n) return;
	const uint4 a = __ldg( reinterpret_cast<const uint4*>( in + i)), b = __ldg( reinterpret_cast<const uint4*>( in + i) + 1);
	// a: variable, format, ordpos, ordend; b: origseg, origpos, endseg, endpos
	const uint32_t ordlen = a.w - a.z, origlen = b.w - b.y;
	if (b.x | b.z | a.y | (a.x >> 12) | (ordlen >> 4) | (origlen >> 16)) *notpackable = 1;
	sp_item_packed o;
	o.var_lens = a.x | (ordlen << 12) | (origlen << 16); o.ordpos = a.z; o.origpos = b.y;
	out[ i] = o;
}

} // namespace sp
#endif
