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

static __global__ void __launch_bounds__(MatchThreads) k_match_documents( MatchParams P)
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
		slot = __shfl_sync( 0xFFFFFFFFu, slot, 0);
		if (slot >= ntodo) break;
		const uint32_t d = P.doc_list ? P.doc_list[ slot] : slot;
		const uint64_t tb = P.tok_off[ d], te = P.tok_off[ d+1];
		if (lane == 0) sp_doc_begin( S);
		__syncwarp();
		uint64_t nevents = 0;
		uint32_t status = 0;
		for (uint64_t ti=tb; ti<te; ++ti)
		{
			if (lane == 0)
			{
				sp_lexem tok = P.tokens[ ti];
				sp_input_begin( A, S, tok, P.origsegs ? P.origsegs[ ti] : 0);
			}
			for (uint32_t fi=0;; ++fi)
			{
				// everything the other lanes need from the state lane 0 mutates is read between two barriers
				__syncwarp();
				status = S.status;
				const uint32_t nfollow = S.nfollow;
				uint32_t ev = 0, b = 0, n = 0;
				if (!status && fi < nfollow)
				{
					ev = A.follow[ fi].event;
					if (ev) {b = sp_evhash( ev); n = S.bucket_size[ b];}
				}
				__syncwarp();
				if (status || fi >= nfollow) break;
				if (ev)
				{
					const uint32_t* evs = A.bucket_event + (size_t)b * A.cap.triggers;
					for (uint32_t base=0; base<n; base+=32)
					{
						uint32_t e = (base + lane < n) ? evs[ base + lane] : 0xFFFFFFFFu;
						unsigned int hits = __ballot_sync( 0xFFFFFFFFu, e == ev);
						if (hits && lane == 0)
						{
							SpEventData data = A.follow[ fi].data;
							while (hits && !S.status)
							{
								int i = __ffs( hits) - 1;
								hits &= hits - 1;
								sp_fire_trigger_at( A, S, b, base + i, data);
							}
						}
					}
				}
				if (lane == 0 && !S.status) sp_event_finish( P.T, A, S, fi);
			}
			if (status) break;
			++nevents;
		}
		__syncwarp();
		if (lane == 0)
		{
			uint32_t nres = 0, nitems = 0;
			if (!S.status) sp_doc_count_output( P.T, A, S, nres, nitems);
			if (S.status) {nres = 0; nitems = 0;}
			uint64_t rb = 0, ib = 0;
			if (nres) rb = atomicAdd( P.cursors + 0, (unsigned long long)nres);
			if (nitems) ib = atomicAdd( P.cursors + 1, (unsigned long long)nitems);
			if (rb + nres > P.res_cap || ib + nitems > P.item_cap)
			{
				S.status = SP_MST_OVF_OUTPUT; nres = 0; nitems = 0;
			}
			else if (nres) sp_doc_write_output( A, S, P.res_pool + rb, P.item_pool + ib);
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
	else {atomicOr( kinds + 0, 1u); list[ atomicAdd( count, 1u)] = d;}
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

} // namespace sp
#endif
