// Island-parallel resolution and scan-based ordinal positions (second half of the lexer pipeline).
//
//   k_isl_reduce / k_isl_carry / k_isl_flags   prefix max of event ends and suffix min of event starts (keys carry the
//                   document in the upper half, so no segmentation is needed); event k starts an island iff
//                   min(start of k..) >= max(end of ..k-1). Also counts the island starts per tile.
//   k_isl_compact   island start list from the flags.
//   k_isl_resolve   one thread per island: reference supersede logic (src/patternLexer.cpp:754-822) in thread local
//                   memory (global scratch for oversized islands); pass 0 counts, pass 1 writes the kept elements densely.
//   ordinal positions (src/patternLexer.cpp:893-945) over the dense elements:
//       scan A (max)     which elements advance the position        -> SP_META_INC
//       scan B (segsum)  position per element, emit decision        -> ordpos / SP_SIZE_EMIT
//       scan C (sum)     output index, lexem records and per document offsets
#ifndef SP_LEXER_ISLANDS_CUH_INCLUDED
#define SP_LEXER_ISLANDS_CUH_INCLUDED
#include "lexer_device.cuh"
#include "tile_scan.cuh"

namespace sp {

struct IslandParams
{
	SpLexerTables T;
	const SpRawEvent* events;
	const uint32_t* ev_doc;		// [nev] document of every event
	uint64_t nev;
	uint8_t* flags;			// [nev] 1 = first event of an island
	uint32_t* tile_nisl;		// [ntiles] island starts per tile
	const uint64_t* tile_isl_off;	// [ntiles+1]
	uint32_t* isl_start;		// [nisl+1] event index
	uint64_t nisl;
	uint32_t* isl_kept;		// [nisl]
	const uint64_t* isl_off;	// [nisl+1] dense element offsets
	SpMatchEvent* scratch;		// [nev*ar_mult] for islands that do not fit thread local memory
	uint32_t ar_mult;
	SpMatchEvent* elem;		// dense kept elements
	uint32_t* elem_doc;
	uint64_t nelem;
	uint32_t* doc_status;		// [ndocs]
	uint8_t* doc_has_content;	// [ndocs]
	uint64_t* doc_elem_begin;	// [ndocs+1] first dense element of every document
	uint32_t ndocs;
	sp_lexem* lexems;
	uint64_t* lex_off;		// [ndocs+1]
};

static __global__ void __launch_bounds__(TsThreads) k_isl_reduce( IslandParams P, uint64_t* aggmax, uint64_t* aggmin)
{
	__shared__ uint64_t smem[ 33];
	uint64_t first = (uint64_t)blockIdx.x * TsTile + (uint64_t)threadIdx.x * TsItems;
	uint64_t mx = 0, mn = ~0ull;
	#pragma unroll
	for (int k=0; k<TsItems; ++k)
	{
		if (first + k >= P.nev) break;
		SpRawEvent ev = P.events[ first + k];
		uint32_t doc = P.ev_doc[ first + k];
		mx = OpMax::apply( mx, sp_event_r_key( ev, doc));
		mn = OpMin::apply( mn, sp_event_lo_key( ev, doc));
	}
	uint64_t excl, total;
	blockScan<OpMax>( mx, smem, excl, total);
	if (threadIdx.x == 0) aggmax[ blockIdx.x] = total;
	blockScan<OpMin>( mn, smem, excl, total);
	if (threadIdx.x == 0) aggmin[ blockIdx.x] = total;
}

static __global__ void __launch_bounds__(TsThreads) k_isl_flags( IslandParams P, const uint64_t* carrymax, const uint64_t* carrymin)
{
	__shared__ uint64_t smem[ 33];
	__shared__ uint64_t mirror[ TsThreads];
	__shared__ uint32_t warpcnt[ TsThreads/32];
	uint64_t first = (uint64_t)blockIdx.x * TsTile + (uint64_t)threadIdx.x * TsItems;
	uint64_t rk[ TsItems], lk[ TsItems];
	uint64_t mx = 0, mn = ~0ull;
	#pragma unroll
	for (int k=0; k<TsItems; ++k)
	{
		rk[k] = 0; lk[k] = ~0ull;
		if (first + k < P.nev)
		{
			SpRawEvent ev = P.events[ first + k];
			uint32_t doc = P.ev_doc[ first + k];
			rk[k] = sp_event_r_key( ev, doc); lk[k] = sp_event_lo_key( ev, doc);
		}
		mx = OpMax::apply( mx, rk[k]); mn = OpMin::apply( mn, lk[k]);
	}
	uint64_t exmax, exmin, total;
	blockScan<OpMax>( mx, smem, exmax, total);
	// suffix min over the threads: prefix scan in mirrored thread order
	mirror[ TsThreads-1-threadIdx.x] = mn;
	__syncthreads();
	uint64_t mv = mirror[ threadIdx.x];
	__syncthreads();
	blockScan<OpMin>( mv, smem, exmin, total);
	mirror[ TsThreads-1-threadIdx.x] = exmin;
	__syncthreads();
	exmin = mirror[ threadIdx.x];
	uint64_t premax[ TsItems];
	uint64_t run = OpMax::apply( carrymax[ blockIdx.x], exmax);
	#pragma unroll
	for (int k=0; k<TsItems; ++k) {premax[k] = run; run = OpMax::apply( run, rk[k]);}
	uint64_t suf = OpMin::apply( carrymin[ blockIdx.x], exmin);
	uint32_t cnt = 0;
	#pragma unroll
	for (int k=TsItems-1; k>=0; --k)
	{
		suf = OpMin::apply( suf, lk[k]);
		if (first + k < P.nev)
		{
			uint8_t f = (suf >= premax[k]) ? 1 : 0;
			P.flags[ first + k] = f;
			cnt += f;
		}
	}
	#pragma unroll
	for (int ofs=16; ofs>0; ofs>>=1) cnt += __shfl_down_sync( 0xFFFFFFFFu, cnt, ofs);
	if ((threadIdx.x & 31) == 0) warpcnt[ threadIdx.x >> 5] = cnt;
	__syncthreads();
	if (threadIdx.x == 0)
	{
		uint32_t t = 0;
		for (int w=0; w<TsThreads/32; ++w) t += warpcnt[ w];
		P.tile_nisl[ blockIdx.x] = t;
	}
}

static __global__ void __launch_bounds__(TsThreads) k_isl_compact( IslandParams P)
{
	__shared__ uint64_t smem[ 33];
	uint64_t first = (uint64_t)blockIdx.x * TsTile + (uint64_t)threadIdx.x * TsItems;
	uint32_t f[ TsItems]; uint64_t cnt = 0;
	#pragma unroll
	for (int k=0; k<TsItems; ++k) {f[k] = (first + k < P.nev) ? P.flags[ first + k] : 0; cnt += f[k];}
	uint64_t excl, total;
	blockScan<OpSum>( cnt, smem, excl, total);
	uint64_t pos = P.tile_isl_off[ blockIdx.x] + excl;
	#pragma unroll
	for (int k=0; k<TsItems; ++k) if (f[k]) P.isl_start[ pos++] = (uint32_t)(first + k);
	if (blockIdx.x == 0 && threadIdx.x == 0) P.isl_start[ P.nisl] = (uint32_t)P.nev;
}

// ------------------------------------------------------------------------------------------------
// Slot pipeline (lexer_device.cuh "slot formulation"): one resolution pass per island, one sum scan over the slots.
//   k_isl_to_slots  one thread per island: supersede logic, COUNTS/INC flags, slot records; lists the islands whose
//                   first element is BindUnique
//   k_isl_fix       the listed islands look at the element in front of them (rare)
//   SlotScan        tile reduce / carry / apply over the slots: low word positions advanced, high word elements emitted
//   k_doc_bases     scan value in front of the first slot of every document -> ordinal position base, lexem offsets
static __global__ void __launch_bounds__(128) k_isl_to_slots( SpSlotParams P)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i >= P.nisl) return;
	uint64_t eb = P.isl_start[ i], ee = P.isl_start[ i+1];
	SpMatchEvent local[ SP_ISLAND_LOCAL];
	// oversized islands work in their own slot range (HBM) instead of thread local memory
	SpMatchEvent* ar = ((ee - eb) * P.ar_mult <= SP_ISLAND_LOCAL) ? local : reinterpret_cast<SpMatchEvent*>( P.slots + eb * P.ar_mult);
	if (sp_island_to_slots( P, i, ar)) P.fix_list[ atomicAdd( P.fix_count, 1u)] = (uint32_t)i;
}

static __global__ void __launch_bounds__(128) k_isl_fix( SpSlotParams P)
{
	const uint32_t nfix = *P.fix_count;
	for (uint32_t f = blockIdx.x * blockDim.x + threadIdx.x; f < nfix; f += gridDim.x * blockDim.x) sp_island_fix_unique( P, P.fix_list[ f]);
}

struct SlotScan
{
	typedef OpSum Op;
	SpSlotParams P;
	__device__ uint64_t load( uint64_t j) const
	{
		SpSlot s = reinterpret_cast<const SpSlot*>( P.slots)[ j];
		return sp_slot_value( P, j, s);
	}
	__device__ void store( uint64_t j, uint64_t excl, uint64_t incl) const
	{
		uint64_t value = incl - excl;
		if (!(value >> 32)) return;
		SpSlot s = P.slots[ j];
		sp_slot_emit( P, j, s, excl, value);
	}
};

static __global__ void __launch_bounds__(256) k_doc_bases( SpSlotParams P, const uint64_t* carry, const uint64_t* total)
{
	const unsigned int lane = threadIdx.x & 31;
	const uint32_t d = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
	if (d > P.ndocs) return;
	const uint64_t nslots = P.nev * P.ar_mult;
	const uint64_t fs = P.doc_ev_begin[ d] * P.ar_mult;
	uint64_t base;
	if (fs >= nslots) base = *total;
	else
	{
		const uint64_t tile = fs / TsTile;
		uint64_t sum = 0;
		for (uint64_t j = tile * TsTile + lane; j < fs; j += 32) sum += sp_slot_value( P, j, P.slots[ j]);
		#pragma unroll
		for (int ofs=16; ofs>0; ofs>>=1) sum += __shfl_down_sync( 0xFFFFFFFFu, sum, ofs);
		base = carry[ tile] + __shfl_sync( 0xFFFFFFFFu, sum, 0);
	}
	if (lane == 0) {P.doc_base[ d] = base; P.lex_off[ d] = base >> 32;}
}

} // namespace sp
#endif
