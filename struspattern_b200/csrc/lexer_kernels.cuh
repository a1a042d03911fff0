// CUDA kernels of the level-1 lexer (sm_100a). Every kernel is a thin grid-stride-free wrapper around a
// per-thread body of lexer_device.cuh; see DESIGN.md "lexer pipeline" for the data flow and the HBM layout.
//
//   k_lex_count    one thread per text chunk: forward DFA scan, counts raw events, records the end state.
//                  Round 0 guesses the start state by scanning `warmup` bytes ahead of the chunk.
//   k_lex_verify   compares every guessed start state with the end state of the chunk before; mismatches are
//                  flagged for another k_lex_count round (host loop until none is left).
//   k_lex_emit     same scan from the verified start states, writes (definition, end offset) events at the
//                  prefix-summed positions -> events are globally ordered by (document, end offset, definition).
//   k_lex_finish   one thread per raw event: reverse DFA walk for the leftmost start (HS_FLAG_SOM_LEFTMOST),
//                  size check, sub expression selection, symbol lookup.
//   (the per-document form of the second half, spk_resolve_doc / spk_write_doc of lexer_device.cuh, is the definition the
//   parallel formulations are tested against on the CPU, tests/hostsim; it has no kernel of its own any more)
#ifndef SP_LEXER_KERNELS_CUH_INCLUDED
#define SP_LEXER_KERNELS_CUH_INCLUDED
#include "lexer_device.cuh"
#include "lexer_fused.cuh"
#include "tile_scan.cuh"
#include <cuda/barrier>

namespace sp {

enum {LexScanThreads = 128, LexEventThreads = 128, LexDocThreads = 64};

// the byte -> column map of the forward automaton is read once per text byte with a different index in every lane:
// as a global load that is a 32 address gather per byte; a copy in shared memory makes it a (mostly conflict free) LDS
__device__ __forceinline__ void stageByteClasses( SpScanParams& P, uint8_t* s_class)
{
	for (unsigned int i=threadIdx.x; i<64; i+=blockDim.x) reinterpret_cast<uint32_t*>( s_class)[ i] = reinterpret_cast<const uint32_t*>( P.T.fwd_class)[ i];
	__syncthreads();
	P.T.fwd_class = s_class;
}

// TMA-staged text tiles (north_star kernel 1): the 128 chunks of a block are 64 KiB of consecutive text. Every thread asks the
// bulk copy engine for its own 512 byte chunk (cp.async.bulk global -> shared, completion counted in bytes on one mbarrier)
// into a row of LexTilePitch = 528 bytes: with that pitch the 16 byte vector reads of the lanes of a quarter warp fall into
// different banks (row r starts at bank 4r mod 32), where a pitch of 512 would put all 32 lanes on the same banks. The scan
// then reads its chunk from shared memory through a pointer that is addressed like the text itself (tile row - begin).
// A chunk that is not a full, 16 byte aligned 512 bytes (the last one of the batch) is read from global memory as before.
enum {LexTilePitch = 528, LexTileBytes = LexScanThreads * LexTilePitch};

__device__ __forceinline__ const uint8_t* stageTextTile( const SpScanParams& P, uint64_t j, uint8_t* s_tile, cuda::barrier<cuda::thread_scope_block>& bar)
{
	const uint64_t begin = j * P.chunk;
	const bool staged = j < P.nchunks && P.chunk == 512 && begin + 512 <= P.total;
	if (threadIdx.x == 0)
	{
		init( &bar, blockDim.x);
		cuda::device::experimental::fence_proxy_async_shared_cta();
	}
	__syncthreads();
	uint8_t* row = s_tile + (size_t)threadIdx.x * LexTilePitch;
	cuda::barrier<cuda::thread_scope_block>::arrival_token token = staged
		? (cuda::device::memcpy_async_tx( row, P.text + begin, cuda::aligned_size_t<16>( 512), bar), cuda::device::barrier_arrive_tx( bar, 1, 512))
		: bar.arrive();
	bar.wait( std::move( token));
	return staged ? row - begin : P.text;
}

template <bool TMA>
static __global__ void __launch_bounds__(LexScanThreads) k_lex_count( SpScanParams P, int round)
{
	__shared__ __align__(16) uint8_t s_class[ 256];
	extern __shared__ __align__(128) uint8_t s_tile[];
	__shared__ cuda::barrier<cuda::thread_scope_block> bar;
	stageByteClasses( P, s_class);
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint8_t* scan_text = P.text;
	if (TMA) scan_text = stageTextTile( P, j, s_tile, bar);
	if (j < P.nchunks) spk_count_chunk( P, j, round, scan_text);
}

static __global__ void __launch_bounds__(256) k_lex_verify( SpScanParams P)
{
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	int bad = (j < P.nchunks) ? spk_verify_chunk( P, j) : 0;
	unsigned int ballot = __ballot_sync( 0xFFFFFFFFu, bad);
	if ((threadIdx.x & 31) == 0 && ballot) atomicAdd( P.nbad, (uint32_t)__popc( ballot));
}

template <bool TMA>
static __global__ void __launch_bounds__(LexScanThreads) k_lex_emit( SpScanParams P)
{
	__shared__ __align__(16) uint8_t s_class[ 256];
	extern __shared__ __align__(128) uint8_t s_tile[];
	__shared__ cuda::barrier<cuda::thread_scope_block> bar;
	stageByteClasses( P, s_class);
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const uint8_t* scan_text = P.text;
	if (TMA) scan_text = stageTextTile( P, j, s_tile, bar);
	if (j < P.nchunks) spk_emit_chunk( P, j, scan_text);
}

static __global__ void __launch_bounds__(LexEventThreads) k_lex_finish( SpScanParams P, uint64_t nev)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nev) spk_finish_event( P, i);
}

// ---- fused segment pipeline (lexer_fused.cuh)
//   k_lex_fused    one thread per text chunk: forward scan from the verified start state, events completed, resolved and
//                  flagged in thread local memory, kept elements written densely into the chunk's slot range
//   k_fused_fix    leading BindPredecessor elements of a range: look back for the answer
//   AggScan        exclusive scan of the per chunk (advanced | emitted << 32) pairs
//   k_fused_docs   per document scan base, lexem offsets, late checks
//   k_fused_write  one warp per chunk: ordinal positions and output indices by ballot scans, 16 byte lexem records
enum {LexFusedThreads = 128, LexWriteThreads = 128};

enum {LexFusedSmem = (SP_FUSED_AR * 16 + SP_FUSED_MEMO * 16 + SP_FUSED_BUF * 4) * LexFusedThreads};

static __global__ void __launch_bounds__(LexFusedThreads) k_lex_fused( SpFusedParams P)
{
	// work memory of the threads in shared memory, element i of thread t at [i*LexFusedThreads + t] (no bank conflicts):
	// open segment elements, reverse walk memos, queued entries
	extern __shared__ uint4 s_dyn[];
	SpMatchEvent* s_ar = reinterpret_cast<SpMatchEvent*>( s_dyn);
	SpStartMemo* s_memo = reinterpret_cast<SpStartMemo*>( s_dyn + SP_FUSED_AR * LexFusedThreads);
	uint32_t* s_buf = reinterpret_cast<uint32_t*>( s_dyn + (SP_FUSED_AR + SP_FUSED_MEMO) * LexFusedThreads);
	const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	SpStridedEvents ar; ar.base = s_ar + threadIdx.x; ar.stride = LexFusedThreads;
	if (j < P.nchunks) spk_fused_chunk( P, j, s_buf + threadIdx.x, LexFusedThreads, ar, s_memo + threadIdx.x);
}

// chunk resolution over completed events (second half of the fused formulation alone): one thread per chunk, open
// segment elements in shared memory, element i of thread t at [i*LexFusedThreads + t]
enum {LexResolveTile = 8};	// events per lane and refill (a power of two up to 32)

struct TileSource	/* the lane's column of the warp's event tile: element k at [k*33] */
{
	const uint4* ev_tile; const uint32_t* doc_tile;
	__device__ __forceinline__ SpRawEvent ev( uint32_t k) const
	{
		const uint4 v = ev_tile[ k * 33];
		SpRawEvent e; e.pat = v.x; e.from = v.y; e.to = v.z; e.aux = v.w;
		return e;
	}
	__device__ __forceinline__ uint32_t df( uint32_t k) const {return doc_tile[ k * 33];}
};

static __global__ void __launch_bounds__(LexFusedThreads, 8) k_lex_resolve_chunks( SpFusedParams P)
{
	__shared__ uint4 s_ar[ SP_FUSED_AR * LexFusedThreads];
	// event tiles of the warps, [warp][k][lane]: every lane consumes its own stream of events, which lies 2-3 KB away from
	// its neighbour's; the warp fetches LexResolveTile events of every stream with coalesced loads (32/LexResolveTile streams
	// x LexResolveTile events per instruction) instead of 32 scattered 16 byte loads per step
	__shared__ uint4 s_ev[ LexFusedThreads/32][ LexResolveTile][ 33];
	__shared__ uint32_t s_doc[ LexFusedThreads/32][ LexResolveTile][ 33];
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	SpStridedEvents ar; ar.base = reinterpret_cast<SpMatchEvent*>( s_ar) + threadIdx.x; ar.stride = LexFusedThreads;
	SpFusedThread st; SpChunkRange R;
	const bool active = j < P.nchunks;
	if (active) sp_resolve_begin( P, j, st, R, ar); else {R.a = R.b = 0;}
	uint64_t pos = R.a;
	for (;;)
	{
		if (!__any_sync( 0xFFFFFFFFu, pos < R.b)) break;
		const uint32_t plo = (uint32_t)pos, phi = (uint32_t)(pos >> 32);
		const uint32_t left = (uint32_t)((R.b - pos) < LexResolveTile ? (R.b - pos) : LexResolveTile);
		#pragma unroll
		for (int r=0; r<LexResolveTile; ++r)
		{
			const unsigned int s = r * (32/LexResolveTile) + (lane % (32/LexResolveTile)), k = lane / (32/LexResolveTile);
			const uint64_t p = ((uint64_t)__shfl_sync( 0xFFFFFFFFu, phi, s) << 32) | __shfl_sync( 0xFFFFFFFFu, plo, s);
			const uint32_t n = __shfl_sync( 0xFFFFFFFFu, left, s);
			if (k < n)
			{
				s_ev[ warp][ k][ s] = *reinterpret_cast<const uint4*>( P.events + p + k);
				s_doc[ warp][ k][ s] = P.ev_doc[ p + k];
			}
		}
		__syncwarp();
		if (left)
		{
			TileSource src; src.ev_tile = &s_ev[ warp][ 0][ lane]; src.doc_tile = &s_doc[ warp][ 0][ lane];
			sp_resolve_view( P, j, st, R, src, left, pos + left == R.b);
		}
		pos += left;
		__syncwarp();
	}
	if (active) sp_resolve_end( P, j, st, R);
}

// island pipeline behind a chunk resolution that gave up: plain document ids again
static __global__ void __launch_bounds__(256) k_evdoc_clear( uint32_t* ev_doc, uint64_t nev)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nev) ev_doc[ i] &= ~SP_EVDOC_CUT;
}

static __global__ void __launch_bounds__(256) k_fused_fix( SpFusedParams P)
{
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (j < P.nchunks) spk_fused_fix_chunk( P, j);
}

struct AggScan
{
	typedef OpSum Op;
	const uint64_t* agg; uint64_t* prefix;
	__device__ uint64_t load( uint64_t i) const {return agg[ i];}
	__device__ void store( uint64_t i, uint64_t excl, uint64_t) const {prefix[ i] = excl;}
};

static __global__ void __launch_bounds__(256) k_fused_docs( SpFusedParams P, const uint64_t* total)
{
	uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
	if (d <= P.ndocs) spk_fused_doc( P, d, *total);
}

static __global__ void __launch_bounds__(LexWriteThreads) k_fused_write( SpFusedParams P)
{
	const unsigned int lane = threadIdx.x & 31;
	const uint64_t j = (uint64_t)blockIdx.x * (LexWriteThreads/32) + (threadIdx.x >> 5);
	if (j >= P.nchunks) return;
	const uint32_t n = P.kept[ j];
	if (!n) return;
	const uint64_t rb = P.region_begin[ j], pre = P.prefix[ j];
	uint32_t incp = (uint32_t)pre;
	uint64_t emitp = pre >> 32;
	const unsigned int lt = (1u << lane) - 1u, le = lt | (1u << lane);
	for (uint32_t k0=0; k0<n; k0+=32)
	{
		const uint32_t k = k0 + lane;
		const bool valid = k < n;
		SpSlot s; s.id = 0; s.origpos = 0; s.info = 0; s.doc = 0;
		if (valid)
		{
			const uint4 v = *reinterpret_cast<const uint4*>( P.slots + rb + k);
			s.id = v.x; s.origpos = v.y; s.info = v.z; s.doc = v.w;
		}
		const unsigned int binc = __ballot_sync( 0xFFFFFFFFu, valid && (s.info & SP_SLOT_INC));
		sp_lexem lx;
		bool emit = false;
		if (valid) emit = sp_fused_lexem( s, incp + __popc( binc & le) - (uint32_t)__ldg( P.doc_base + s.doc), lx);
		const unsigned int bemit = __ballot_sync( 0xFFFFFFFFu, emit);
		if (emit) *reinterpret_cast<uint4*>( P.lexems + emitp + __popc( bemit & lt)) = make_uint4( lx.id, lx.ordpos, lx.origpos, lx.origsize);
		incp += __popc( binc); emitp += __popc( bemit);
	}
}

} // namespace sp
#endif
