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
//   k_lex_resolve  one thread per document: the reference's order dependent supersede logic
//                  (src/patternLexer.cpp:754-822) over the document's events, then counts the lexems.
//   k_lex_write    one thread per document: ordinal position assignment (src/patternLexer.cpp:893-945) into the
//                  prefix-summed output range.
#ifndef SP_LEXER_KERNELS_CUH_INCLUDED
#define SP_LEXER_KERNELS_CUH_INCLUDED
#include "lexer_device.cuh"
#include "lexer_fused.cuh"
#include "tile_scan.cuh"

namespace sp {

enum {LexScanThreads = 128, LexEventThreads = 256, LexDocThreads = 64};

static __global__ void __launch_bounds__(LexScanThreads) k_lex_count( SpScanParams P, int round)
{
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (j < P.nchunks) spk_count_chunk( P, j, round);
}

static __global__ void __launch_bounds__(256) k_lex_verify( SpScanParams P)
{
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	int bad = (j < P.nchunks) ? spk_verify_chunk( P, j) : 0;
	unsigned int ballot = __ballot_sync( 0xFFFFFFFFu, bad);
	if ((threadIdx.x & 31) == 0 && ballot) atomicAdd( P.nbad, (uint32_t)__popc( ballot));
}

static __global__ void __launch_bounds__(LexScanThreads) k_lex_emit( SpScanParams P)
{
	uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (j < P.nchunks) spk_emit_chunk( P, j);
}

static __global__ void __launch_bounds__(LexEventThreads) k_lex_finish( SpScanParams P, uint64_t nev)
{
	uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if (i < nev) spk_finish_event( P, i);
}

static __global__ void __launch_bounds__(LexDocThreads) k_lex_resolve( SpScanParams P)
{
	uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
	if (d < P.ndocs) spk_resolve_doc( P, d);
}

static __global__ void __launch_bounds__(LexDocThreads) k_lex_write( SpScanParams P)
{
	uint32_t d = blockIdx.x * blockDim.x + threadIdx.x;
	if (d < P.ndocs) spk_write_doc( P, d);
}

// ---- fused segment pipeline (lexer_fused.cuh)
//   k_lex_fused    one thread per text chunk: forward scan from the verified start state, events completed, resolved and
//                  flagged in thread local memory, kept elements written densely into the chunk's slot range
//   k_fused_fix    leading BindPredecessor elements of a range: look back for the answer
//   AggScan        exclusive scan of the per chunk (advanced | emitted << 32) pairs
//   k_fused_docs   per document scan base, lexem offsets, late checks
//   k_fused_write  one warp per chunk: ordinal positions and output indices by ballot scans, 16 byte lexem records
enum {LexFusedThreads = 128, LexWriteThreads = 256};

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
