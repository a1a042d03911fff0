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

} // namespace sp
#endif
