// Generic three phase scan over 64 bit values with an arbitrary associative operator (max, min, sum,
// segmented sum): per tile reduce -> single block scan of the tile aggregates -> per tile apply.
// Used by the lexer to find independent event islands and to assign ordinal positions in parallel.
#ifndef SP_TILE_SCAN_CUH_INCLUDED
#define SP_TILE_SCAN_CUH_INCLUDED
#include "device_util.cuh"

namespace sp {

enum {TsThreads = 256, TsItems = 4, TsTile = TsThreads * TsItems};

struct OpMax {static __device__ __forceinline__ uint64_t identity() {return 0;} static __device__ __forceinline__ uint64_t apply( uint64_t a, uint64_t b) {return a > b ? a : b;}};
struct OpMin {static __device__ __forceinline__ uint64_t identity() {return ~0ull;} static __device__ __forceinline__ uint64_t apply( uint64_t a, uint64_t b) {return a < b ? a : b;}};
struct OpSum {static __device__ __forceinline__ uint64_t identity() {return 0;} static __device__ __forceinline__ uint64_t apply( uint64_t a, uint64_t b) {return a + b;}};
// values are (segment << 32 | count); segments ascend along the array; a (earlier) is added to b only inside one segment
struct OpSegSum
{
	static __device__ __forceinline__ uint64_t identity() {return ~0ull;}	// two sided identity (no real value looks like it)
	static __device__ __forceinline__ uint64_t apply( uint64_t a, uint64_t b)
	{
		if (a == ~0ull) return b;
		if (b == ~0ull) return a;
		return ((a >> 32) == (b >> 32)) ? b + (a & 0xFFFFFFFFull) : b;
	}
};

// Inclusive scan of one value per thread over the block (thread order); `a` is the earlier operand of Op::apply.
// smem: 33 values. Returns the inclusive result, `excl` the exclusive one, `total` the block aggregate.
template <class Op>
__device__ __forceinline__ uint64_t blockScan( uint64_t v, uint64_t* smem, uint64_t& excl, uint64_t& total)
{
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
	uint64_t incl = v;
	#pragma unroll
	for (int ofs=1; ofs<32; ofs<<=1)
	{
		uint64_t up = __shfl_up_sync( 0xFFFFFFFFu, incl, ofs);
		if (lane >= (unsigned int)ofs) incl = Op::apply( up, incl);
	}
	uint64_t prev = __shfl_up_sync( 0xFFFFFFFFu, incl, 1);
	if (lane == 0) prev = Op::identity();
	if (lane == 31) smem[ warp] = incl;
	__syncthreads();
	if (warp == 0)
	{
		uint64_t w = (lane < nwarps) ? smem[ lane] : Op::identity(), wi = w;
		#pragma unroll
		for (int ofs=1; ofs<32; ofs<<=1)
		{
			uint64_t up = __shfl_up_sync( 0xFFFFFFFFu, wi, ofs);
			if (lane >= (unsigned int)ofs) wi = Op::apply( up, wi);
		}
		uint64_t wprev = __shfl_up_sync( 0xFFFFFFFFu, wi, 1);
		if (lane == 0) wprev = Op::identity();
		smem[ lane] = wprev;
		if (lane == nwarps-1) smem[ 32] = wi;
	}
	__syncthreads();
	uint64_t base = smem[ warp];
	total = smem[ 32];
	__syncthreads();
	excl = Op::apply( base, prev);
	return Op::apply( base, incl);
}

// In place exclusive scan of the tile aggregates (single block); reversed: suffix scan.
template <class Op>
static __global__ void __launch_bounds__(1024) k_tile_carry( uint64_t* agg, uint64_t ntiles, int reversed, uint64_t* total_out)
{
	__shared__ uint64_t smem[ 33];
	uint64_t carry = Op::identity();
	for (uint64_t base=0; base<ntiles; base+=1024)
	{
		uint64_t i = base + threadIdx.x;
		uint64_t idx = reversed ? ntiles - 1 - i : i;
		uint64_t v = (i < ntiles) ? agg[ idx] : Op::identity(), excl, total;
		blockScan<Op>( v, smem, excl, total);
		if (i < ntiles) agg[ idx] = Op::apply( carry, excl);
		carry = Op::apply( carry, total);
	}
	if (total_out && threadIdx.x == 0) *total_out = carry;
}

// F provides: typedef Op; __device__ uint64_t load( uint64_t i) const; __device__ void store( uint64_t i, uint64_t excl, uint64_t incl) const;
template <class F>
static __global__ void __launch_bounds__(TsThreads) k_tile_reduce( F f, uint64_t n, uint64_t* agg)
{
	typedef typename F::Op Op;
	__shared__ uint64_t smem[ 33];
	uint64_t first = (uint64_t)blockIdx.x * TsTile + (uint64_t)threadIdx.x * TsItems;
	uint64_t acc = Op::identity();
	#pragma unroll
	for (int k=0; k<TsItems; ++k) if (first + k < n) acc = Op::apply( acc, f.load( first + k));
	uint64_t excl, total;
	blockScan<Op>( acc, smem, excl, total);
	if (threadIdx.x == 0) agg[ blockIdx.x] = total;
}

template <class F>
static __global__ void __launch_bounds__(TsThreads) k_tile_apply( F f, uint64_t n, const uint64_t* carry)
{
	typedef typename F::Op Op;
	__shared__ uint64_t smem[ 33];
	uint64_t first = (uint64_t)blockIdx.x * TsTile + (uint64_t)threadIdx.x * TsItems;
	uint64_t v[ TsItems];
	uint64_t acc = Op::identity();
	#pragma unroll
	for (int k=0; k<TsItems; ++k) {v[k] = (first + k < n) ? f.load( first + k) : Op::identity(); acc = Op::apply( acc, v[k]);}
	uint64_t excl, total;
	blockScan<Op>( acc, smem, excl, total);
	uint64_t run = Op::apply( carry[ blockIdx.x], excl);
	#pragma unroll
	for (int k=0; k<TsItems; ++k)
	{
		uint64_t incl = Op::apply( run, v[k]);
		if (first + k < n) f.store( first + k, run, incl);
		run = incl;
	}
}

// Forward scan driver: 3 launches on `stream`. agg must hold ceil(n/TsTile) values.
template <class F>
inline void tileScan( const F& f, uint64_t n, uint64_t* agg, cudaStream_t stream, uint64_t* total_out=0)
{
	if (!n) return;
	uint64_t ntiles = (n + TsTile - 1) / TsTile;
	k_tile_reduce<F><<<(unsigned int)ntiles, TsThreads, 0, stream>>>( f, n, agg);
	k_tile_carry<typename F::Op><<<1, 1024, 0, stream>>>( agg, ntiles, 0, total_out);
	k_tile_apply<F><<<(unsigned int)ntiles, TsThreads, 0, stream>>>( f, n, agg);
	SP_CUDA( cudaGetLastError());
}

} // namespace sp
#endif
