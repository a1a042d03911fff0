// Small CUDA plumbing shared by the lexer and matcher: error checks, grow-only device buffers, and an
// exclusive prefix sum (u32 counts -> u64 offsets) used to place variable sized per-chunk / per-document output.
#ifndef SP_DEVICE_UTIL_CUH_INCLUDED
#define SP_DEVICE_UTIL_CUH_INCLUDED
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <stdexcept>
#include <cstdio>

namespace sp {

struct CudaError :public std::runtime_error
{
	explicit CudaError( const std::string& msg) :std::runtime_error(msg){}
};

inline void cudaCheck( cudaError_t rc, const char* what, const char* file, int line)
{
	if (rc == cudaSuccess) return;
	char buf[ 512];
	snprintf( buf, sizeof(buf), "CUDA error %s (%s) at %s:%d", cudaGetErrorString( rc), what, file, line);
	throw CudaError( buf);
}
#define SP_CUDA( call) ::sp::cudaCheck( (call), #call, __FILE__, __LINE__)

// Device memory that only grows; contents are not preserved on growth.
class DeviceBuffer
{
public:
	DeviceBuffer() :m_ptr(0),m_cap(0){}
	~DeviceBuffer() {if (m_ptr) cudaFree( m_ptr);}
	DeviceBuffer( const DeviceBuffer&) = delete;
	DeviceBuffer& operator=( const DeviceBuffer&) = delete;

	void* reserve( std::size_t bytes)
	{
		if (bytes > m_cap)
		{
			if (m_ptr) {cudaFree( m_ptr); m_ptr = 0; m_cap = 0;}
			std::size_t want = bytes + bytes / 8 + 256;
			SP_CUDA( cudaMalloc( &m_ptr, want));
			m_cap = want;
		}
		return m_ptr;
	}
	template <typename T> T* as( std::size_t count) {return static_cast<T*>( reserve( count * sizeof(T) + 16));}
	void* ptr() const {return m_ptr;}
	std::size_t capacity() const {return m_cap;}
private:
	void* m_ptr;
	std::size_t m_cap;
};

// ---- scalar read back: a few words from device memory into PINNED host memory, written by a kernel (pinned memory
// is addressable from the device under unified addressing) instead of a copy command: the device-to-host copy engine
// serves the copies of all streams in the order they were issued, so a 16 byte read back of one context waited behind
// the gigabyte of results another context was fetching - tens of milliseconds per synchronisation point when several
// contexts share a device (end-to-end pipeline).
static __global__ void k_read_back( uint32_t* dst_pinned, const uint32_t* __restrict__ src, uint32_t nwords)
{
	if (threadIdx.x < nwords) dst_pinned[ threadIdx.x] = src[ threadIdx.x];
	__threadfence_system();
}
inline void readBackAsync( void* dst_pinned, const void* src_device, std::size_t bytes, cudaStream_t stream)
{
	if (bytes % 4 || bytes > 128) throw std::logic_error( "read back of scalars: size not supported");
	k_read_back<<<1, 32, 0, stream>>>( static_cast<uint32_t*>( dst_pinned), static_cast<const uint32_t*>( src_device), (uint32_t)(bytes / 4));
	SP_CUDA( cudaGetLastError());
}

// ---- exclusive prefix sum: out[i] = sum of in[0..i), out[n] = total
enum {ScanThreads = 256, ScanItems = 8, ScanTile = ScanThreads * ScanItems};

__device__ __forceinline__ uint64_t blockExclusiveScan( uint64_t v, uint64_t* warpsums, uint64_t& blocktotal)
{
	const unsigned int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
	uint64_t incl = v;
	#pragma unroll
	for (int ofs=1; ofs<32; ofs<<=1)
	{
		uint64_t up = __shfl_up_sync( 0xFFFFFFFFu, incl, ofs);
		if (lane >= (unsigned int)ofs) incl += up;
	}
	if (lane == 31) warpsums[ warp] = incl;
	__syncthreads();
	if (warp == 0)
	{
		uint64_t w = (lane < blockDim.x / 32) ? warpsums[ lane] : 0, wi = w;
		#pragma unroll
		for (int ofs=1; ofs<32; ofs<<=1)
		{
			uint64_t up = __shfl_up_sync( 0xFFFFFFFFu, wi, ofs);
			if (lane >= (unsigned int)ofs) wi += up;
		}
		warpsums[ lane] = wi - w;
		if (lane == 31) warpsums[ 32] = wi;
	}
	__syncthreads();
	uint64_t rt = warpsums[ warp] + incl - v;
	blocktotal = warpsums[ 32];
	__syncthreads();
	return rt;
}

static __global__ void __launch_bounds__(ScanThreads) k_scan_tile_sums( const uint32_t* __restrict__ in, uint64_t n, uint64_t* __restrict__ tilesum)
{
	__shared__ uint64_t warpsums[ 33];
	uint64_t base = (uint64_t)blockIdx.x * ScanTile;
	uint64_t sum = 0;
	#pragma unroll
	for (int k=0; k<ScanItems; ++k)
	{
		uint64_t i = base + (uint64_t)k * ScanThreads + threadIdx.x;
		if (i < n) sum += in[ i];
	}
	uint64_t total;
	blockExclusiveScan( sum, warpsums, total);
	if (threadIdx.x == 0) tilesum[ blockIdx.x] = total;
}

static __global__ void __launch_bounds__(1024) k_scan_tile_offsets( uint64_t* __restrict__ tilesum, uint64_t ntiles, uint64_t* __restrict__ total_out)
{
	__shared__ uint64_t warpsums[ 33];
	uint64_t carry = 0;
	for (uint64_t base=0; base<ntiles; base+=1024)
	{
		uint64_t i = base + threadIdx.x;
		uint64_t v = (i < ntiles) ? tilesum[ i] : 0, total;
		uint64_t ex = blockExclusiveScan( v, warpsums, total);
		if (i < ntiles) tilesum[ i] = carry + ex;
		carry += total;
	}
	if (threadIdx.x == 0) *total_out = carry;
}

static __global__ void __launch_bounds__(ScanThreads) k_scan_write( const uint32_t* __restrict__ in, uint64_t n, const uint64_t* __restrict__ tileofs, uint64_t* __restrict__ out)
{
	__shared__ uint64_t warpsums[ 33];
	// thread t owns ScanItems consecutive elements so that one block scan suffices
	uint64_t first = (uint64_t)blockIdx.x * ScanTile + (uint64_t)threadIdx.x * ScanItems;
	uint32_t v[ ScanItems];
	uint64_t sum = 0;
	#pragma unroll
	for (int k=0; k<ScanItems; ++k) {v[k] = (first + k < n) ? in[ first + k] : 0; sum += v[k];}
	uint64_t total;
	uint64_t ofs = tileofs[ blockIdx.x] + blockExclusiveScan( sum, warpsums, total);
	#pragma unroll
	for (int k=0; k<ScanItems; ++k)
	{
		if (first + k < n) out[ first + k] = ofs;
		ofs += v[k];
	}
}

// out must hold n+1 elements; tmp is scratch. Enqueues 3 kernels on `stream`.
inline void exclusiveScan( const uint32_t* in, uint64_t n, uint64_t* out, DeviceBuffer& tmp, cudaStream_t stream)
{
	uint64_t ntiles = (n + ScanTile - 1) / ScanTile;
	if (ntiles == 0) {SP_CUDA( cudaMemsetAsync( out, 0, sizeof(uint64_t), stream)); return;}
	uint64_t* tilesum = tmp.as<uint64_t>( ntiles);
	k_scan_tile_sums<<<(unsigned int)ntiles, ScanThreads, 0, stream>>>( in, n, tilesum);
	k_scan_tile_offsets<<<1, 1024, 0, stream>>>( tilesum, ntiles, out + n);
	k_scan_write<<<(unsigned int)ntiles, ScanThreads, 0, stream>>>( in, n, tilesum, out);
	SP_CUDA( cudaGetLastError());
}

} // namespace sp
#endif
