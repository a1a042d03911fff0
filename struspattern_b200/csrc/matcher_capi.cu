// C ABI of the level-2 matcher (include/strus_pattern_b200.h): instance = flattened rule graph in HBM,
// context = per-warp arenas, output pools and result buffers of one in-flight batch. There is no CPU path:
// every run launches the sm_100a kernels of matcher_kernels.cuh and fails with SP_ERR_CUDA if no device is usable.
#include "strus_pattern_b200.h"
#include "matcher_program.hpp"
#include "matcher_kernels.cuh"
#include "matcher_eval_kernels.cuh"
#include "device_util.cuh"
#include <vector>
#include <algorithm>
#include <string>
#include <cstring>
#include <cstdlib>

using namespace sp;

struct sp_matcher
{
	MatcherProgram prog;
	std::string lasterror;
	bool uploaded;
	int device;
	SpMatcherTables dev;
	std::vector<void*> allocations;

	sp_matcher() :uploaded(false),device(-1) {std::memset( &dev, 0, sizeof(dev));}
	~sp_matcher() {release();}

	void release()
	{
		for (void* p: allocations) cudaFree( p);
		allocations.clear();
		uploaded = false;
	}
	template <typename T> const T* put( const std::vector<T>& v)
	{
		void* p = 0;
		std::size_t bytes = (v.size() ? v.size() : 1) * sizeof(T);
		SP_CUDA( cudaMalloc( &p, bytes + 16));
		allocations.push_back( p);
		if (!v.empty()) SP_CUDA( cudaMemcpy( p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
		return static_cast<const T*>( p);
	}
	void invalidate() {if (uploaded) release();}
	void upload()
	{
		if (uploaded) return;
		SP_CUDA( cudaGetDevice( &device));
		const MatcherTablesHost& H = prog.tables();
		dev = H.view;
		dev.programs = put( H.programs); dev.trigger_defs = put( H.trigger_defs);
		dev.program_triggers = put( H.program_triggers); dev.event_slots = put( H.event_slots);
		dev.eval_slots = put( H.eval_slots); dev.eval_follows = put( H.eval_follows); dev.eval_xlist = put( H.eval_xlist); dev.prog_flags = put( H.prog_flags);
		dev.eval_entries = put( H.eval_entries); dev.eval_vars = put( H.eval_vars);
		dev.eval_agg = put( H.eval_agg); dev.eval_indlist = put( H.eval_indlist);
		dev.eval_pairs = put( H.eval_pairs); dev.eval_pairlist = put( H.eval_pairlist);
		uploaded = true;
	}
};

struct UnpackableError :public std::runtime_error
{
	UnpackableError() :std::runtime_error( "the records of this run do not fit the packed wire format (segments, long spans, format handles on items): use sp_matcher_fetch"){}
};

static int matcherFail( std::string& lasterror, const std::exception& e)
{
	lasterror = e.what();
	if (dynamic_cast<const UnpackableError*>( &e)) return SP_ERR_UNSUPPORTED;
	if (dynamic_cast<const CudaError*>( &e)) return SP_ERR_CUDA;
	if (dynamic_cast<const std::bad_alloc*>( &e)) return SP_ERR_NOMEM;
	return SP_ERR_ARGUMENT;
}
#define SP_MATCHER_TRY( obj, ...) try {__VA_ARGS__; return SP_OK;} catch (const std::exception& e) {return matcherFail( (obj)->lasterror, e);}

extern "C" {

sp_matcher* sp_matcher_create(void) {try {return new sp_matcher();} catch (...) {return 0;}}
void sp_matcher_destroy( sp_matcher* m) {delete m;}
const char* sp_matcher_last_error( const sp_matcher* m) {return m->lasterror.c_str();}

int sp_matcher_define_option( sp_matcher* m, const char* name, double value)
	{SP_MATCHER_TRY( m, m->invalidate(); m->prog.defineOption( name, value))}
int sp_matcher_define_term_frequency( sp_matcher* m, uint32_t termid, double df)
	{SP_MATCHER_TRY( m, m->prog.defineTermFrequency( termid, df))}
int sp_matcher_push_term( sp_matcher* m, uint32_t termid)
	{SP_MATCHER_TRY( m, m->prog.pushTerm( termid))}
int sp_matcher_push_expression( sp_matcher* m, int joinop, size_t argc, uint32_t range, uint32_t cardinality)
	{SP_MATCHER_TRY( m, m->invalidate(); m->prog.pushExpression( joinop, argc, range, cardinality))}
int sp_matcher_push_pattern( sp_matcher* m, const char* name)
	{SP_MATCHER_TRY( m, m->prog.pushPattern( name))}
int sp_matcher_attach_variable( sp_matcher* m, const char* name)
	{SP_MATCHER_TRY( m, m->prog.attachVariable( name))}
int sp_matcher_define_pattern( sp_matcher* m, const char* name, const char* formatstring, int visible)
	{SP_MATCHER_TRY( m, m->invalidate(); m->prog.definePattern( name, formatstring ? formatstring : "", visible != 0))}
int sp_matcher_compile( sp_matcher* m)
	{SP_MATCHER_TRY( m, m->invalidate(); m->prog.compile())}
int sp_matcher_upload( sp_matcher* m)
	{SP_MATCHER_TRY( m, m->upload())}
const char* sp_matcher_pattern_name( const sp_matcher* m, uint32_t handle) {return m->prog.patternName( handle);}
const char* sp_matcher_variable_name( const sp_matcher* m, uint32_t handle) {return m->prog.variableName( handle);}
const char* sp_matcher_format_string( const sp_matcher* m, uint32_t handle) {return m->prog.formatString( handle);}

static int64_t valueOut( const std::string& val, char* buf, uint64_t capacity)
{
	if (buf && capacity) std::memcpy( buf, val.data(), (std::size_t)std::min<uint64_t>( capacity, val.size()));
	return (int64_t)val.size();
}
int64_t sp_matcher_result_value( const sp_matcher* m, const sp_result* result, const sp_item* doc_items, char* buf, uint64_t capacity)
{
	try
	{
		std::string val;
		if (!m->prog.resultValue( *result, doc_items, val)) return SP_NO_VALUE;
		return valueOut( val, buf, capacity);
	}
	catch (const std::exception& e) {const_cast<sp_matcher*>( m)->lasterror = e.what(); return SP_ERR_ARGUMENT;}
}
int64_t sp_matcher_item_value( const sp_matcher* m, const sp_item* item, char* buf, uint64_t capacity)
{
	try
	{
		std::string val;
		if (!m->prog.itemValue( item, val)) return SP_NO_VALUE;
		return valueOut( val, buf, capacity);
	}
	catch (const std::exception& e) {const_cast<sp_matcher*>( m)->lasterror = e.what(); return SP_ERR_ARGUMENT;}
}
int64_t sp_resolve_value( const char* value, uint64_t value_len, const uint8_t* source, uint64_t source_len, char* buf, uint64_t capacity)
{
	try
	{
		std::string val;
		resolveValue( value, (std::size_t)value_len, source, (std::size_t)source_len, val);
		return valueOut( val, buf, capacity);
	}
	catch (const std::exception&) {return SP_ERR_ARGUMENT;}
}

int64_t sp_matcher_table_info( const sp_matcher* m, const char* key)
{
	std::string k( key);
	try
	{
		const MatcherTablesHost& H = const_cast<sp_matcher*>( m)->prog.tables();
		if (k == "programs") return H.view.nprograms;
		if (k == "trigger_defs") return H.view.ntrigger_defs;
		if (k == "program_triggers") return H.view.nprogram_triggers;
		if (k == "stopwords") return H.view.nstopwords;
		if (k == "event_slots") return (int64_t)H.view.event_mask + 1;
		if (k == "max_trigger_defs") return H.view.max_trigger_defs;
		if (k == "table_bytes") return (int64_t)(H.programs.size() * sizeof(SpProgram) + H.trigger_defs.size() * sizeof(SpTriggerDef)
							+ H.program_triggers.size() * sizeof(SpProgramTrigger) + H.event_slots.size() * sizeof(SpEventSlot));
		if (k == "uploaded") return m->uploaded ? 1 : 0;
		if (k == "eval_capable") return H.view.eval_capable;
		if (k == "eval_follows") return H.view.nfollows;
	}
	catch (...) {}
	return -1;
}

} // extern "C"

#include "matcher_context.inl"

// ---- Count pass of the window evaluator without any allocation (include/strus_pattern_b200.h: sp_matcher_count_noalloc).
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it). The scratch arrays the count pass
// needs (event slot, work items, document, the three prefix sums and their tile sums, result bits, work queue) are
// carved from ONE file-scope device array of SP_NOALLOC_SCRATCH_BYTES that every GPU loading this library holds until
// exit. All calls on a device share it: calls that overlap on the GPU race; everything that must start at zero is
// cleared at the beginning of every call (the array keeps the contents of the last call).
static __device__ __align__(256) unsigned char g_noalloc_scratch[ SP_NOALLOC_SCRATCH_BYTES];

namespace {
enum {NoAllocMaxDevices = 64};
unsigned char* g_noalloc_base[ NoAllocMaxDevices];	// device address of g_noalloc_scratch per device (0 = not asked yet)
int g_noalloc_nsm[ NoAllocMaxDevices];

// In host code `g_noalloc_scratch` is only a handle: the device address comes from cudaGetSymbolAddress, once per device,
// with that device current.
unsigned char* noallocScratchBase( int dev, int& nsm)
{
	if (dev < 0 || dev >= NoAllocMaxDevices) throw std::runtime_error( "device ordinal out of range for the no-alloc entry point");
	if (!g_noalloc_base[ dev])
	{
		void* p = 0;
		SP_CUDA( cudaGetSymbolAddress( &p, g_noalloc_scratch));
		int n = 0;
		SP_CUDA( cudaDeviceGetAttribute( &n, cudaDevAttrMultiProcessorCount, dev));
		g_noalloc_nsm[ dev] = n;
		g_noalloc_base[ dev] = static_cast<unsigned char*>( p);
	}
	nsm = g_noalloc_nsm[ dev];
	return g_noalloc_base[ dev];
}

struct NoAllocCarver
{
	unsigned char* base; std::size_t used;
	explicit NoAllocCarver( unsigned char* base_) :base(base_),used(0){}
	template <typename T> T* take( std::size_t count)
	{
		T* rt = reinterpret_cast<T*>( base + used);
		used += (count * sizeof(T) + 255) & ~(std::size_t)255;
		return rt;
	}
};

// exclusiveScan (device_util.cuh) with the tile sums in caller provided memory
void noallocScan( const uint32_t* in, uint64_t n, uint64_t* out, uint64_t* tilesum, cudaStream_t stream)
{
	const uint64_t ntiles = (n + ScanTile - 1) / ScanTile;
	if (!ntiles) {SP_CUDA( cudaMemsetAsync( out, 0, sizeof(uint64_t), stream)); return;}
	k_scan_tile_sums<<<(unsigned int)ntiles, ScanThreads, 0, stream>>>( in, n, tilesum);
	k_scan_tile_offsets<<<1, 1024, 0, stream>>>( tilesum, ntiles, out + n);
	k_scan_write<<<(unsigned int)ntiles, ScanThreads, 0, stream>>>( in, n, tilesum, out);
	SP_CUDA( cudaGetLastError());
}

void matcherCountNoAlloc( sp_matcher* m, const sp_lexem* d_tokens, const uint64_t* d_tok_off, uint32_t ndocs, uint64_t ntokens,
			  uint32_t* d_doc_results, uint32_t* d_doc_items, sp_matcher_stats* d_doc_stats, uint32_t* d_doc_flags, cudaStream_t stream)
{
	if (!m->uploaded) throw std::runtime_error( "rule tables are not on the device: call sp_matcher_upload first (the no-alloc entry point does not allocate them)");
	if (!m->dev.eval_capable) throw std::runtime_error( "the rule set needs the state machine, which works on allocated arenas: no no-alloc count for it");
	if (ntokens > SP_NOALLOC_MAX_TOKENS) throw std::runtime_error( "batch beyond the limit of the no-alloc entry point: more than SP_NOALLOC_MAX_TOKENS tokens");
	if (ndocs > SP_NOALLOC_MAX_DOCS) throw std::runtime_error( "batch beyond the limit of the no-alloc entry point: more than SP_NOALLOC_MAX_DOCS documents");
	if (!ndocs) return;
	if (!d_tokens || !d_tok_off || !d_doc_results || !d_doc_items || !d_doc_stats || !d_doc_flags) throw std::runtime_error( "null pointer passed to the no-alloc entry point");
	int dev = -1, nsm = 0;
	SP_CUDA( cudaGetDevice( &dev));
	if (dev != m->device) throw std::runtime_error( "matcher tables live on another device than the current one");
	NoAllocCarver S( noallocScratchBase( dev, nsm));

	EvalParams E; std::memset( &E, 0, sizeof(E));
	E.T = m->dev; E.tokens = d_tokens; E.origsegs = 0; E.tok_off = d_tok_off; E.ndocs = ndocs; E.ntokens = ntokens;
	E.tok_slot = S.take<uint32_t>( SP_NOALLOC_MAX_TOKENS); E.tok_work = S.take<uint32_t>( SP_NOALLOC_MAX_TOKENS); E.tok_doc = S.take<uint32_t>( SP_NOALLOC_MAX_TOKENS);
	uint64_t* d_tok_work_off = S.take<uint64_t>( SP_NOALLOC_MAX_TOKENS + 1);
	E.tok_nres = S.take<uint32_t>( SP_NOALLOC_MAX_TOKENS); E.tok_nitems = S.take<uint32_t>( SP_NOALLOC_MAX_TOKENS);
	uint64_t* d_tok_res_off = S.take<uint64_t>( SP_NOALLOC_MAX_TOKENS + 1);
	uint64_t* d_tok_item_off = S.take<uint64_t>( SP_NOALLOC_MAX_TOKENS + 1);
	E.tok_work_off = d_tok_work_off; E.tok_res_off = d_tok_res_off; E.tok_item_off = d_tok_item_off;
	uint64_t* d_tilesum = S.take<uint64_t>( (SP_NOALLOC_MAX_TOKENS + ScanTile - 1) / ScanTile);
	E.resbits = S.take<uint32_t>( SP_NOALLOC_MAX_RULES / 32 + 2);
	E.queue = S.take<unsigned long long>( SP_NOALLOC_QUEUE); E.queue_cap = SP_NOALLOC_QUEUE;
	uint32_t* d_status = S.take<uint32_t>( SP_NOALLOC_MAX_DOCS);
	uint32_t* d_flag_list = S.take<uint32_t>( SP_NOALLOC_MAX_DOCS);
	uint64_t* d_scalars = S.take<uint64_t>( 4);		// [0] results [1] items [2] flagged documents [3] queued rules
	if (S.used > SP_NOALLOC_SCRATCH_BYTES) throw std::logic_error( "scratch layout of the no-alloc entry point exceeds its array");
	E.queue_count = reinterpret_cast<unsigned long long*>( d_scalars + 3);
	E.stats = d_doc_stats; E.flags = d_doc_flags;

	// everything that has to start at zero (the array holds what the last call left)
	SP_CUDA( cudaMemsetAsync( d_doc_flags, 0, (std::size_t)ndocs * sizeof(uint32_t), stream));
	SP_CUDA( cudaMemsetAsync( d_doc_stats, 0, (std::size_t)ndocs * sizeof(sp_matcher_stats), stream));
	SP_CUDA( cudaMemsetAsync( d_scalars, 0, 4 * sizeof(uint64_t), stream));
	uint64_t host[ 4] = {0,0,0,0};
	if (ntokens)
	{
		k_eval_prepare<<<(unsigned int)((ntokens + 255) / 256), 256, 0, stream>>>( E);
		noallocScan( E.tok_work, ntokens, d_tok_work_off, d_tilesum, stream);
		SP_CUDA( cudaMemcpyAsync( host, d_tok_work_off + ntokens, sizeof(uint64_t), cudaMemcpyDeviceToHost, stream));
		SP_CUDA( cudaStreamSynchronize( stream));
		const uint64_t nwork = host[ 0];
		if (nwork > SP_NOALLOC_MAX_RULES) throw std::runtime_error( "batch beyond the limit of the no-alloc entry point: more than SP_NOALLOC_MAX_RULES rule instances");
		SP_CUDA( cudaMemsetAsync( E.resbits, 0, (nwork / 32 + 2) * sizeof(uint32_t), stream));
		const uint64_t ntiles = (ntokens + EvalTile - 1) / EvalTile;
		k_eval_count_join<8><<<(unsigned int)((ntiles + EvalWarpsPerBlock - 1) / EvalWarpsPerBlock), EvalThreads, 0, stream>>>( E);
		k_eval_count_full<<<(unsigned int)nsm * 8, 128, 0, stream>>>( E);
		noallocScan( E.tok_nres, ntokens, d_tok_res_off, d_tilesum, stream);
		noallocScan( E.tok_nitems, ntokens, d_tok_item_off, d_tilesum, stream);
	}
	else
	{
		SP_CUDA( cudaMemsetAsync( d_tok_res_off, 0, sizeof(uint64_t), stream));
		SP_CUDA( cudaMemsetAsync( d_tok_item_off, 0, sizeof(uint64_t), stream));
	}
	k_eval_docs<<<(ndocs + 255) / 256, 256, 0, stream>>>( E, d_doc_results, d_doc_items, d_status, d_flag_list,
							     reinterpret_cast<uint32_t*>( d_scalars + 2), reinterpret_cast<unsigned long long*>( d_scalars));
	SP_CUDA( cudaGetLastError());
	SP_CUDA( cudaMemcpyAsync( host, d_scalars, 4 * sizeof(uint64_t), cudaMemcpyDeviceToHost, stream));
	SP_CUDA( cudaStreamSynchronize( stream));
	if (host[ 3] > SP_NOALLOC_QUEUE) throw std::runtime_error( "batch beyond the limit of the no-alloc entry point: more than SP_NOALLOC_QUEUE rules need the full fold");
}
} // namespace

extern "C" int sp_matcher_count_noalloc( sp_matcher* m, const sp_lexem* tokens_device, const uint64_t* tok_offsets_device, uint32_t ndocs, uint64_t ntokens,
					 uint32_t* doc_results_device, uint32_t* doc_items_device, sp_matcher_stats* doc_stats_device, uint32_t* doc_flags_device, void* cuda_stream)
{
	SP_MATCHER_TRY( m, matcherCountNoAlloc( m, tokens_device, tok_offsets_device, ndocs, ntokens, doc_results_device, doc_items_device, doc_stats_device, doc_flags_device, (cudaStream_t)cuda_stream))
}
