// C ABI of the level-2 matcher (include/strus_pattern_b200.h): instance = flattened rule graph in HBM,
// context = per-warp arenas, output pools and result buffers of one in-flight batch. There is no CPU path:
// every run launches the sm_100a kernels of matcher_kernels.cuh and fails with SP_ERR_CUDA if no device is usable.
#include "strus_pattern_b200.h"
#include "matcher_program.hpp"
#include "matcher_kernels.cuh"
#include "matcher_eval_kernels.cuh"
#include "device_util.cuh"
#include <vector>
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
		uploaded = true;
	}
};

static int matcherFail( std::string& lasterror, const std::exception& e)
{
	lasterror = e.what();
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
