// ORACLE (test infrastructure, not product code).
// C entry points around the UNMODIFIED reference level-2 implementation
// (/root/reference/src/patternMatcher.cpp + ruleMatcherAutomaton.cpp, compiled in place by
// oracle/Makefile into oracle/_ref/libsp_ref_matcher.so). The call sequence per document is the one of
// tests/randomTokenPatternMatch/src/testRandomTokenPatternMatch.cpp:138-161
// (createContext, putInput per token, fetchResults, getStatistics); threading is the reference's own
// model (:326-345): N threads over a static partition of the documents sharing one compiled instance.
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this.
#include "patternMatcher.hpp"
#include "strus/patternMatcherInstanceInterface.hpp"
#include "strus/patternMatcherContextInterface.hpp"
#include "strus/errorBufferInterface.hpp"
#include "strus/lib/error.hpp"
#include "strus_pattern_b200.h"
#include <vector>
#include <string>
#include <thread>
#include <map>
#include <memory>
#include <cstring>
#include <cmath>

namespace {
struct RefMatcher
{
	strus::ErrorBufferInterface* errorhnd;
	strus::PatternMatcherInterface* factory;
	strus::PatternMatcherInstanceInterface* inst;
	std::string lasterror;
	// name tables filled while collecting results (name pointer -> dense index)
	std::vector<std::string> names;
	std::map<std::string,uint32_t> nameidx;
	// last batch output
	std::vector<sp_result> results;
	std::vector<uint64_t> result_offsets;
	std::vector<sp_item> items;
	std::vector<uint64_t> item_offsets;
	std::vector<sp_matcher_stats> stats;
	std::vector<uint32_t> status;
	// value strings (PatternMatcherResult::value / PatternMatcherResultItem::value) parallel to results / items; NULL values have no entry in *_hasvalue
	std::vector<std::string> result_values, item_values;
	std::vector<char> result_hasvalue, item_hasvalue;

	RefMatcher() :errorhnd(0),factory(0),inst(0){}
	~RefMatcher() {delete inst; delete factory; delete errorhnd;}
	bool check()
	{
		if (errorhnd->hasError()) {lasterror = errorhnd->fetchError(); return false;}
		return true;
	}
	uint32_t nameIndex( const char* nam)
	{
		std::string key( nam ? nam : "");
		std::map<std::string,uint32_t>::const_iterator it = nameidx.find( key);
		if (it != nameidx.end()) return it->second;
		names.push_back( key);
		return nameidx[ key] = (uint32_t)names.size();
	}
};

struct DocOut
{
	std::vector<sp_result> results;
	std::vector<sp_item> items;
	std::vector<std::string> resnames;	// parallel to results
	std::vector<std::string> itemnames;	// parallel to items
	std::vector<std::string> resvalues, itemvalues;
	std::vector<char> reshasvalue, itemhasvalue;
	sp_matcher_stats stats;
	uint32_t status;
};
}

extern "C" {

void* refm_create()
{
	RefMatcher* rm = new RefMatcher();
	rm->errorhnd = strus::createErrorBuffer_standard( 0, 64, 0);
	rm->factory = new strus::PatternMatcher( rm->errorhnd);
	rm->inst = rm->factory->createInstance();
	return rm;
}
void refm_destroy( void* h) {delete (RefMatcher*)h;}
const char* refm_last_error( void* h) {return ((RefMatcher*)h)->lasterror.c_str();}

int refm_define_option( void* h, const char* name, double value)
	{RefMatcher* rm=(RefMatcher*)h; rm->inst->defineOption( name, value); return rm->check()?0:-1;}
int refm_define_term_frequency( void* h, uint32_t termid, double df)
{
	RefMatcher* rm=(RefMatcher*)h;
	try {rm->inst->defineTermFrequency( termid, df);}	// not wrapped by the reference (:361-364)
	catch (const std::exception& e) {rm->lasterror = e.what(); return -1;}
	return rm->check()?0:-1;
}
int refm_push_term( void* h, uint32_t termid)
	{RefMatcher* rm=(RefMatcher*)h; rm->inst->pushTerm( termid); return rm->check()?0:-1;}
int refm_push_expression( void* h, int joinop, size_t argc, uint32_t range, uint32_t cardinality)
{
	RefMatcher* rm=(RefMatcher*)h;
	rm->inst->pushExpression( (strus::PatternMatcherInstanceInterface::JoinOperation)joinop, argc, range, cardinality);
	return rm->check()?0:-1;
}
int refm_push_pattern( void* h, const char* name)
	{RefMatcher* rm=(RefMatcher*)h; rm->inst->pushPattern( name); return rm->check()?0:-1;}
int refm_attach_variable( void* h, const char* name)
	{RefMatcher* rm=(RefMatcher*)h; rm->inst->attachVariable( name); return rm->check()?0:-1;}
int refm_define_pattern( void* h, const char* name, const char* formatstring, int visible)
	{RefMatcher* rm=(RefMatcher*)h; rm->inst->definePattern( name, formatstring?formatstring:"", visible!=0); return rm->check()?0:-1;}
int refm_compile( void* h)
	{RefMatcher* rm=(RefMatcher*)h; bool ok = rm->inst->compile(); return (rm->check() && ok)?0:-1;}

static void runDocument( const strus::PatternMatcherInstanceInterface* inst, strus::ErrorBufferInterface* errorhnd,
			 const sp_lexem* tokens, const uint32_t* origsegs, uint64_t tb, uint64_t te, bool with_results, DocOut& out)
{
	std::unique_ptr<strus::PatternMatcherContextInterface> ctx( inst->createContext());
	out.status = 0;
	std::memset( &out.stats, 0, sizeof(out.stats));
	for (uint64_t ti=tb; ti<te; ++ti)
	{
		const sp_lexem& t = tokens[ ti];
		ctx->putInput( strus::analyzer::PatternLexem( t.id, t.ordpos, strus::analyzer::Position( origsegs?origsegs[ti]:0, t.origpos), t.origsize));
		if (errorhnd->hasError()) {(void)errorhnd->fetchError(); out.status = (uint32_t)-SP_ERR_RUNTIME; break;}
	}
	std::vector<strus::analyzer::PatternMatcherResult> results = ctx->fetchResults();
	if (errorhnd->hasError()) {(void)errorhnd->fetchError(); out.status = (uint32_t)-SP_ERR_RUNTIME;}
	strus::analyzer::PatternMatcherStatistics st = ctx->getStatistics();
	std::vector<strus::analyzer::PatternMatcherStatistics::Item>::const_iterator si = st.items().begin(), se = st.items().end();
	double avgActive = 0.0;
	for (; si != se; ++si)
	{
		if (0==std::strcmp( si->name(), "nofProgramsInstalled")) out.stats.nof_programs_installed = (uint64_t)si->value();
		else if (0==std::strcmp( si->name(), "nofAltKeyProgramsInstalled")) out.stats.nof_altkey_programs_installed = (uint64_t)si->value();
		else if (0==std::strcmp( si->name(), "nofSignalsFired")) out.stats.nof_signals_fired = (uint64_t)si->value();
		else if (0==std::strcmp( si->name(), "nofTriggersAvgActive")) avgActive = si->value();
	}
	// nofTriggersAvgActive = nofOpenPatterns / nofEvents (src/patternMatcher.cpp:311-314); recover the integer sum
	uint64_t nofEvents = (out.status == 0) ? (te - tb) : 0;
	if (out.status == 0)
	{
		out.stats.nof_events = nofEvents;
		out.stats.nof_open_patterns = (uint64_t)std::llround( avgActive * (double)nofEvents);
	}
	if (!with_results) {out.stats.nof_events |= 0; out.results.resize( results.size()); return;}
	std::vector<strus::analyzer::PatternMatcherResult>::const_iterator ri = results.begin(), re = results.end();
	for (; ri != re; ++ri)
	{
		sp_result r;
		r.name = 0; r.format = 0;
		r.ordpos = ri->ordpos(); r.ordend = ri->ordend();
		r.origseg = ri->origpos().seg(); r.origpos = ri->origpos().ofs();
		r.endseg = ri->origend().seg(); r.endpos = ri->origend().ofs();
		r.item_begin = (uint32_t)out.items.size(); r.item_count = (uint32_t)ri->items().size();
		out.results.push_back( r);
		out.resnames.push_back( ri->name() ? ri->name() : "");
		out.resvalues.push_back( ri->value() ? ri->value() : ""); out.reshasvalue.push_back( ri->value() ? 1 : 0);
		std::vector<strus::analyzer::PatternMatcherResultItem>::const_iterator ii = ri->items().begin(), ie = ri->items().end();
		for (; ii != ie; ++ii)
		{
			sp_item it;
			it.variable = 0; it.format = 0;
			it.ordpos = ii->ordpos(); it.ordend = ii->ordend();
			it.origseg = ii->origpos().seg(); it.origpos = ii->origpos().ofs();
			it.endseg = ii->origend().seg(); it.endpos = ii->origend().ofs();
			out.items.push_back( it);
			out.itemnames.push_back( ii->name() ? ii->name() : "");
			out.itemvalues.push_back( ii->value() ? ii->value() : ""); out.itemhasvalue.push_back( ii->value() ? 1 : 0);
		}
	}
}

// Run a batch. with_results=0: only counts and statistics are kept (timing runs).
int refm_run_batch( void* h, const sp_lexem* tokens, const uint32_t* origsegs, const uint64_t* tok_offsets, uint32_t ndocs, int nthreads, int with_results)
{
	RefMatcher* rm=(RefMatcher*)h;
	if (nthreads < 1) nthreads = 1;
	std::vector<DocOut> outs( ndocs);
	std::vector<std::thread> threads;
	const strus::PatternMatcherInstanceInterface* inst = rm->inst;
	strus::ErrorBufferInterface* errorhnd = rm->errorhnd;
	for (int t=0; t<nthreads; ++t)
	{
		uint32_t db = (uint32_t)((uint64_t)ndocs * t / nthreads), de = (uint32_t)((uint64_t)ndocs * (t+1) / nthreads);
		threads.push_back( std::thread( [=,&outs]() {
			for (uint32_t d=db; d<de; ++d) runDocument( inst, errorhnd, tokens, origsegs, tok_offsets[d], tok_offsets[d+1], with_results!=0, outs[d]);
		}));
	}
	for (std::size_t t=0; t<threads.size(); ++t) threads[t].join();
	rm->results.clear(); rm->items.clear();
	rm->result_offsets.assign( 1, 0); rm->item_offsets.assign( 1, 0);
	rm->stats.clear(); rm->status.clear();
	rm->result_values.clear(); rm->item_values.clear(); rm->result_hasvalue.clear(); rm->item_hasvalue.clear();
	for (uint32_t d=0; d<ndocs; ++d)
	{
		DocOut& o = outs[d];
		if (with_results)
		{
			for (std::size_t i=0; i<o.results.size(); ++i) {o.results[i].name = rm->nameIndex( o.resnames[i].c_str()); rm->results.push_back( o.results[i]);}
			for (std::size_t i=0; i<o.items.size(); ++i) {o.items[i].variable = rm->nameIndex( o.itemnames[i].c_str()); rm->items.push_back( o.items[i]);}
			rm->result_offsets.push_back( rm->results.size());
			rm->result_values.insert( rm->result_values.end(), o.resvalues.begin(), o.resvalues.end());
			rm->result_hasvalue.insert( rm->result_hasvalue.end(), o.reshasvalue.begin(), o.reshasvalue.end());
			rm->item_values.insert( rm->item_values.end(), o.itemvalues.begin(), o.itemvalues.end());
			rm->item_hasvalue.insert( rm->item_hasvalue.end(), o.itemhasvalue.begin(), o.itemhasvalue.end());
		}
		else
		{
			rm->result_offsets.push_back( rm->result_offsets.back() + o.results.size());
		}
		rm->item_offsets.push_back( rm->items.size());
		rm->stats.push_back( o.stats);
		rm->status.push_back( o.status);
	}
	return 0;
}
uint64_t refm_result_count( void* h) {return ((RefMatcher*)h)->result_offsets.empty()?0:((RefMatcher*)h)->result_offsets.back();}
uint64_t refm_item_count( void* h) {return ((RefMatcher*)h)->items.size();}
int refm_fetch( void* h, sp_result* results, uint64_t* result_offsets, sp_item* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status)
{
	RefMatcher* rm=(RefMatcher*)h;
	if (results && !rm->results.empty()) std::memcpy( results, rm->results.data(), rm->results.size()*sizeof(sp_result));
	if (result_offsets) std::memcpy( result_offsets, rm->result_offsets.data(), rm->result_offsets.size()*sizeof(uint64_t));
	if (items && !rm->items.empty()) std::memcpy( items, rm->items.data(), rm->items.size()*sizeof(sp_item));
	if (item_offsets) std::memcpy( item_offsets, rm->item_offsets.data(), rm->item_offsets.size()*sizeof(uint64_t));
	if (stats && !rm->stats.empty()) std::memcpy( stats, rm->stats.data(), rm->stats.size()*sizeof(sp_matcher_stats));
	if (status && !rm->status.empty()) std::memcpy( status, rm->status.data(), rm->status.size()*sizeof(uint32_t));
	return 0;
}
// Value string of result / item `idx` (index into the fetched arrays) of the last batch: length (the first `capacity` bytes
// are written), SP_NO_VALUE if the reference delivered NULL.
static int64_t valueOut( const std::vector<std::string>& values, const std::vector<char>& has, uint64_t idx, char* buf, uint64_t capacity)
{
	if (idx >= values.size()) return SP_ERR_ARGUMENT;
	if (!has[ idx]) return SP_NO_VALUE;
	const std::string& v = values[ idx];
	if (buf && capacity) std::memcpy( buf, v.data(), (std::size_t)(capacity < v.size() ? capacity : v.size()));
	return (int64_t)v.size();
}
int64_t refm_result_value( void* h, uint64_t idx, char* buf, uint64_t capacity)
	{RefMatcher* rm=(RefMatcher*)h; return valueOut( rm->result_values, rm->result_hasvalue, idx, buf, capacity);}
int64_t refm_item_value( void* h, uint64_t idx, char* buf, uint64_t capacity)
	{RefMatcher* rm=(RefMatcher*)h; return valueOut( rm->item_values, rm->item_hasvalue, idx, buf, capacity);}

// Name of a result (sp_result.name) or item (sp_item.variable) index of this oracle's string table.
const char* refm_name( void* h, uint32_t idx)
{
	RefMatcher* rm=(RefMatcher*)h;
	return (idx == 0 || idx > rm->names.size()) ? "" : rm->names[ idx-1].c_str();
}

} // extern "C"
