// TEST INFRASTRUCTURE (not product code): runs the per-document state machine of
// struspattern_b200/csrc/matcher_device.cuh as a plain loop on the CPU, with the tables produced by the product's
// host rule-graph builder (matcher_program.hpp). It lets the "not gpu" test suite check the flattened tables and
// the state machine against the reference (oracle/_ref/libsp_ref_matcher.so) while no GPU is present. Nothing
// under struspattern_b200/ links or loads this file; the shipped library has no CPU path.
#include <cstdint>
#include <cstring>
#include "matcher_program.hpp"
#include "matcher_device.cuh"
#include <vector>
#include <string>

namespace {
struct MatcherSim
{
	sp::MatcherProgram prog;
	std::string lasterror;
	SpArenaCaps cap;
	std::vector<sp_result> results;
	std::vector<uint64_t> result_offsets;
	std::vector<sp_item> items;
	std::vector<uint64_t> item_offsets;
	std::vector<sp_matcher_stats> stats;
	std::vector<uint32_t> status;
	uint64_t retries;

	MatcherSim() :retries(0)
	{
		cap.rules = 256; cap.triggers = 256; cap.items = 512; cap.datarefs = 256; cap.results = 256;
		cap.follow = 64; cap.dispose = 256; cap.heap = 64;
	}
};
}

#define MSIM_TRY(...) MatcherSim* ms = (MatcherSim*)h; try {__VA_ARGS__; return 0;} \
	catch (const std::exception& e) {ms->lasterror = e.what(); return -1;}

extern "C" {

void* msim_create() {return new MatcherSim();}
void msim_destroy( void* h) {delete (MatcherSim*)h;}
const char* msim_last_error( void* h) {return ((MatcherSim*)h)->lasterror.c_str();}
int msim_define_option( void* h, const char* name, double value) {MSIM_TRY( ms->prog.defineOption( name, value))}
int msim_define_term_frequency( void* h, uint32_t termid, double df) {MSIM_TRY( ms->prog.defineTermFrequency( termid, df))}
int msim_push_term( void* h, uint32_t termid) {MSIM_TRY( ms->prog.pushTerm( termid))}
int msim_push_expression( void* h, int joinop, size_t argc, uint32_t range, uint32_t cardinality) {MSIM_TRY( ms->prog.pushExpression( joinop, argc, range, cardinality))}
int msim_push_pattern( void* h, const char* name) {MSIM_TRY( ms->prog.pushPattern( name))}
int msim_attach_variable( void* h, const char* name) {MSIM_TRY( ms->prog.attachVariable( name))}
int msim_define_pattern( void* h, const char* name, const char* fmt, int visible) {MSIM_TRY( ms->prog.definePattern( name, fmt ? fmt : "", visible != 0))}
int msim_compile( void* h) {MSIM_TRY( ms->prog.compile())}
// initial arena capacities (small on purpose: the overflow / retry path is part of what is tested)
int msim_set_caps( void* h, uint32_t rules, uint32_t triggers, uint32_t items, uint32_t datarefs, uint32_t results, uint32_t follow, uint32_t dispose, uint32_t heap)
{
	MatcherSim* ms = (MatcherSim*)h;
	ms->cap.rules = rules; ms->cap.triggers = triggers; ms->cap.items = items; ms->cap.datarefs = datarefs;
	ms->cap.results = results; ms->cap.follow = follow; ms->cap.dispose = dispose; ms->cap.heap = heap;
	return 0;
}

int msim_run_batch( void* h, const sp_lexem* tokens, const uint32_t* origsegs, const uint64_t* tok_offsets, uint32_t ndocs, int, int)
{
	MSIM_TRY({
		const sp::MatcherTablesHost& H = ms->prog.tables();
		const SpMatcherTables& T = H.view;
		ms->results.clear(); ms->items.clear(); ms->stats.clear(); ms->status.clear();
		ms->result_offsets.assign( 1, 0); ms->item_offsets.assign( 1, 0);
		SpDocState S; std::memset( &S, 0, sizeof(S));
		SpArenaCaps cap = ms->cap;
		SpArenaLayout L; sp_arena_layout( cap, T.nstopwords, L);
		std::vector<uint8_t> mem( L.bytes, 0);
		SpArena A; sp_arena_bind( L, mem.data(), A);
		for (uint32_t d=0; d<ndocs; ++d)
		{
			uint64_t nevents = 0; uint32_t nres = 0, nitems = 0;
			for (;;)
			{
				nevents = sp_match_document_seq( T, A, S, tokens, origsegs, tok_offsets[ d], tok_offsets[ d+1]);
				if (!S.status) sp_doc_count_output( T, A, S, nres, nitems);
				if (S.status < SP_MST_OVERFLOW) break;
				if (S.status == SP_MST_OVF_DEPTH) throw std::runtime_error( "nesting of result items too deep");
				// arena too small: the document runs again with four times the memory
				++ms->retries;
				cap.rules *= 4; cap.triggers *= 4; cap.items *= 4; cap.datarefs *= 4; cap.results *= 4; cap.follow *= 4; cap.dispose *= 4; cap.heap *= 4;
				sp_arena_layout( cap, T.nstopwords, L);
				mem.assign( L.bytes, 0);
				sp_arena_bind( L, mem.data(), A);
				std::memset( &S, 0, sizeof(S));
			}
			sp_matcher_stats st; std::memset( &st, 0, sizeof(st));
			if (!S.status)
			{
				std::size_t rb = ms->results.size(), ib = ms->items.size();
				ms->results.resize( rb + nres); ms->items.resize( ib + nitems);
				sp_doc_write_output( A, S, ms->results.data() + rb, ms->items.data() + ib);
				sp_doc_stats( S, nevents, st);
			}
			else
			{
				// like the oracle wrapper: counters as they were at the failing token, no event statistics
				sp_doc_stats( S, 0, st); st.nof_open_patterns = 0;
			}
			ms->stats.push_back( st);
			ms->status.push_back( S.status);
			ms->result_offsets.push_back( ms->results.size());
			ms->item_offsets.push_back( ms->items.size());
		}
	})
}
uint64_t msim_result_count( void* h) {return ((MatcherSim*)h)->results.size();}
uint64_t msim_item_count( void* h) {return ((MatcherSim*)h)->items.size();}
uint64_t msim_retries( void* h) {return ((MatcherSim*)h)->retries;}
int msim_fetch( void* h, sp_result* results, uint64_t* result_offsets, sp_item* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status)
{
	MatcherSim* ms = (MatcherSim*)h;
	if (results && !ms->results.empty()) std::memcpy( results, ms->results.data(), ms->results.size()*sizeof(sp_result));
	if (result_offsets) std::memcpy( result_offsets, ms->result_offsets.data(), ms->result_offsets.size()*sizeof(uint64_t));
	if (items && !ms->items.empty()) std::memcpy( items, ms->items.data(), ms->items.size()*sizeof(sp_item));
	if (item_offsets) std::memcpy( item_offsets, ms->item_offsets.data(), ms->item_offsets.size()*sizeof(uint64_t));
	if (stats && !ms->stats.empty()) std::memcpy( stats, ms->stats.data(), ms->stats.size()*sizeof(sp_matcher_stats));
	if (status && !ms->status.empty()) std::memcpy( status, ms->status.data(), ms->status.size()*sizeof(uint32_t));
	return 0;
}
const char* msim_pattern_name( void* h, uint32_t handle) {const char* s = ((MatcherSim*)h)->prog.patternName( handle); return s ? s : "";}
const char* msim_variable_name( void* h, uint32_t handle) {const char* s = ((MatcherSim*)h)->prog.variableName( handle); return s ? s : "";}
int64_t msim_info( void* h, const char* key)
{
	MatcherSim* ms = (MatcherSim*)h;
	std::string k( key);
	try
	{
		const sp::MatcherTablesHost& H = ms->prog.tables();
		if (k == "programs") return H.view.nprograms;
		if (k == "trigger_defs") return H.view.ntrigger_defs;
		if (k == "program_triggers") return H.view.nprogram_triggers;
		if (k == "stopwords") return H.view.nstopwords;
		if (k == "event_slots") return (int64_t)H.view.event_mask + 1;
	}
	catch (...) {}
	return -1;
}

} // extern "C"
