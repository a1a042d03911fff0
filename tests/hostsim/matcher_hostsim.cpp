// TEST INFRASTRUCTURE (not product code): runs the per-document state machine of
// struspattern_b200/csrc/matcher_device.cuh as a plain loop on the CPU, with the tables produced by the product's
// host rule-graph builder (matcher_program.hpp). It lets the "not gpu" test suite check the flattened tables and
// the state machine against the reference (oracle/_ref/libsp_ref_matcher.so) while no GPU is present. Nothing
// under struspattern_b200/ links or loads this file; the shipped library has no CPU path.
#include <cstdint>
#include <cstring>
#include "matcher_program.hpp"
#include "matcher_device.cuh"
#include "matcher_eval.cuh"
#include <vector>
#include <string>
#include <cstdio>
#include <cstdlib>

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
	int mode;			// 0 = state machine, 1 = window evaluator (documents it flags run on the state machine)
	uint64_t eval_docs, eval_flagged;

	MatcherSim() :retries(0),mode(0),eval_docs(0),eval_flagged(0)
	{
		cap.rules = 256; cap.triggers = 256; cap.items = 512; cap.datarefs = 256; cap.results = 256;
		cap.follow = 64; cap.dispose = 256; cap.heap = 64;
	}
};
}

static uint64_t g_lite_written = 0, g_results_written = 0;

// One document through the window evaluator (matcher_eval.cuh), sites in the order of the CUDA kernels: token by token,
// the token event first, then its static follow events, every program list in list order. Returns false if the
// document has to run on the state machine.
static bool evalDocument( const SpMatcherTables& T, const sp_lexem* tokens, const uint32_t* segs, uint64_t tb, uint64_t te,
			  std::vector<sp_result>& results, std::vector<sp_item>& items, sp_matcher_stats& st)
{
	std::memset( &st, 0, sizeof(st));
	for (uint64_t i=tb; i<te; ++i)
	{
		if (sp_eval_token_irregular( tokens[ i], segs ? segs[ i] : 0, i == tb, i == tb ? 0 : tokens[ i-1].ordpos)) return false;
	}
	const std::size_t item_base = items.size();
	for (uint64_t i=tb; i<te; ++i)
	{
		const uint32_t slotidx = sp_find_event_index( T, tokens[ i].id);
		if (slotidx == SP_EVAL_NOSLOT) continue;
		SpEvalSite base;
		sp_eval_site_token( T, tokens[ i], segs ? segs[ i] : 0, i, slotidx, base);
		for (uint32_t c=0; c<=base.follow_count; ++c)
		{
			SpEvalSite site = base;
			uint32_t listslot = slotidx;
			if (c > 0)
			{
				const SpEvalFollow& F = T.eval_follows[ base.follow_begin + c - 1];
				if (F.eslot == SP_EVAL_NOSLOT) continue;
				sp_eval_site_follow( base, c - 1, F, site);
				listslot = F.eslot;
			}
			const SpEventSlot& es = T.event_slots[ listslot];
			for (uint32_t k=0; k<es.prog_count; ++k)
			{
				const SpProgramTrigger pt = T.program_triggers[ es.prog_begin + k];
				SpEvalRule R;
				SpEvalTake takes[ SP_EVAL_MAX_TAKES];
				sp_eval_fold<true,0>( T, tokens, segs, tb, te, site, pt, false, R, takes);
				if (R.flag)
				{
					if (getenv( "SP_EVAL_DEBUG"))
					{
						const SpProgram& P = T.programs[ pt.program];
						fprintf( stderr, "flag: doc token %llu cpos %u program %u past %u range %u initsigval %u count %u defs:", (unsigned long long)(i-tb), site.cpos, pt.program, pt.past_event, P.range, P.initsigval, P.initcount);
						for (uint32_t d=0; d<P.trig_count; ++d) {const SpTriggerDef& f = T.trigger_defs[ P.trig_begin+d]; fprintf( stderr, " (ev %u sig %u val %u key %u var %u)", f.event, f.sigtype, f.sigval, f.is_key, f.variable);}
						fprintf( stderr, "\n");
					}
					return false;
				}
				// the counting instance must agree with the collecting one (the kernels run them in two passes)
				SpEvalRule C;
				sp_eval_fold<false,0>( T, tokens, segs, tb, te, site, pt, false, C, 0);
				// the compiled list entry must give the same answer whenever it claims to know it
				SpEvalRule Qt;
				if (sp_eval_entry_run( T, T.eval_entries[ es.prog_begin + k], tokens, te, site, Qt)
				    && (Qt.installed != R.installed || Qt.signals != R.signals || Qt.open != R.open || Qt.has_result != R.has_result
					|| (R.has_result && sp_eval_result_items( Qt) != sp_eval_result_items( R)) || R.altkey || R.flag))
				{
					throw std::runtime_error( "list entry of the evaluator disagrees with the fold");
				}
				if (C.flag || C.nitems != R.nitems || C.ntakes != R.ntakes || C.has_result != R.has_result || C.signals != R.signals || C.open != R.open)
				{
					throw std::runtime_error( "count pass and write pass of the evaluator disagree");
				}
				st.nof_programs_installed += R.installed; st.nof_altkey_programs_installed += R.altkey;
				st.nof_signals_fired += R.signals; st.nof_open_patterns += R.open;
				if (!R.has_result) continue;
				const uint32_t want = sp_eval_result_items( R);
				const std::size_t ib = items.size();
				items.resize( ib + want);
				const uint32_t got = want ? sp_eval_write_items( T, tokens, segs, tb, te, takes, R.ntakes, items.data() + ib) : 0;
				if (got != want) throw std::runtime_error( "evaluator item count differs from the items written");
				sp_result rec;
				sp_eval_result_record( T.programs[ pt.program], R, (uint32_t)(ib - item_base), want, rec);
				results.push_back( rec);
				// the write pass from the compiled list entry must produce the same record and items
				{
					const SpEvalEntry& en = T.eval_entries[ es.prog_begin + k];
					const SpEvalVars& V = T.eval_vars[ es.prog_begin + k];
					SpEvalTrace tr;
					if (sp_eval_lite_trace( en, V, tokens, tb, te, site, tr))
					{
						const SpEvalFollow* F = site.cpos ? &T.eval_follows[ site.follow_begin + site.cpos - 1] : 0;
						uint32_t extra = 0;
						bool ok = !F || sp_eval_lite_follow_items( T, tokens, site.tok, te, en, *F, extra);
						const uint32_t lwant = (tr.flags & SP_LITE_DREFEMIT) ? tr.nitems + extra : 0;
						std::vector<sp_item> litems( lwant + 64);
						uint32_t lgot = 0;
						if (ok && lwant) lgot = sp_eval_lite_write_items( T, tokens, segs, tb, te, site, V, tr, F, litems.data(), ok);
						if (ok)
						{
							sp_result lrec;
							sp_eval_lite_result_record( T.programs[ pt.program], tokens, segs, tb, tr, (uint32_t)(ib - item_base), lwant, lrec);
							if (lwant != want || lgot != lwant || std::memcmp( &lrec, &rec, sizeof(rec)) != 0
							    || (want && std::memcmp( litems.data(), items.data() + ib, want * sizeof(sp_item)) != 0))
							{
								throw std::runtime_error( "write pass from the list entry differs from the full fold");
							}
							++g_lite_written;
						}
					}
					++g_results_written;
				}
			}
		}
	}
	st.nof_events = te - tb;
	return true;
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

int msim_set_mode( void* h, int mode) {((MatcherSim*)h)->mode = mode; return 0;}

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
			if (ms->mode == 1 && T.eval_capable)
			{
				std::size_t rb = ms->results.size(), ib = ms->items.size();
				sp_matcher_stats est;
				++ms->eval_docs;
				if (evalDocument( T, tokens, origsegs, tok_offsets[ d], tok_offsets[ d+1], ms->results, ms->items, est))
				{
					ms->stats.push_back( est);
					ms->status.push_back( 0);
					ms->result_offsets.push_back( ms->results.size());
					ms->item_offsets.push_back( ms->items.size());
					continue;
				}
				++ms->eval_flagged;
				ms->results.resize( rb); ms->items.resize( ib);
			}
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
static int64_t valueOut( const std::string& val, char* buf, uint64_t capacity)
{
	if (buf && capacity) std::memcpy( buf, val.data(), (std::size_t)(capacity < val.size() ? capacity : val.size()));
	return (int64_t)val.size();
}
int64_t msim_result_value( void* h, const sp_result* result, const sp_item* doc_items, char* buf, uint64_t capacity)
{
	try {std::string val; if (!((MatcherSim*)h)->prog.resultValue( *result, doc_items, val)) return SP_NO_VALUE; return valueOut( val, buf, capacity);}
	catch (const std::exception& e) {((MatcherSim*)h)->lasterror = e.what(); return SP_ERR_ARGUMENT;}
}
int64_t msim_item_value( void* h, const sp_item* item, char* buf, uint64_t capacity)
{
	try {std::string val; if (!((MatcherSim*)h)->prog.itemValue( item, val)) return SP_NO_VALUE; return valueOut( val, buf, capacity);}
	catch (const std::exception& e) {((MatcherSim*)h)->lasterror = e.what(); return SP_ERR_ARGUMENT;}
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
		if (k == "eval_capable") return H.view.eval_capable;
		if (k == "eval_follows") return H.view.nfollows;
		if (k == "eval_docs") return (int64_t)ms->eval_docs;
		if (k == "eval_flagged") return (int64_t)ms->eval_flagged;
		if (k == "lite_written") return (int64_t)g_lite_written;
		if (k == "results_written") return (int64_t)g_results_written;
	}
	catch (...) {}
	return -1;
}

} // extern "C"
