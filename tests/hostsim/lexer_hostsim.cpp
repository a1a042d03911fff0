// TEST INFRASTRUCTURE (not product code): runs the kernel bodies of struspattern_b200/csrc/lexer_device.cuh
// as plain loops on the CPU, with the tables produced by the product's host compiler. It lets the
// "not gpu" test suite check the compiled tables and the per-thread logic against the oracle while no GPU
// is present. Nothing under struspattern_b200/ links or loads this file; the shipped library has no CPU path.
#include <cstdint>
#include <cstring>
struct uint4 {uint32_t x, y, z, w;};
struct uint2 {uint32_t x, y;};
#include "lexer_compile.hpp"
#include "lexer_device.cuh"
#include "lexer_fused.cuh"
#include <vector>
#include <string>

namespace {
struct HostSim
{
	sp::LexerProgram prog;
	std::string lasterror;
	std::vector<sp_lexem> lexems;
	std::vector<uint64_t> offsets;
	std::vector<uint32_t> status;
	std::vector<SpRawEvent> events;
	std::vector<uint64_t> doc_ev_begin;
	uint64_t fixups, rounds;
	int mode;	// 0: one loop per document (reference order), 1: island + slot formulation, 2: fused segment formulation
	uint64_t nislands, nfixed, fused, legacy, cutstates, leadfixed;
	HostSim() :fixups(0),rounds(0),mode(0),nislands(0),nfixed(0),fused(0),legacy(0),cutstates(0),leadfixed(0){}
};
}

extern "C" {

void* hsim_create() {return new HostSim();}
void hsim_destroy( void* h) {delete (HostSim*)h;}
const char* hsim_last_error( void* h) {return ((HostSim*)h)->lasterror.c_str();}

#define HSIM_TRY(stmt) HostSim* hs = (HostSim*)h; try {stmt; return 0;} \
	catch (const sp::UnsupportedError& e) {hs->lasterror = std::string("unsupported: ") + e.what(); return -6;} \
	catch (const sp::SyntaxError& e) {hs->lasterror = e.what(); return -2;} \
	catch (const std::exception& e) {hs->lasterror = e.what(); return -1;}

int hsim_define_option( void* h, const char* name, double) {HSIM_TRY( hs->prog.defineOption( name))}
int hsim_define_lexem( void* h, uint32_t id, const char* expr, size_t exprlen, uint32_t resultIndex, uint32_t level, uint32_t posbind)
	{HSIM_TRY( hs->prog.defineLexem( id, std::string( expr, exprlen), resultIndex, level, posbind))}
int hsim_define_symbol( void* h, uint32_t symbolid, uint32_t lexemid, const char* name, size_t namelen)
	{HSIM_TRY( hs->prog.defineSymbol( symbolid, lexemid, std::string( name, namelen)))}
int hsim_compile( void* h) {HSIM_TRY( hs->prog.compile())}
// image of the compiled program (LexerProgram::serialize / deserialize)
int64_t hsim_serialize( void* h, uint8_t* image, uint64_t capacity)
{
	HostSim* hs = (HostSim*)h;
	try
	{
		std::vector<uint8_t> bytes;
		hs->prog.serialize( bytes);
		if (image && capacity >= bytes.size()) std::memcpy( image, bytes.data(), bytes.size());
		return (int64_t)bytes.size();
	}
	catch (const std::exception& e) {hs->lasterror = e.what(); return -1;}
}
int hsim_deserialize( void* h, const uint8_t* image, uint64_t size) {HSIM_TRY( hs->prog.deserialize( image, (std::size_t)size))}
int hsim_set_mode( void* h, int mode) {((HostSim*)h)->mode = mode; return 0;}
int64_t hsim_info( void* h, const char* key)
{
	HostSim* hs = (HostSim*)h;
	std::string k( key);
	if (k == "fwd_states") return hs->prog.forwardStates();
	if (k == "fwd_stride") return hs->prog.tables().view.fwd_stride;
	if (k == "rev_states") return hs->prog.reverseStates();
	if (k == "fixups") return (int64_t)hs->fixups;
	if (k == "rounds") return (int64_t)hs->rounds;
	if (k == "raw_events") return (int64_t)hs->events.size();
	if (k == "islands") return (int64_t)hs->nislands;
	if (k == "fixed_islands") return (int64_t)hs->nfixed;
	if (k == "fused") return (int64_t)hs->fused;
	if (k == "fused_legacy") return (int64_t)hs->legacy;
	if (k == "lead_fixed") return (int64_t)hs->leadfixed;
	if (k == "cut_states") return (int64_t)hs->prog.tables().view.fwd_cut_n;
	if (k == "cut_lo") return (int64_t)hs->prog.tables().view.fwd_cut_lo;
	if (k == "acc_base") return (int64_t)hs->prog.tables().view.fwd_acc_base;
	return -1;
}

// how often the scan reads the row of every forward state (tools/state_histogram.py: which rows are worth a place in shared memory)
int hsim_state_histogram( void* h, const uint8_t* text, const uint64_t* doc_off, uint32_t ndocs, uint64_t* hist)
{
	HostSim* hs = (HostSim*)h;
	if (!hs->prog.compiled()) {hs->lasterror = "called scan without calling 'compile'"; return -1;}
	const SpLexerTables& T = hs->prog.tables().view;
	for (uint32_t d=0; d<ndocs; ++d)
	{
		uint32_t state = T.fwd_init;
		for (uint64_t i=doc_off[d]; i<doc_off[d+1]; ++i)
		{
			++hist[ state];
			state = T.fwd_next[ state * T.fwd_stride + T.fwd_class[ text[i]]];
		}
	}
	return 0;
}

// text must be 16 byte aligned for the vector path to be exercised (numpy arrays are)
int hsim_scan_batch( void* h, const uint8_t* text, const uint64_t* doc_off, uint32_t ndocs, uint32_t chunk, uint32_t warmup)
{
	HostSim* hs = (HostSim*)h;
	if (!hs->prog.compiled()) {hs->lasterror = "called scan without calling 'compile'"; return -1;}
	SpScanParams P; std::memset( &P, 0, sizeof(P));
	P.T = hs->prog.tables().view;
	P.text = text; P.doc_off = doc_off; P.ndocs = ndocs; P.total = doc_off[ ndocs];
	P.chunk = chunk; P.warmup = warmup;
	P.nchunks = (P.total + chunk - 1) / chunk;
	std::vector<uint32_t> start_state( P.nchunks+1), end_state( P.nchunks+1), ev_count( P.nchunks+1);
	std::vector<uint8_t> known( P.nchunks+1), redo( P.nchunks+1);
	std::vector<uint64_t> ev_off( P.nchunks+1, 0);
	P.start_state = start_state.data(); P.end_state = end_state.data(); P.ev_count = ev_count.data();
	P.known = known.data(); P.redo = redo.data();
	hs->fixups = 0; hs->rounds = 0;
	for (uint64_t j=0; j<P.nchunks; ++j) spk_count_chunk( P, j, 0, P.text);
	for (;;)
	{
		uint64_t nbad = 0;
		for (uint64_t j=0; j<P.nchunks; ++j) nbad += spk_verify_chunk( P, j);
		if (!nbad) break;
		hs->fixups += nbad; hs->rounds += 1;
		for (uint64_t j=0; j<P.nchunks; ++j) spk_count_chunk( P, j, 1, P.text);
	}
	for (uint64_t j=0; j<P.nchunks; ++j) ev_off[ j+1] = ev_off[ j] + ev_count[ j];
	P.ev_off = ev_off.data();
	uint64_t nev = ev_off[ P.nchunks];
	hs->events.assign( nev + 1, SpRawEvent());
	hs->doc_ev_begin.assign( ndocs + 1, 0);
	hs->doc_ev_begin[ ndocs] = nev;
	P.events = hs->events.data(); P.doc_ev_begin = hs->doc_ev_begin.data();
	std::vector<uint32_t> head( P.nchunks + 1, 0), ev_docflag( nev + 1, 0);
	P.head = head.data(); P.ev_doc = ev_docflag.data();
	for (uint64_t j=0; j<P.nchunks; ++j) spk_emit_chunk( P, j, P.text);
	hs->events.resize( nev);
	for (uint64_t i=0; i<nev; ++i) spk_finish_event( P, i);
	P.ar_mult = P.T.has_symbols ? 2 : 1;
	std::vector<SpMatchEvent> ar( nev * P.ar_mult + 2);
	std::vector<uint32_t> doc_kept( ndocs+1), doc_nlex( ndocs+1);
	std::vector<uint64_t> lex_off( ndocs+1, 0);
	hs->status.assign( ndocs, 0);
	uint32_t dummy = 0;
	P.ar = ar.data(); P.doc_kept = doc_kept.data(); P.doc_nlex = doc_nlex.data(); P.status = ndocs ? hs->status.data() : &dummy;
	hs->fused = 0; hs->legacy = 0; hs->leadfixed = 0;
	bool fusedDone = false;
	if ((hs->mode == 3 && P.T.fwd_cut_n && P.nchunks)
	||  (hs->mode == 2 && P.T.fwd_cut_n && P.T.npat <= SP_FUSED_MAXPAT && P.nchunks && P.T.fwd_max_acc + 1 <= SP_FUSED_BUF))
	{
		// the fused segment pipeline of lexer_fused.cuh as plain loops (same per-thread functions as the kernels)
		const uint64_t nch = P.nchunks;
		std::vector<SpSlot> slots( nev * P.ar_mult + 1);
		std::vector<uint64_t> region_begin( nch), agg( nch), prefix( nch + 1, 0), doc_local( ndocs + 1, 0), doc_base( ndocs + 1, 0);
		std::vector<uint32_t> kept( nch), lead_pred( nch), inc_tail( nch), doc_chunk( ndocs + 1, hs->mode == 3 ? 0xFFFFFFFFu : 0);

		std::vector<uint8_t> has_docstart( nch), has_content( ndocs + 1, 0), succ_early( ndocs + 1, 0);
		uint32_t need_legacy = 0;
		SpFusedParams F; std::memset( &F, 0, sizeof(F));
		F.T = P.T; F.text = text; F.doc_off = doc_off; F.ndocs = ndocs; F.total = P.total; F.chunk = chunk; F.nchunks = nch;
		F.start_state = start_state.data(); F.ev_off = ev_off.data(); F.ar_mult = P.ar_mult; F.slots = slots.data();
		F.region_begin = region_begin.data(); F.kept = kept.data(); F.agg = agg.data(); F.lead_pred = lead_pred.data();
		F.inc_tail = inc_tail.data(); F.has_docstart = has_docstart.data(); F.prefix = prefix.data();
		F.doc_chunk = doc_chunk.data(); F.doc_local = doc_local.data(); F.doc_has_content = has_content.data();
		F.doc_succ_early = succ_early.data(); F.doc_base = doc_base.data(); F.lex_off = lex_off.data(); F.need_legacy = &need_legacy;
		F.events = hs->events.data(); F.ev_doc = ev_docflag.data(); F.head = head.data(); F.nev = nev;
		for (uint64_t j=0; j<nch; ++j)
		{
			uint32_t buf[ SP_FUSED_BUF]; SpMatchEvent arlocal[ SP_FUSED_AR]; SpStartMemo memo[ SP_FUSED_MEMO];
			SpStridedEvents ar; ar.base = arlocal; ar.stride = 1;
			if (hs->mode == 3) spk_resolve_chunk( F, j, ar); else spk_fused_chunk( F, j, buf, 1, ar, memo);
		}
		for (uint64_t j=0; j<nch; ++j) {uint32_t before = lead_pred[ j]; spk_fused_fix_chunk( F, j); if (before) hs->leadfixed += 1;}
		for (uint64_t j=0; j<nch; ++j) prefix[ j+1] = prefix[ j] + agg[ j];
		for (uint32_t d=0; d<=ndocs; ++d) spk_fused_doc( F, d, prefix[ nch]);
		if (need_legacy) hs->legacy = 1;
		else
		{
			hs->lexems.assign( (prefix[ nch] >> 32) + 1, sp_lexem());
			F.lexems = hs->lexems.data();
			for (uint64_t j=0; j<nch; ++j) spk_fused_write_chunk( F, j);
			hs->lexems.resize( prefix[ nch] >> 32);
			hs->fused = 1;
			fusedDone = true;
		}
	}
	if (fusedDone) {}
	else if (hs->mode == 0)
	{
		for (uint32_t d=0; d<ndocs; ++d) spk_resolve_doc( P, d);
		for (uint32_t d=0; d<ndocs; ++d) lex_off[ d+1] = lex_off[ d] + doc_nlex[ d];
		hs->lexems.assign( lex_off[ ndocs] + 1, sp_lexem());
		P.lex_off = lex_off.data(); P.lexems = hs->lexems.data();
		for (uint32_t d=0; d<ndocs; ++d) spk_write_doc( P, d);
		hs->lexems.resize( lex_off[ ndocs]);
	}
	else
	{
		// the slot pipeline of lexer_islands.cuh as plain loops (same per-thread functions)
		std::vector<uint32_t> ev_doc( nev + 1);
		for (uint32_t d=0; d<ndocs; ++d) for (uint64_t e=hs->doc_ev_begin[ d]; e<hs->doc_ev_begin[ d+1]; ++e) ev_doc[ e] = d;
		std::vector<uint64_t> sufmin( nev + 1, ~0ull);
		for (uint64_t k=nev; k>0; --k) {uint64_t lo = sp_event_lo_key( hs->events[ k-1], ev_doc[ k-1]); sufmin[ k-1] = lo < sufmin[ k] ? lo : sufmin[ k];}
		std::vector<uint32_t> isl_start;
		uint64_t premax = 0;
		for (uint64_t k=0; k<nev; ++k)
		{
			if (sufmin[ k] >= premax) isl_start.push_back( (uint32_t)k);
			uint64_t r = sp_event_r_key( hs->events[ k], ev_doc[ k]);
			if (r > premax) premax = r;
		}
		uint64_t nisl = isl_start.size();
		isl_start.push_back( (uint32_t)nev);
		hs->nislands = nisl; hs->nfixed = 0;
		const uint64_t nslots = nev * P.ar_mult;
		std::vector<SpSlot> slots( nslots + 1);
		std::vector<uint8_t> has_content( ndocs + 1, 0);
		std::vector<uint32_t> first_content( ndocs + 1, SP_NO_SLOT), fix_list;
		std::vector<uint64_t> doc_base( ndocs + 1, 0);
		SpSlotParams S; std::memset( &S, 0, sizeof(S));
		S.T = P.T; S.events = hs->events.data(); S.ev_doc = ev_doc.data(); S.nev = nev; S.ar_mult = P.ar_mult;
		S.slots = slots.data(); S.isl_start = isl_start.data(); S.nisl = nisl;
		S.doc_status = P.status; S.doc_has_content = has_content.data(); S.doc_first_content = first_content.data();
		S.doc_ev_begin = hs->doc_ev_begin.data(); S.ndocs = ndocs; S.doc_base = doc_base.data(); S.lex_off = lex_off.data();
		for (uint64_t i=0; i<nisl; ++i)
		{
			uint64_t cap = (uint64_t)(isl_start[ i+1] - isl_start[ i]) * P.ar_mult;
			SpMatchEvent local[ SP_ISLAND_LOCAL];
			// small islands in "thread local" memory, oversized ones in their own slot range, like the kernel
			SpMatchEvent* work = cap <= SP_ISLAND_LOCAL ? local : reinterpret_cast<SpMatchEvent*>( slots.data() + (uint64_t)isl_start[ i] * P.ar_mult);
			if (sp_island_to_slots( S, i, work)) fix_list.push_back( (uint32_t)i);
		}
		for (uint32_t i: fix_list) sp_island_fix_unique( S, i);
		hs->nfixed = fix_list.size();
		std::vector<uint64_t> excl( nslots + 1, 0);
		for (uint64_t j=0; j<nslots; ++j) excl[ j+1] = excl[ j] + sp_slot_value( S, j, slots[ j]);
		for (uint32_t d=0; d<=ndocs; ++d)
		{
			uint64_t fs = hs->doc_ev_begin[ d] * P.ar_mult;
			doc_base[ d] = excl[ fs < nslots ? fs : nslots];
			lex_off[ d] = doc_base[ d] >> 32;
		}
		hs->lexems.assign( (excl[ nslots] >> 32) + 1, sp_lexem());
		S.lexems = hs->lexems.data();
		for (uint64_t j=0; j<nslots; ++j) sp_slot_emit( S, j, slots[ j], excl[ j], excl[ j+1] - excl[ j]);
		hs->lexems.resize( excl[ nslots] >> 32);
	}
	hs->offsets = lex_off;
	return 0;
}
uint64_t hsim_result_count( void* h) {return ((HostSim*)h)->lexems.size();}
uint64_t hsim_raw_event_count( void* h) {return ((HostSim*)h)->events.size();}
int hsim_fetch( void* h, sp_lexem* lexems, uint64_t* offsets, uint32_t* status)
{
	HostSim* hs = (HostSim*)h;
	if (lexems && !hs->lexems.empty()) std::memcpy( lexems, hs->lexems.data(), hs->lexems.size()*sizeof(sp_lexem));
	if (offsets) std::memcpy( offsets, hs->offsets.data(), hs->offsets.size()*sizeof(uint64_t));
	if (status && !hs->status.empty()) std::memcpy( status, hs->status.data(), hs->status.size()*sizeof(uint32_t));
	return 0;
}
// raw events after start resolution as (definition index 1-based, from, to) triples + per document offsets
int hsim_fetch_raw( void* h, uint32_t* triples, uint64_t* offsets)
{
	HostSim* hs = (HostSim*)h;
	for (std::size_t i=0; i<hs->events.size(); ++i)
	{
		triples[ 3*i] = hs->events[i].pat + 1; triples[ 3*i+1] = hs->events[i].from; triples[ 3*i+2] = hs->events[i].to;
	}
	if (offsets) std::memcpy( offsets, hs->doc_ev_begin.data(), hs->doc_ev_begin.size()*sizeof(uint64_t));
	return 0;
}

} // extern "C"
