// Window evaluator of the level-2 matcher: the same results as the event-driven state machine of
// matcher_device.cuh, computed WITHOUT mutable per-document state.
//
// The reference (src/ruleMatcherAutomaton.cpp) installs a rule instance for every occurrence of a key event
// (installProgram :1168-1270), files its triggers in hash buckets, and lets later events find them (getTriggers
// :229-257, fireSignal :772-979) until the rule is finished or its position window has passed (setCurrentPos
// :1084-1135). What one rule instance does is a pure function of the events between its installation and its
// expiry: a fold of fireSignal over the window. This file is that fold: one (site, program) pair per thread, the
// document's tokens read-only, no arena, no bucket arrays, no atomics. The CUDA kernels (matcher_eval_kernels.cuh)
// run it for all (token, keyed program) pairs of a batch in parallel; tests/hostsim runs it on the CPU.
//
// What the fold cannot know is the ORDER of the reference's trigger arrays (append + swap-with-last removal, i.e.
// the whole history of the document). Everything a rule instance computes is independent of that order except when
// one event hits two live triggers of the same rule; the fold then raises `flag` and the document is run again by
// the state machine (exact in every order). The results of a document come out ordered by (key token, program list
// order) instead of emission order; the contract is the multiset per document with ordered items.
//
// Events that are not term events: only follow events of programs whose triggers are all of type "any" and wait for
// term events. Such a program emits when it is installed (first signal on a fresh slot, :794-806), so the follow
// events of a token are a static list per term event (matcher_tables.h: SpEvalFollow) in the order of the program
// list (:1137-1147). Rule sets outside of this (matcher_program.hpp: eval_capable == 0) use the state machine.
//
// Statistics (src/patternMatcher.cpp:303-318) are sums over rule instances: programs installed, alternative-key
// programs installed, signals fired (trigger hits + key signals + replayed signals), and the live triggers summed
// over the input tokens (:992, sampled once per doTransition call) = sum over rules of triggers x tokens survived.
#ifndef SP_MATCHER_EVAL_CUH_INCLUDED
#define SP_MATCHER_EVAL_CUH_INCLUDED
#include "matcher_device.cuh"

#define SP_EVAL_MAX_TAKES 32
#ifdef __CUDACC__
#define sp_ctz( m) (__ffs( (int)(m)) - 1)
#define sp_popc( m) __popc( m)
#else
#define sp_ctz( m) __builtin_ctz( m)
#define sp_popc( m) __builtin_popcount( m)
#endif
#define SP_EVAL_NOSLOT 0xFFFFFFFFu

/* One element taken by a rule (fireSignal, takeEventData :900-934): what the rule's item list gets. */
typedef struct SpEvalTake
{
	uint32_t variable;		/* variable of the trigger; 0 = the items of the follow event were joined (:911-918) */
	uint32_t tokrel;		/* token of the event, relative to the first token of the document */
	uint32_t follow;		/* 0 = the token event itself, else 1 + index into eval_follows */
} SpEvalTake;

/* A place where programs are installed: a token event or one of its static follow events. */
typedef struct SpEvalSite
{
	uint64_t tok;			/* token index (absolute) */
	uint32_t cpos;			/* 0 = the token event; c = follow event number c-1 of the token */
	uint32_t keyevent;
	uint32_t follow_begin, follow_count;	/* static follow events of the token (eval_follows range) */
	SpEventData data;		/* EventData of the site event; subdataref is used as a flag here */
} SpEvalSite;

/* Slot (ActionSlot :87 + Rule :171) and outcome of one rule instance. */
typedef struct SpEvalRule
{
	uint32_t value, count, end_ordpos, start_ordpos, start_origseg, start_origpos;
	uint32_t done, disposed, has_dataref;
	uint32_t installed, altkey, signals, flag;
	uint64_t open;			/* live triggers x input tokens seen */
	uint32_t has_result, dataref_at_emit;
	uint32_t r_start_ordpos, r_end_ordpos, r_start_origseg, r_start_origpos, r_end_origseg, r_end_origpos;
	uint32_t nitems;		/* length of the rule's item list including the expansion of referenced sub lists */
	uint32_t ntakes;
} SpEvalRule;

SP_MHD uint32_t sp_find_event_index( const SpMatcherTables& T, uint32_t event)
{
	uint32_t key = event + 1;
	for (uint32_t h = sp_event_slot_hash( event) & T.event_mask;; h = (h + 1) & T.event_mask)
	{
		uint32_t k = SP_MLDG( &T.event_slots[ h].key);
		if (k == key) return h;
		if (k == 0) return SP_EVAL_NOSLOT;
	}
}

SP_MHD void sp_eval_token_data( const sp_lexem& tok, uint32_t origseg, SpEventData& data)
{
	data.start_origseg = origseg; data.start_origpos = tok.origpos; data.end_origseg = origseg; data.end_origpos = tok.origpos + tok.origsize;
	data.start_ordpos = tok.ordpos; data.end_ordpos = tok.ordpos + 1; data.subdataref = 0; data.formathandle = 0;
}

/* Tokens the evaluator hands to the state machine (putInput checks, src/patternMatcher.cpp:136-156, and position 0,
 * which the slots use as "not set", src/ruleMatcherAutomaton.cpp:919). */
SP_MHD bool sp_eval_token_irregular( const sp_lexem& tok, uint32_t origseg, bool first, uint32_t prev_ordpos)
{
	if (tok.ordpos == 0 || tok.id >= (1u<<29)) return true;
	if (tok.origsize >= 0x7FFFFFFFu || tok.origpos >= 0x7FFFFFFFu || origseg >= 0x7FFFFFFFu) return true;
	return !first && tok.ordpos < prev_ordpos;
}

SP_MHD void sp_eval_site_token( const SpMatcherTables& T, const sp_lexem& tok, uint32_t origseg, uint64_t i, uint32_t slotidx, SpEvalSite& site)
{
	site.tok = i; site.cpos = 0; site.keyevent = tok.id;
	site.follow_begin = site.follow_count = 0;
	if (slotidx != SP_EVAL_NOSLOT)
	{
		site.follow_begin = SP_MLDG( &T.eval_slots[ slotidx].follow_begin);
		site.follow_count = SP_MLDG( &T.eval_slots[ slotidx].follow_count);
	}
	sp_eval_token_data( tok, origseg, site.data);
}

/* The follow event F of a token event: same span as the token (slot start = data start on the first signal, :919-924;
 * end_ordpos = data.end_ordpos, :801-804; follow EventData :941). */
SP_MHD void sp_eval_follow_data( const SpEventData& tokdata, const SpEvalFollow& F, SpEventData& data)
{
	data = tokdata;
	data.subdataref = F.sub;
	data.formathandle = F.format;
}

SP_MHD void sp_eval_site_follow( const SpEvalSite& base, uint32_t c, const SpEvalFollow& F, SpEvalSite& site)
{
	site = base;
	site.cpos = c + 1; site.keyevent = F.event;
	sp_eval_follow_data( base.data, F, site.data);
}

SP_MHD void sp_eval_rule_init( SpEvalRule& R)
{
	R.value = R.count = R.end_ordpos = R.start_ordpos = R.start_origseg = R.start_origpos = 0;
	R.done = R.disposed = R.has_dataref = 0;
	R.installed = R.altkey = R.signals = R.flag = 0; R.open = 0;
	R.has_result = R.dataref_at_emit = 0;
	R.r_start_ordpos = R.r_end_ordpos = R.r_start_origseg = R.r_start_origpos = R.r_end_origseg = R.r_end_origpos = 0;
	R.nitems = R.ntakes = 0;
}

/* fireSignal (src/ruleMatcherAutomaton.cpp:772-979) on the private slot of the rule. Returns true if the rule takes
 * the event data. The follow event of the rule itself is not produced here: programs that emit events somebody waits
 * for are the static "any" programs (see above). */
SP_MHD bool sp_eval_fire( SpEvalRule& R, const SpProgram& P, uint32_t sigtype, uint32_t sigval, uint32_t variable, const SpEventData& data, bool replay)
{
	bool match = false, take = false, finished = false;
	R.signals += 1;
	switch (sigtype)
	{
		case SP_SIG_ANY:
			take = true;
			if (R.count > 0)
			{
				match = true;
				R.count -= 1;
				finished = (R.count == 0);
				if (R.end_ordpos < data.end_ordpos) R.end_ordpos = data.end_ordpos;
			}
			break;
		case SP_SIG_SEQUENCE:
		case SP_SIG_SEQUENCE_IMM:
		case SP_SIG_WITHIN:
		{
			bool accept;
			if (sigtype == SP_SIG_WITHIN) accept = (sigval & R.value) != 0 && R.end_ordpos <= data.start_ordpos;
			else if (sigtype == SP_SIG_SEQUENCE) accept = sigval == R.value && R.end_ordpos <= data.start_ordpos;
			else accept = sigval == R.value && R.end_ordpos == data.start_ordpos;
			if (accept)
			{
				R.end_ordpos = data.end_ordpos;
				if (sigtype == SP_SIG_WITHIN) R.value &= ~sigval; else R.value = sigval - 1;
				if (R.count > 0)
				{
					R.count -= 1;
					match = (R.count == 0);
				}
				else match = true;
				finished = (R.value == 0);
				take = true;
			}
			break;
		}
		case SP_SIG_DEL:
			R.count = 0;
			R.value = 0;
			R.disposed = 1;
			return false;
		default:	/* SP_SIG_AND: such rule sets are not eval_capable */
			R.flag = 1;
			return false;
	}
	if (take)
	{
		if (variable || data.subdataref) R.has_dataref = 1;
		if (R.start_ordpos == 0)
		{
			R.start_ordpos = data.start_ordpos;
			R.start_origseg = data.start_origseg;
			R.start_origpos = data.start_origpos;
		}
		else if (R.start_ordpos > data.start_ordpos)
		{
			R.start_ordpos = data.start_ordpos;
			if (R.start_origseg > data.start_origseg || (R.start_origseg == data.start_origseg && R.start_origpos > data.start_origpos))
			{
				R.start_origseg = data.start_origseg;
				R.start_origpos = data.start_origpos;
			}
		}
	}
	if (match)
	{
		if (!R.done)
		{
			if (P.event && replay) R.flag = 1;	/* "encountered past trigger with follow" (:1328-1332) */
			if (P.resultHandle)
			{
				R.has_result = 1; R.dataref_at_emit = R.has_dataref;
				R.r_start_ordpos = R.start_ordpos; R.r_end_ordpos = R.end_ordpos;
				R.r_start_origseg = R.start_origseg; R.r_start_origpos = R.start_origpos;
				R.r_end_origseg = data.end_origseg; R.r_end_origpos = data.end_origpos;
			}
			R.done = 1;
		}
		if (finished) R.disposed = 1;
	}
	return take;
}

template <bool COLLECT, int DEPTH>
SP_MHD void sp_eval_fold( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
			  const SpEvalSite& site, const SpProgramTrigger pt, bool install_only, SpEvalRule& R, SpEvalTake* takes);

/* Item list bookkeeping of a taken event (:900-918). */
template <bool COLLECT, int DEPTH>
SP_MHD void sp_eval_take( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
			  SpEvalRule& R, SpEvalTake* takes, uint32_t variable, uint64_t tok, uint32_t follow, const SpEventData& data)
{
	if (!variable && !data.subdataref) return;
	if (R.ntakes >= SP_EVAL_MAX_TAKES) {R.flag = 1; return;}
	if (COLLECT)
	{
		SpEvalTake& t = takes[ R.ntakes];
		t.variable = variable; t.tokrel = (uint32_t)(tok - tb); t.follow = follow;
	}
	R.ntakes += 1;
	if (variable) R.nitems += 1;
	if (DEPTH == 0 && follow && data.subdataref && (!variable || !data.formathandle))
	{
		// the item list of the rule that emitted the follow event: referenced (variable: read when the results are
		// fetched, i.e. complete) or joined (no variable: copied now, i.e. as it is after the installation)
		const SpEvalFollow F = T.eval_follows[ follow-1];
		SpEvalSite qsite;
		sp_eval_site_token( T, toks[ tok], segs ? segs[ tok] : 0, tok, SP_EVAL_NOSLOT, qsite);
		SpEvalRule Q;
		sp_eval_fold<false,1>( T, toks, segs, tb, te, qsite, T.program_triggers[ F.qpt], variable == 0, Q, 0);
		R.nitems += Q.nitems;
		if (Q.flag) R.flag = 1;
	}
}

SP_MHD bool sp_eval_rule_same( const SpEvalRule& a, const SpEvalRule& b)
{
	return a.value == b.value && a.count == b.count && a.end_ordpos == b.end_ordpos && a.start_ordpos == b.start_ordpos
		&& a.start_origseg == b.start_origseg && a.start_origpos == b.start_origpos && a.done == b.done && a.disposed == b.disposed
		&& a.has_dataref == b.has_dataref && a.flag == b.flag && a.has_result == b.has_result && a.dataref_at_emit == b.dataref_at_emit
		&& a.r_start_ordpos == b.r_start_ordpos && a.r_end_ordpos == b.r_end_ordpos && a.r_start_origseg == b.r_start_origseg
		&& a.r_start_origpos == b.r_start_origpos && a.r_end_origseg == b.r_end_origseg && a.r_end_origpos == b.r_end_origpos
		&& a.nitems == b.nitems && a.ntakes == b.ntakes;
}

/* Signals of one event on the live triggers of the rule (getTriggers + fireSignal loop, :1015-1024). The triggers fire
 * in the order of the bucket array, which the fold does not know. One hit (the usual case) has no order. Two hits are
 * applied in both orders: if the rule ends up the same (a sequence accepts only the trigger whose signal value is due,
 * and the first one accepted moves end_ordpos past the event), the order does not matter; otherwise, and with more
 * than two hits, the document is flagged for the state machine. */
template <bool COLLECT, int DEPTH>
SP_MHD void sp_eval_hits( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
			  const SpProgram& P, uint32_t installed_mask, uint32_t event, const SpEventData& data, uint64_t tok, uint32_t follow,
			  SpEvalRule& R, SpEvalTake* takes)
{
	uint32_t nhits = 0, d1 = 0, d2 = 0;
	for (uint32_t m = installed_mask; m; m &= m - 1)
	{
		const uint32_t d = (uint32_t)sp_ctz( m);
		if (SP_MLDG( &T.trigger_defs[ P.trig_begin + d].event) != event) continue;
		if (nhits == 0) d1 = d; else if (nhits == 1) d2 = d;
		++nhits;
	}
	if (nhits == 0) return;
	if (nhits == 1)
	{
		const SpTriggerDef def = T.trigger_defs[ P.trig_begin + d1];
		if (sp_eval_fire( R, P, def.sigtype, def.sigval, def.variable, data, false))
		{
			sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, def.variable, tok, follow, data);
		}
		return;
	}
	if (nhits > 2)
	{
		// more than two hits: order independent if they are all sequence triggers (signal values of one rule are distinct, so
		// at most one of them is due, and the one accepted moves end_ordpos past the event: the others change nothing)
		bool sequences = true;
		for (uint32_t m = installed_mask; m; m &= m - 1)
		{
			const SpTriggerDef def = T.trigger_defs[ P.trig_begin + (uint32_t)sp_ctz( m)];
			if (def.event == event && def.sigtype != SP_SIG_SEQUENCE && def.sigtype != SP_SIG_SEQUENCE_IMM) sequences = false;
		}
		if (!sequences) {R.flag = 1; R.signals += nhits; return;}
		for (uint32_t m = installed_mask; m; m &= m - 1)
		{
			const SpTriggerDef def = T.trigger_defs[ P.trig_begin + (uint32_t)sp_ctz( m)];
			if (def.event != event) continue;
			if (sp_eval_fire( R, P, def.sigtype, def.sigval, def.variable, data, false))
			{
				sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, def.variable, tok, follow, data);
			}
		}
		return;
	}
	const SpTriggerDef defA = T.trigger_defs[ P.trig_begin + d1], defB = T.trigger_defs[ P.trig_begin + d2];
	SpEvalRule Rb = R;
	uint32_t vb[ 2] = {0xFFFFFFFFu, 0xFFFFFFFFu}, va[ 2] = {0xFFFFFFFFu, 0xFFFFFFFFu};
	if (sp_eval_fire( Rb, P, defB.sigtype, defB.sigval, defB.variable, data, false))
	{
		vb[ 0] = defB.variable;
		sp_eval_take<false,DEPTH>( T, toks, segs, tb, te, Rb, 0, defB.variable, tok, follow, data);
	}
	if (sp_eval_fire( Rb, P, defA.sigtype, defA.sigval, defA.variable, data, false))
	{
		vb[ vb[ 0] == 0xFFFFFFFFu ? 0 : 1] = defA.variable;
		sp_eval_take<false,DEPTH>( T, toks, segs, tb, te, Rb, 0, defA.variable, tok, follow, data);
	}
	if (sp_eval_fire( R, P, defA.sigtype, defA.sigval, defA.variable, data, false))
	{
		va[ 0] = defA.variable;
		sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, defA.variable, tok, follow, data);
	}
	if (sp_eval_fire( R, P, defB.sigtype, defB.sigval, defB.variable, data, false))
	{
		va[ va[ 0] == 0xFFFFFFFFu ? 0 : 1] = defB.variable;
		sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, defB.variable, tok, follow, data);
	}
	if (!sp_eval_rule_same( R, Rb) || va[ 0] != vb[ 0] || va[ 1] != vb[ 1]) R.flag = 1;
}

/* One rule instance from its installation to its end.
 * install_only: stop behind the signals of the installation (the item list as a joining rule sees it). */
template <bool COLLECT, int DEPTH>
SP_MHD void sp_eval_fold( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
			  const SpEvalSite& site, const SpProgramTrigger pt, bool install_only, SpEvalRule& R, SpEvalTake* takes)
{
	sp_eval_rule_init( R);
	const SpProgram P = T.programs[ pt.program];
	const uint32_t curpos = SP_MLDG( &toks[ site.tok].ordpos);
	const uint32_t expiry = site.data.start_ordpos + P.range;
	if (expiry < curpos) return;	// installProgram :1171: the rule cannot match anymore
	R.installed = 1;
	R.value = P.initsigval; R.count = P.initcount & 0xFFFFu;

	// which definitions become live triggers, which are signalled directly (:1204-1250)
	uint32_t installed_mask = 0, key_mask = 0;
	{
		bool hasKeyEvent = false;
		for (uint32_t d=0; d<P.trig_count; ++d)
		{
			const SpTriggerDef def = T.trigger_defs[ P.trig_begin + d];
			bool install = true;
			if (def.event == site.keyevent)
			{
				key_mask |= 1u << d;
				const bool needsInstall = (def.sigtype == SP_SIG_ANY && R.count > 1);
				if (def.is_key && !hasKeyEvent) {hasKeyEvent = true; install = needsInstall;}
				else if (def.sigtype == SP_SIG_DEL) install = needsInstall;
			}
			if (install) installed_mask |= 1u << d;
		}
	}
	// replayPastEvent (:1272-1334): the latest occurrence of the original key event (a stop word) is signalled
	if (pt.past_event)
	{
		R.altkey = 1;
		SpEventSlot pes;
		if (sp_find_event( T, pt.past_event, pes) && pes.stopword)
		{
			// The log holds every event handled before the site event: the tokens in front of the site token with
			// their follow events, the site token itself if the site is one of its follow events, and the follow
			// events in front of the site.
			const bool pastIsTerm = (pt.past_event >> 29) == 0;
			uint64_t j = site.cpos ? site.tok + 1 : site.tok;
			uint32_t jfollow = 0;		// 1 + index into eval_follows if the logged event is a follow event
			bool found = false;
			while (j > tb && !found)
			{
				--j;
				const uint32_t jpos = SP_MLDG( &toks[ j].ordpos);
				if (jpos + P.range < curpos) break;
				const uint32_t jid = SP_MLDG( &toks[ j].id);
				if (pastIsTerm) {found = (jid == pt.past_event); continue;}
				uint32_t fb, fc;
				if (j == site.tok) {fb = site.follow_begin; fc = site.cpos - 1;}
				else
				{
					const uint32_t si = sp_find_event_index( T, jid);
					if (si == SP_EVAL_NOSLOT) continue;
					fb = T.eval_slots[ si].follow_begin; fc = T.eval_slots[ si].follow_count;
				}
				for (uint32_t c=fc; c>0 && !found; --c)
				{
					if (T.eval_follows[ fb + c-1].event == pt.past_event) {found = true; jfollow = fb + c;}
				}
			}
			if (found)
			{
				SpEventData past;
				sp_eval_token_data( toks[ j], segs ? segs[ j] : 0, past);
				if (jfollow)
				{
					SpEventData tokdata = past;
					sp_eval_follow_data( tokdata, T.eval_follows[ jfollow-1], past);
				}
				uint32_t del_mask = 0;
				for (uint32_t d=P.trig_count; d>0; --d)		// trigger list of the rule: last installed first
				{
					if (!(installed_mask & (1u << (d-1)))) continue;
					const SpTriggerDef def = T.trigger_defs[ P.trig_begin + d-1];
					if (def.sigtype == SP_SIG_DEL) del_mask |= 1u << (d-1);
					if (def.event == pt.past_event)
					{
						if (sp_eval_fire( R, P, def.sigtype, def.sigval, def.variable, past, true))
						{
							sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, def.variable, j, jfollow, past);
						}
					}
				}
				// a structure delimiter (a term event, matcher_program.hpp: buildEvalTables) logged after the replayed
				// event kills the rule (:1306-1321)
				const uint64_t kend = site.cpos ? site.tok + 1 : site.tok;
				for (uint32_t d=P.trig_count; d>0 && del_mask && !R.disposed; --d)
				{
					if (!(del_mask & (1u << (d-1)))) continue;
					const uint32_t delevent = T.trigger_defs[ P.trig_begin + d-1].event;
					SpEventSlot ds;
					if (!sp_find_event( T, delevent, ds) || !ds.stopword) continue;
					for (uint64_t k=j+1; k<kend; ++k)
					{
						if (SP_MLDG( &toks[ k].id) == delevent) {R.disposed = 1; break;}
					}
				}
			}
		}
	}
	// the key event itself (:1259-1269)
	if (key_mask && !R.disposed)
	{
		const uint32_t follow = site.cpos ? site.follow_begin + site.cpos : 0;
		for (uint32_t m = key_mask; m; m &= m - 1)
		{
			const SpTriggerDef def = T.trigger_defs[ P.trig_begin + (uint32_t)sp_ctz( m)];
			if (sp_eval_fire( R, P, def.sigtype, def.sigval, def.variable, site.data, false))
			{
				sp_eval_take<COLLECT,DEPTH>( T, toks, segs, tb, te, R, takes, def.variable, site.tok, follow, site.data);
			}
		}
	}
	if (install_only || R.disposed || !installed_mask) return;

	const uint32_t ntrig = (uint32_t)sp_popc( installed_mask);
	const bool listens = DEPTH == 0 && (SP_MLDG( &T.prog_flags[ pt.program]) & SP_PROGFLAG_LISTENS_FOLLOW) != 0;
	// the follow events of the site token behind the site
	if (listens)
	{
		for (uint32_t c=site.cpos; c<site.follow_count && !R.disposed; ++c)
		{
			const SpEvalFollow F = T.eval_follows[ site.follow_begin + c];
			SpEventData fdata;
			SpEventData tokdata = site.data; tokdata.subdataref = 0; tokdata.formathandle = 0;
			sp_eval_follow_data( tokdata, F, fdata);
			sp_eval_hits<COLLECT,DEPTH>( T, toks, segs, tb, te, P, installed_mask, F.event, fdata, site.tok, site.follow_begin + c + 1, R, takes);
		}
		if (R.disposed) return;
	}
	// the tokens behind the site while the rule lives: alive while curpos <= expiry (setCurrentPos :1084-1135)
	for (uint64_t k=site.tok+1; k<te; ++k)
	{
		const sp_lexem tk = toks[ k];
		if (tk.ordpos > expiry) break;
		R.open += ntrig;
		SpEventData data;
		sp_eval_token_data( tk, segs ? segs[ k] : 0, data);
		sp_eval_hits<COLLECT,DEPTH>( T, toks, segs, tb, te, P, installed_mask, tk.id, data, k, 0, R, takes);
		if (R.disposed) break;		// deactivated at the end of the event (:1028-1034)
		if (listens)
		{
			const uint32_t si = sp_find_event_index( T, tk.id);
			if (si != SP_EVAL_NOSLOT)
			{
				const uint32_t fb = T.eval_slots[ si].follow_begin, fc = T.eval_slots[ si].follow_count;
				for (uint32_t c=0; c<fc && !R.disposed; ++c)
				{
					const SpEvalFollow F = T.eval_follows[ fb + c];
					SpEventData fdata;
					sp_eval_follow_data( data, F, fdata);
					sp_eval_hits<COLLECT,DEPTH>( T, toks, segs, tb, te, P, installed_mask, F.event, fdata, k, fb + c + 1, R, takes);
				}
				if (R.disposed) break;
			}
		}
	}
}

/* A rule instance followed through its window from its compiled list entry alone (matcher_tables.h: SpEvalEntry): no
 * loads of the program or its trigger definitions, a handful of registers of state. Gives the counters and whether the
 * rule has a result and how many items it delivers; the result record and the items themselves are produced by the
 * full fold (write pass). */
typedef struct SpEvalLite
{
	uint32_t value, count, endpos;	/* slot */
	uint32_t flags;			/* SP_LITE_* */
	uint32_t nitems, signals, alive_tokens;
} SpEvalLite;
#define SP_LITE_ALIVE     1u
#define SP_LITE_DONE      2u
#define SP_LITE_DATAREF   4u
#define SP_LITE_RESULT    8u
#define SP_LITE_DREFEMIT 16u
#define SP_LITE_FULL     32u	/* an event hit two live triggers: the full fold decides (and may flag the document) */

SP_MHD void sp_eval_lite_init( const SpEvalEntry& en, uint32_t site_ordpos, SpEvalLite& L)
{
	L.value = en.value0; L.count = SP_EVAL_M2_COUNT( en.meta2);
	L.endpos = (en.meta2 & SP_EVAL_M2_END0) ? site_ordpos + 1 : 0;
	L.flags = (en.meta & SP_EVAL_META_DEAD) ? 0u : SP_LITE_ALIVE;
	if (en.meta2 & SP_EVAL_M2_DONE0) L.flags |= SP_LITE_DONE;
	if (en.meta2 & SP_EVAL_M2_DATAREF0) L.flags |= SP_LITE_DATAREF;
	if (en.meta2 & SP_EVAL_M2_RESULT0) L.flags |= SP_LITE_RESULT;
	if (en.meta2 & SP_EVAL_M2_DREFEMIT0) L.flags |= SP_LITE_DREFEMIT;
	L.nitems = SP_EVAL_M2_NITEMS( en.meta2); L.signals = SP_EVAL_META_NKEY( en.meta); L.alive_tokens = 0;
}

/* fireSignal of live trigger j (:772-979) for a token at `pos`. Returns SP_LITE_TOOK | SP_LITE_EMITTED bits. */
#define SP_LITE_TOOK    1u	/* the rule has taken the event */
#define SP_LITE_EMITTED 2u	/* the rule has emitted its result with this event */
SP_MHD uint32_t sp_eval_lite_fire( const SpEvalEntry& en, uint32_t j, uint32_t pos, SpEvalLite& L)
{
	const uint32_t code = (en.codes >> (8 * j)) & 0xFFu, type = code >> 5, sig = code & 31u;
	bool take = false, match = false, finished = false;
	L.signals += 1;
	if (type == SP_SIG_ANY)
	{
		take = true;
		if (L.count > 0)
		{
			match = true; L.count -= 1; finished = (L.count == 0);
			if (L.endpos < pos + 1) L.endpos = pos + 1;
		}
	}
	else if (type == SP_SIG_DEL)
	{
		L.count = 0; L.value = 0; L.flags &= ~SP_LITE_ALIVE;
		return 0;
	}
	else
	{
		bool accept;
		if (type == SP_SIG_WITHIN) accept = ((1u << sig) & L.value) != 0 && L.endpos <= pos;
		else if (type == SP_SIG_SEQUENCE) accept = sig + 1 == L.value && L.endpos <= pos;
		else accept = sig + 1 == L.value && L.endpos == pos;
		if (accept)
		{
			L.endpos = pos + 1;
			if (type == SP_SIG_WITHIN) L.value &= ~(1u << sig); else L.value = sig;
			if (L.count > 0) {L.count -= 1; match = (L.count == 0);} else match = true;
			finished = (L.value == 0);
			take = true;
		}
	}
	uint32_t what = take ? SP_LITE_TOOK : 0u;
	if (take && ((SP_EVAL_M2_VARMASK( en.meta2) >> j) & 1u)) {L.nitems += 1; L.flags |= SP_LITE_DATAREF;}
	if (match)
	{
		if (!(L.flags & SP_LITE_DONE))
		{
			if (en.meta2 & SP_EVAL_M2_HASRES) {L.flags |= SP_LITE_RESULT | ((L.flags & SP_LITE_DATAREF) ? SP_LITE_DREFEMIT : 0u); what |= SP_LITE_EMITTED;}
			L.flags |= SP_LITE_DONE;
		}
		if (finished) L.flags &= ~SP_LITE_ALIVE;
	}
	return what;
}

/* one window token for the rule: expiry, live trigger count sample (:992), signals.
 * Returns (SP_LITE_TOOK | SP_LITE_EMITTED bits) | live trigger << 4 if a live trigger was signalled, else 0. */
SP_MHD uint32_t sp_eval_lite_step( const SpEvalEntry& en, uint32_t expiry, uint32_t id, uint32_t pos, SpEvalLite& L)
{
	if (!(L.flags & SP_LITE_ALIVE)) return 0;
	if (pos > expiry) {L.flags &= ~SP_LITE_ALIVE; return 0;}
	L.alive_tokens += 1;
	if (id != en.ev[0] && id != en.ev[1] && id != en.ev[2] && id != en.ev[3]) return 0;
	const uint32_t m = (id == en.ev[0] ? 1u : 0u) | (id == en.ev[1] ? 2u : 0u) | (id == en.ev[2] ? 4u : 0u) | (id == en.ev[3] ? 8u : 0u);
	if (m & (m - 1)) {L.flags = (L.flags | SP_LITE_FULL) & ~SP_LITE_ALIVE; return 0;}
	const uint32_t j = (uint32_t)sp_ctz( m);
	return sp_eval_lite_fire( en, j, pos, L) | (j << 4);
}

SP_MHD void sp_eval_lite_result( const SpEvalEntry& en, const SpEvalLite& L, SpEvalRule& R)
{
	sp_eval_rule_init( R);
	R.installed = 1; R.signals = L.signals;
	R.open = (uint64_t)L.alive_tokens * SP_EVAL_META_NTRIG( en.meta);
	R.has_result = (L.flags & SP_LITE_RESULT) ? 1 : 0;
	R.dataref_at_emit = (L.flags & SP_LITE_DREFEMIT) ? 1 : 0;
	R.nitems = L.nitems;
}

/* A rule keyed by follow event F has taken F with a variable `nkv` times at its installation (SP_EVAL_M2_NITEMS of an
 * entry with SP_EVAL_META_ITEMSFULL): each of these items refers to the item list of the rule Q that emitted F, and
 * the items of that list follow it in the result (gatherResultItems, src/patternMatcher.cpp:185-188) unless F has a
 * format. `extra` = number of those items: Q followed to its end from its own list entry. False: the full fold decides. */
SP_MHD bool sp_eval_lite_follow_items( const SpMatcherTables& T, const sp_lexem* toks, uint64_t tok, uint64_t te, const SpEvalEntry& en, const SpEvalFollow& F, uint32_t& extra)
{
	extra = 0;
	const uint32_t nkv = SP_EVAL_M2_NITEMS( en.meta2);
	if (!nkv || !F.sub || F.format) return true;
	const SpEvalEntry q = T.eval_entries[ F.qpt];
	if (q.meta & SP_EVAL_META_GENERIC) return false;
	SpEvalLite Q;
	const uint32_t spos = SP_MLDG( &toks[ tok].ordpos);
	sp_eval_lite_init( q, spos, Q);
	const uint32_t expiry = spos + SP_EVAL_META_RANGE( q.meta);
	for (uint64_t k=tok+1; k<te && (Q.flags & SP_LITE_ALIVE); ++k)
	{
		sp_eval_lite_step( q, expiry, SP_MLDG( &toks[ k].id), SP_MLDG( &toks[ k].ordpos), Q);
	}
	if (Q.flags & SP_LITE_FULL) return false;
	extra = nkv * Q.nitems;
	return true;
}

/* per entry form of what k_eval_count does with 32 rules in lock step (tests/hostsim). Returns false if the rule needs the full fold. */
SP_MHD bool sp_eval_entry_run( const SpMatcherTables& T, const SpEvalEntry& en, const sp_lexem* toks, uint64_t te, const SpEvalSite& site, SpEvalRule& R)
{
	if (en.meta & SP_EVAL_META_GENERIC) return false;
	SpEvalLite L;
	sp_eval_lite_init( en, site.data.start_ordpos, L);
	const uint32_t expiry = site.data.start_ordpos + SP_EVAL_META_RANGE( en.meta);
	for (uint64_t k=site.tok+1; k<te && (L.flags & SP_LITE_ALIVE); ++k)
	{
		sp_eval_lite_step( en, expiry, SP_MLDG( &toks[ k].id), SP_MLDG( &toks[ k].ordpos), L);
	}
	if (L.flags & SP_LITE_FULL) return false;
	uint32_t extra = 0;
	if ((L.flags & SP_LITE_RESULT) && (en.meta & SP_EVAL_META_ITEMSFULL))
	{
		if (!site.cpos || !sp_eval_lite_follow_items( T, toks, site.tok, te, en, T.eval_follows[ site.follow_begin + site.cpos - 1], extra)) return false;
	}
	sp_eval_lite_result( en, L, R);
	R.nitems += extra;
	return true;
}

/* ---- write pass from the compiled list entry: the rule is followed once more with the taken events recorded ---- */
#define SP_LITE_MAXTAKES 8
typedef struct SpEvalTrace
{
	uint32_t flags;				/* SP_LITE_* of the ended rule */
	uint32_t ntakes;			/* live triggers that took an event with a variable */
	uint32_t take_rel[ SP_LITE_MAXTAKES];	/* token of the event, relative to the first token of the document */
	uint32_t take_var[ SP_LITE_MAXTAKES];
	uint32_t first_rel, emit_rel;		/* token of the first event taken / of the event that emitted the result */
	uint32_t emit_end;			/* end_ordpos of the slot at the emission */
	uint32_t nitems;			/* items of the rule's list without the items of referenced lists */
} SpEvalTrace;

/* Follows the rule of list entry `en` (variables V) installed at `site` and records what the result needs.
 * False: not a rule the entry describes completely (the full fold does it). */
SP_MHD bool sp_eval_lite_trace( const SpEvalEntry& en, const SpEvalVars& V, const sp_lexem* toks, uint64_t tb, uint64_t te, const SpEvalSite& site, SpEvalTrace& tr)
{
	if ((en.meta & SP_EVAL_META_GENERIC) || V.nkeyvar > SP_EVAL_NKEYVAR) return false;
	SpEvalLite L;
	const uint32_t spos = site.data.start_ordpos;
	sp_eval_lite_init( en, spos, L);
	const uint32_t expiry = spos + SP_EVAL_META_RANGE( en.meta);
	const uint32_t none = 0xFFFFFFFFu;
	tr.ntakes = 0;
	tr.first_rel = (en.meta2 & SP_EVAL_M2_END0) ? (uint32_t)(site.tok - tb) : none;
	tr.emit_rel = (uint32_t)(site.tok - tb); tr.emit_end = spos + 1;
	bool overflow = false;
	for (uint64_t k=site.tok+1; k<te && (L.flags & SP_LITE_ALIVE); ++k)
	{
		const uint32_t what = sp_eval_lite_step( en, expiry, SP_MLDG( &toks[ k].id), SP_MLDG( &toks[ k].ordpos), L);
		if (what & SP_LITE_TOOK)
		{
			const uint32_t var = V.livevar[ (what >> 4) & 3u];
			if (tr.first_rel == none) tr.first_rel = (uint32_t)(k - tb);
			if (var)
			{
				if (tr.ntakes < SP_LITE_MAXTAKES) {tr.take_rel[ tr.ntakes] = (uint32_t)(k - tb); tr.take_var[ tr.ntakes] = var; tr.ntakes += 1;}
				else overflow = true;
			}
		}
		if (what & SP_LITE_EMITTED) {tr.emit_rel = (uint32_t)(k - tb); tr.emit_end = L.endpos;}
	}
	tr.flags = L.flags; tr.nitems = L.nitems;
	return !overflow && !(L.flags & SP_LITE_FULL) && (L.flags & SP_LITE_RESULT) && tr.first_rel != none;
}

SP_MHD void sp_eval_item_record( const sp_lexem& tok, uint32_t origseg, uint32_t variable, uint32_t format, sp_item& o);

/* Items of a traced rule in reading order (last appended first, src/patternMatcher.cpp:164-190): the events taken by live
 * triggers, latest first, then the key event once per key definition that took it with a variable; a key event that is
 * a follow event (F != 0) is followed by the items of the rule that emitted it. Returns the number of items written;
 * `ok` = false if the emitting rule is not one its list entry describes (nothing usable written). */
SP_MHD uint32_t sp_eval_lite_write_items( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
					  const SpEvalSite& site, const SpEvalVars& V, const SpEvalTrace& tr, const SpEvalFollow* F, sp_item* out, bool& ok)
{
	uint32_t n = 0;
	ok = true;
	for (uint32_t i=tr.ntakes; i>0; --i)
	{
		const uint64_t k = tb + tr.take_rel[ i-1];
		sp_eval_item_record( toks[ k], segs ? segs[ k] : 0, tr.take_var[ i-1], 0, out[ n++]);
	}
	const uint32_t siteseg = segs ? segs[ site.tok] : 0;
	for (uint32_t i=V.nkeyvar; i>0; --i)
	{
		sp_eval_item_record( toks[ site.tok], siteseg, V.keyvar[ i-1], F ? F->format : 0, out[ n++]);
		if (F && F->sub && !F->format)
		{
			const SpEvalEntry qen = T.eval_entries[ F->qpt];
			const SpEvalVars qv = T.eval_vars[ F->qpt];
			SpEvalSite qsite = site;
			qsite.cpos = 0; qsite.keyevent = SP_MLDG( &toks[ site.tok].id);
			SpEvalLite Q;
			if ((qen.meta & SP_EVAL_META_GENERIC) || qv.nkeyvar > SP_EVAL_NKEYVAR) {ok = false; return n;}
			// the rule that emitted the follow event, to its end: its live takes (latest first) ...
			const uint32_t spos = site.data.start_ordpos, expiry = spos + SP_EVAL_META_RANGE( qen.meta);
			sp_eval_lite_init( qen, spos, Q);
			uint32_t qrel[ SP_LITE_MAXTAKES], qvar[ SP_LITE_MAXTAKES], qn = 0;
			for (uint64_t k=site.tok+1; k<te && (Q.flags & SP_LITE_ALIVE); ++k)
			{
				const uint32_t what = sp_eval_lite_step( qen, expiry, SP_MLDG( &toks[ k].id), SP_MLDG( &toks[ k].ordpos), Q);
				const uint32_t var = (what & SP_LITE_TOOK) ? qv.livevar[ (what >> 4) & 3u] : 0;
				if (var)
				{
					if (qn >= SP_LITE_MAXTAKES) {ok = false; return n;}
					qrel[ qn] = (uint32_t)(k - tb); qvar[ qn] = var; ++qn;
				}
			}
			if (Q.flags & SP_LITE_FULL) {ok = false; return n;}
			for (uint32_t q=qn; q>0; --q)
			{
				const uint64_t k = tb + qrel[ q-1];
				sp_eval_item_record( toks[ k], segs ? segs[ k] : 0, qvar[ q-1], 0, out[ n++]);
			}
			// ... and its key event
			for (uint32_t q=qv.nkeyvar; q>0; --q) sp_eval_item_record( toks[ site.tok], siteseg, qv.keyvar[ q-1], 0, out[ n++]);
		}
	}
	return n;
}

SP_MHD void sp_eval_lite_result_record( const SpProgram& P, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, const SpEvalTrace& tr, uint32_t item_begin, uint32_t item_count, sp_result& o)
{
	const uint64_t kf = tb + tr.first_rel, ke = tb + tr.emit_rel;
	const sp_lexem first = toks[ kf], last = toks[ ke];
	o.name = P.resultHandle; o.format = P.formatHandle;
	o.ordpos = first.ordpos; o.ordend = tr.emit_end;
	o.origseg = segs ? segs[ kf] : 0; o.origpos = first.origpos;
	o.endseg = segs ? segs[ ke] : 0; o.endpos = last.origpos + last.origsize;
	o.item_begin = item_begin; o.item_count = item_count;
}

SP_MHD void sp_eval_item_record( const sp_lexem& tok, uint32_t origseg, uint32_t variable, uint32_t format, sp_item& o)
{
	o.variable = variable; o.format = format;
	o.ordpos = tok.ordpos; o.ordend = tok.ordpos + 1;
	o.origseg = origseg; o.origpos = tok.origpos;
	o.endseg = origseg; o.endpos = tok.origpos + tok.origsize;
}

/* gatherResultItems (src/patternMatcher.cpp:164-190) over the takes of a rule: the item list is read last appended
 * first; the items of a referenced sub list follow their item; a joined list was appended element by element in its
 * own reading order, which reverses it (joinEventData :755-770). Returns the number of items written. */
SP_MHD uint32_t sp_eval_write_items( const SpMatcherTables& T, const sp_lexem* toks, const uint32_t* segs, uint64_t tb, uint64_t te,
				     const SpEvalTake* takes, uint32_t ntakes, sp_item* out)
{
	uint32_t n = 0;
	for (uint32_t ti=ntakes; ti>0; --ti)
	{
		const SpEvalTake tk = takes[ ti-1];
		const uint64_t k = tb + tk.tokrel;
		const uint32_t seg = segs ? segs[ k] : 0;
		uint32_t format = 0, sub = 0, qpt = 0;
		if (tk.follow)
		{
			const SpEvalFollow F = T.eval_follows[ tk.follow-1];
			format = F.format; sub = F.sub; qpt = F.qpt;
		}
		if (tk.variable) sp_eval_item_record( toks[ k], seg, tk.variable, format, out[ n++]);
		if (tk.follow && sub && (!tk.variable || !format))
		{
			SpEvalSite qsite;
			sp_eval_site_token( T, toks[ k], seg, k, SP_EVAL_NOSLOT, qsite);
			SpEvalRule Q;
			SpEvalTake qtakes[ SP_EVAL_MAX_TAKES];
			sp_eval_fold<true,1>( T, toks, segs, tb, te, qsite, T.program_triggers[ qpt], tk.variable == 0, Q, qtakes);
			if (tk.variable)
			{
				for (uint32_t qi=Q.ntakes; qi>0; --qi)
				{
					const uint64_t qk = tb + qtakes[ qi-1].tokrel;
					sp_eval_item_record( toks[ qk], segs ? segs[ qk] : 0, qtakes[ qi-1].variable, 0, out[ n++]);
				}
			}
			else
			{
				for (uint32_t qi=0; qi<Q.ntakes; ++qi)
				{
					const uint64_t qk = tb + qtakes[ qi].tokrel;
					sp_eval_item_record( toks[ qk], segs ? segs[ qk] : 0, qtakes[ qi].variable, 0, out[ n++]);
				}
			}
		}
	}
	return n;
}

SP_MHD void sp_eval_result_record( const SpProgram& P, const SpEvalRule& R, uint32_t item_begin, uint32_t item_count, sp_result& o)
{
	o.name = P.resultHandle; o.format = P.formatHandle;
	o.ordpos = R.r_start_ordpos; o.ordend = R.r_end_ordpos;
	o.origseg = R.r_start_origseg; o.origpos = R.r_start_origpos;
	o.endseg = R.r_end_origseg; o.endpos = R.r_end_origpos;
	o.item_begin = item_begin; o.item_count = item_count;
}

/* items a result of rule R delivers */
SP_MHD uint32_t sp_eval_result_items( const SpEvalRule& R) {return (R.has_result && R.dataref_at_emit) ? R.nitems : 0;}

#endif
