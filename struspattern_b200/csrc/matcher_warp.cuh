// Warp-cooperative event processing of the level-2 matcher (CUDA only). One warp runs one document; inside one event
// of the follow list the work fans out over the 32 lanes:
//   warpFireHits     the triggers of the event's bucket that wait for the event fire on their rules concurrently (one
//                    lane per hit); what they append to the follow list, the results and the dispose list is committed
//                    in trigger array order by ballot + popc ranks
//   warpInstall      the programs keyed by the event are installed concurrently (one lane per program): rule records
//                    and trigger records come from the atomic pools, positions in the 16 trigger buckets and in the
//                    expiry ring are ranks in program order (packed per-bucket counters, prefix sum over the lanes)
//   warpDeactivate   rules are deactivated concurrently; their triggers are listed in order and then removed by 16
//                    lanes, one per bucket (swap-with-last removal only ever touches one bucket)
//   warpSetCurrentPos  rule expiry with warpDeactivate per ring slot
// Every order the reference makes observable (src/ruleMatcherAutomaton.cpp: bucket arrays :114-152, trigger lists,
// expiry lists :1066-1135, follow list / results / dispose list :981-1064) is the order of the sequential functions of
// matcher_device.cuh; cases the fan-out does not cover (two hits on one rule, alternative-key programs with replay,
// several key definitions, expiry beyond the ring, more than 7 trigger definitions) run those functions on lane 0.
#ifndef SP_MATCHER_WARP_CUH_INCLUDED
#define SP_MATCHER_WARP_CUH_INCLUDED
#include "matcher_device.cuh"

namespace sp {

#define SP_FULL 0xFFFFFFFFu

__device__ __forceinline__ unsigned int laneLt( unsigned int lane) {return (1u << lane) - 1u;}

// Appends of the lanes' signals in lane order: follow events, results, dispose entries (all lanes call).
__device__ __forceinline__ void warpCommit( SpArena& A, SpDocState& S, unsigned int lane, bool valid, const SpFireOut& out)
{
	const unsigned int fm = __ballot_sync( SP_FULL, valid && out.has_follow);
	const unsigned int rm = __ballot_sync( SP_FULL, valid && out.has_result);
	const unsigned int dm = __ballot_sync( SP_FULL, valid && out.dispose_rule);
	if (!(fm | rm | dm)) return;
	const uint32_t nf = S.nfollow, nr = S.nresults, nd = S.ndispose;
	__syncwarp();
	if (valid && out.has_follow)
	{
		uint32_t pos = nf + __popc( fm & laneLt( lane));
		if (pos < A.cap.follow) A.follow[ pos] = out.follow; else sp_fail( S, SP_MST_OVF_FOLLOW);
	}
	if (valid && out.has_result)
	{
		uint32_t pos = nr + __popc( rm & laneLt( lane));
		if (pos < A.cap.results) A.results[ pos] = out.result; else sp_fail( S, SP_MST_OVF_RESULTS);
	}
	if (valid && out.dispose_rule)
	{
		uint32_t pos = nd + __popc( dm & laneLt( lane));
		if (pos < A.cap.dispose) A.dispose[ pos] = out.dispose_rule; else sp_fail( S, SP_MST_OVF_DISPOSE);
	}
	if (lane == 0)
	{
		S.nfollow = min( nf + __popc( fm), A.cap.follow);
		S.nresults = min( nr + __popc( rm), A.cap.results);
		S.ndispose = min( nd + __popc( dm), A.cap.dispose);
	}
	__syncwarp();
}

// Fires the live triggers of bucket b that wait for event ev (n = fill of the bucket, read by the caller).
__device__ __forceinline__ void warpFireHits( SpArena& A, SpDocState& S, unsigned int lane, uint32_t ev, uint32_t b, uint32_t n, const SpEventData& data)
{
	const uint32_t* evs = A.bucket_event + (size_t)b * A.cap.triggers;
	const uint32_t* trs = A.bucket_trigger + (size_t)b * A.cap.triggers;
	for (uint32_t base=0; base<n; base+=32)
	{
		const uint32_t idx = base + lane;
		const bool hit = idx < n && evs[ idx] == ev;
		const unsigned int hits = __ballot_sync( SP_FULL, hit);
		if (!hits) continue;
		SpTrigger trg; trg.rule = 0x80000000u | lane; trg.sigtype = trg.sigval = trg.variable = 0;
		if (hit) trg = A.triggers[ trs[ idx] - 1];
		// two hits on one rule (e.g. sequence(A,A,A) keeps two live A triggers on one slot) must fire one after the other
		const unsigned int same = __match_any_sync( SP_FULL, trg.rule);
		if (__ballot_sync( SP_FULL, hit && same != (1u << lane)))
		{
			if (lane == 0)
			{
				for (unsigned int h = hits; h && !S.status; h &= h - 1) sp_fire_trigger_at( A, S, b, base + __ffs( h) - 1, data);
			}
			__syncwarp();
			continue;
		}
		SpFireOut out; out.dispose_rule = 0; out.has_follow = 0; out.has_result = 0;
		if (hit) sp_fire_core( A, S, trg.rule, trg.sigtype, trg.sigval, trg.variable, data, false, out);
		__syncwarp();
		warpCommit( A, S, lane, hit, out);
		if (lane == 0) S.signals += __popc( hits);
	}
	__syncwarp();
}

// Deactivates the rules list[0..n) in list order (reverse: from the back), optionally returning the rule records to
// their pool (rule expiry). All lanes call.
__device__ __forceinline__ void warpDeactivate( SpArena& A, SpDocState& S, unsigned int lane, const uint32_t* list, uint32_t n, bool reverse, bool freeRule)
{
	for (uint32_t base=0; base<n; base+=32)
	{
		const uint32_t i = base + lane;
		const bool has = i < n;
		const uint32_t r = has ? list[ reverse ? n - 1 - i : i] : (0x80000000u | lane);
		// a rule may be queued twice for deactivation: the first entry does the work
		const unsigned int same = __match_any_sync( SP_FULL, r);
		const bool first = has && !(same & laneLt( lane));
		const bool act = first && A.rules[ r-1].active;
		uint32_t cnt = 0;
		if (act) for (uint32_t t = A.rules[ r-1].trig_head; t; t = A.triggers[ t-1].next) ++cnt;
		uint32_t incl = cnt;
		#pragma unroll
		for (int ofs=1; ofs<32; ofs<<=1)
		{
			uint32_t up = __shfl_up_sync( SP_FULL, incl, ofs);
			if (lane >= (unsigned int)ofs) incl += up;
		}
		const uint32_t total = __shfl_sync( SP_FULL, incl, 31);
		if (act)
		{
			SpRule& rule = A.rules[ r-1];
			uint32_t k = incl - cnt;
			for (uint32_t t = rule.trig_head; t; t = A.triggers[ t-1].next) A.scratch[ k++] = (A.triggers[ t-1].link & 0xF0000000u) | t;
			rule.active = 0; rule.trig_head = 0;
			if (rule.dataref) {sp_dataref_release( A, S, rule.dataref); rule.dataref = 0;}
		}
		if (freeRule && first) sp_pool_free( &S.rule_nfree, A.free_rules, r);
		__syncwarp();
		if (total)
		{
			// swap-with-last removal in list order, one lane per bucket
			if (lane < SP_NOF_BUCKETS)
			{
				uint32_t* bev = A.bucket_event + (size_t)lane * A.cap.triggers;
				uint32_t* btr = A.bucket_trigger + (size_t)lane * A.cap.triggers;
				uint32_t fill = S.bucket_size[ lane];
				for (uint32_t k=0; k<total; ++k)
				{
					const uint32_t e = A.scratch[ k];
					if ((e >> 28) != lane) continue;
					const uint32_t t = e & 0x0FFFFFFFu;
					const uint32_t link = A.triggers[ t-1].link, pos = link & 0x0FFFFFFFu;
					--fill;
					if (pos != fill)
					{
						const uint32_t moved = btr[ fill];
						bev[ pos] = bev[ fill];
						btr[ pos] = moved;
						A.triggers[ moved-1].link = link;
					}
					sp_pool_free( &S.trig_nfree, A.free_triggers, t);
				}
				S.bucket_size[ lane] = fill;
			}
			if (lane == 31) S.nof_triggers -= total;
		}
		__syncwarp();
	}
}

// setCurrentPos (src/ruleMatcherAutomaton.cpp:1084-1135) with the ring slots disposed by the whole warp.
__device__ __forceinline__ void warpSetCurrentPos( SpArena& A, SpDocState& S, unsigned int lane, uint32_t pos)
{
	uint32_t cur = S.curpos;
	__syncwarp();
	if (cur == pos) return;
	uint32_t wcnt = 0;
	for (; wcnt < SP_DISPOSE_WINDOW && cur < pos && !S.status; ++wcnt, ++cur)
	{
		const uint32_t w = cur % SP_DISPOSE_WINDOW;
		const uint32_t nheap = S.nheap;
		__syncwarp();
		if (w == 0 && nheap)
		{
			int moved = 0;
			if (lane == 0) {S.curpos = cur; moved = sp_heap_to_window( A, S) ? 1 : 0;}
			if (__shfl_sync( SP_FULL, moved, 0)) wcnt = 0;
		}
		__syncwarp();
		const uint32_t nrules = S.win_count[ w];
		if (nrules)
		{
			warpDeactivate( A, S, lane, A.win_rules + (size_t)w * A.cap.rules, nrules, true, true);
			if (lane == 0) S.win_count[ w] = 0;
		}
		__syncwarp();
	}
	if (lane == 0)
	{
		S.curpos = cur;
		if (cur < pos)
		{
			S.curpos = pos;
			while (S.nheap && A.heap[ 0].pos < S.curpos)
			{
				uint32_t r = A.heap[ 0].rule;
				sp_rule_dispose( A, S, r);
				sp_heap_pop( A, S);
			}
		}
	}
	__syncwarp();
}


// 16 counters of 8 bits, one per trigger bucket
struct BucketCounts
{
	uint64_t lo, hi;
	__device__ __forceinline__ void add( uint32_t b) {if (b < 8) lo += 1ull << (8*b); else hi += 1ull << (8*(b-8));}
	__device__ __forceinline__ uint32_t get( uint32_t b) const {return (uint32_t)(((b < 8) ? (lo >> (8*b)) : (hi >> (8*(b-8)))) & 0xFF);}
};

// does the trigger definition get a live trigger when its program is installed by `keyevent`? (installProgram,
// src/ruleMatcherAutomaton.cpp:1200-1247, for programs with exactly one definition on the key event)
__device__ __forceinline__ bool installsTrigger( const SpTriggerDef& def, uint32_t keyevent, uint32_t slotcount)
{
	if (def.event != keyevent) return true;
	const bool needsInstall = (def.sigtype == SP_SIG_ANY && slotcount > 1);
	if (def.is_key || def.sigtype == SP_SIG_DEL) return needsInstall;
	return true;
}

// installEventPrograms: programs [es.prog_begin, +prog_count) keyed by `event`, one lane per program. All lanes call.
__device__ __forceinline__ void warpInstall( const SpMatcherTables& T, SpArena& A, SpDocState& S, unsigned int lane, uint32_t event, const SpEventSlot& es, const SpEventData& data)
{
	const uint32_t curpos = S.curpos;
	for (uint32_t base=0; base<es.prog_count && !S.status; base+=32)
	{
		__syncwarp();
		const uint32_t i = base + lane;
		const bool has = i < es.prog_count;
		SpProgramTrigger pt; pt.program = 0; pt.past_event = 0;
		SpProgram P; P.trig_count = 0; P.range = 0; P.trig_begin = 0;
		if (has) {pt = T.program_triggers[ es.prog_begin + i]; P = T.programs[ pt.program];}
		const uint32_t expiry = data.start_ordpos + P.range;
		const bool active = has && !(expiry < curpos);
		uint32_t nkey = 0, keydef = 0;
		if (active && P.trig_count <= 7)
		{
			for (uint32_t j=0; j<P.trig_count; ++j) if (T.trigger_defs[ P.trig_begin + j].event == event) {++nkey; keydef = P.trig_begin + j;}
		}
		const bool hard = active && (expiry >= curpos + SP_DISPOSE_WINDOW || P.trig_count > 7 || nkey != 1);
		if (__ballot_sync( SP_FULL, hard))
		{
			// expiry beyond the ring, several key definitions, many definitions: the sequential path, in list order
			if (lane == 0)
			{
				const uint32_t end = min( base + 32, es.prog_count);
				for (uint32_t k=base; k<end && !S.status; ++k) sp_install_program( T, A, S, event, T.program_triggers[ es.prog_begin + k], data);
			}
			__syncwarp();
			continue;
		}
		const unsigned int amask = __ballot_sync( SP_FULL, active);
		if (!amask) continue;
		// 1. rule record
		uint32_t r = 0;
		if (active)
		{
			r = sp_pool_alloc( &S.rule_nfree, &S.rule_top, A.free_rules, A.cap.rules);
			if (r)
			{
				SpRule& rule = A.rules[ r-1];
				rule.value = P.initsigval; rule.count = P.initcount & 0xFFFFu; rule.event = P.event;
				rule.resultHandle = P.resultHandle; rule.formatHandle = P.formatHandle;
				rule.start_ordpos = rule.end_ordpos = rule.start_origseg = rule.start_origpos = 0;
				rule.trig_head = 0; rule.dataref = 0; rule.active = 1; rule.done = 0; rule.next = 0;
			}
		}
		if (__ballot_sync( SP_FULL, active && !r)) {sp_fail( S, SP_MST_OVF_RULES); break;}
		// 2. expiry ring: position in the slot's list = rank among the lanes with the same slot
		{
			const uint32_t w = active ? expiry % SP_DISPOSE_WINDOW : (0x80000000u | lane);
			const unsigned int grp = __match_any_sync( SP_FULL, w);
			const uint32_t fill = active ? S.win_count[ w] : 0;
			const bool ovf = active && fill + __popc( grp) > A.cap.rules;
			__syncwarp();
			if (__ballot_sync( SP_FULL, ovf)) {sp_fail( S, SP_MST_OVF_RULES); break;}
			if (active)
			{
				A.win_rules[ (size_t)w * A.cap.rules + fill + __popc( grp & laneLt( lane))] = r;
				if (!(grp >> (lane + 1))) S.win_count[ w] = fill + __popc( grp);	// highest lane of the group
			}
		}
		// 2b. alternative-key program: replay of the logged latest occurrence of its original key event
		// (replayPastEvent, src/ruleMatcherAutomaton.cpp:1272-1334) on the triggers the rule is about to get, last
		// installed first. A rule that the replay finishes or that a later delimiter has cut off is deactivated at once;
		// its triggers would be the last of their buckets and leave them unchanged, so they are not installed at all.
		const uint32_t slotcount = P.initcount & 0xFFFFu;
		bool dead = false;
		uint32_t nreplayed = 0;
		if (active && pt.past_event)
		{
			SpEventSlot ps;
			if (sp_find_event( T, pt.past_event, ps) && ps.stopword)
			{
				const SpStopLog log = A.stoplog[ ps.stopword-1];
				if (log.stamp == S.stamp && !(log.data.start_ordpos + P.range < curpos))
				{
					bool cut = false;
					for (uint32_t j=P.trig_count; j>0 && !S.status; --j)
					{
						const SpTriggerDef def = T.trigger_defs[ P.trig_begin + j - 1];
						if (!installsTrigger( def, event, slotcount)) continue;
						if (def.sigtype == SP_SIG_DEL)
						{
							SpEventSlot ds;
							if (sp_find_event( T, def.event, ds) && ds.stopword)
							{
								const SpStopLog& dl = A.stoplog[ ds.stopword-1];
								if (dl.stamp == S.stamp && dl.timestmp > log.timestmp) cut = true;
							}
						}
						if (def.event == pt.past_event)
						{
							SpFireOut ro;
							sp_fire_core( A, S, r, def.sigtype, def.sigval, def.variable, log.data, true, ro);
							if (ro.dispose_rule) dead = true;
							++nreplayed;
						}
					}
					if (cut) dead = true;
				}
			}
		}
		const unsigned int replayed = __ballot_sync( SP_FULL, active && pt.past_event != 0);
		__syncwarp();
		if (dead)
		{
			// deactivateRule of a rule without live triggers (free-only step: no lane allocates here)
			SpRule& rule = A.rules[ r-1];
			rule.active = 0;
			if (rule.dataref) {sp_dataref_release( A, S, rule.dataref); rule.dataref = 0;}
		}
		__syncwarp();
		const bool live = active && !dead;
		// 3. live triggers: per bucket the positions follow program order, then definition order
		BucketCounts mine; mine.lo = mine.hi = 0;
		if (live)
		{
			for (uint32_t j=0; j<P.trig_count; ++j)
			{
				const SpTriggerDef def = T.trigger_defs[ P.trig_begin + j];
				if (installsTrigger( def, event, slotcount)) mine.add( sp_evhash( def.event));
			}
		}
		BucketCounts incl = mine;
		#pragma unroll
		for (int ofs=1; ofs<32; ofs<<=1)
		{
			uint64_t ulo = __shfl_up_sync( SP_FULL, incl.lo, ofs), uhi = __shfl_up_sync( SP_FULL, incl.hi, ofs);
			if (lane >= (unsigned int)ofs) {incl.lo += ulo; incl.hi += uhi;}
		}
		BucketCounts total; total.lo = __shfl_sync( SP_FULL, incl.lo, 31); total.hi = __shfl_sync( SP_FULL, incl.hi, 31);
		BucketCounts off; off.lo = incl.lo - mine.lo; off.hi = incl.hi - mine.hi;
		const uint32_t myfill = (lane < SP_NOF_BUCKETS) ? S.bucket_size[ lane] : 0;
		const bool bovf = lane < SP_NOF_BUCKETS && myfill + total.get( lane) > A.cap.triggers;
		__syncwarp();
		if (__ballot_sync( SP_FULL, bovf)) {sp_fail( S, SP_MST_OVF_TRIGGERS); break;}
		bool tfail = false;
		if (live)
		{
			SpRule& rule = A.rules[ r-1];
			for (uint32_t j=0; j<P.trig_count; ++j)
			{
				const SpTriggerDef def = T.trigger_defs[ P.trig_begin + j];
				if (!installsTrigger( def, event, slotcount)) continue;
				const uint32_t b = sp_evhash( def.event);
				const uint32_t pos = S.bucket_size[ b] + off.get( b);
				off.add( b);
				const uint32_t t = sp_pool_alloc( &S.trig_nfree, &S.trig_top, A.free_triggers, A.cap.triggers);
				if (!t) {tfail = true; break;}
				SpTrigger& rec = A.triggers[ t-1];
				rec.link = (b << 28) | pos; rec.rule = r; rec.sigtype = def.sigtype; rec.sigval = def.sigval; rec.variable = def.variable;
				rec.next = rule.trig_head;
				rule.trig_head = t;
				A.bucket_event[ (size_t)b * A.cap.triggers + pos] = def.event;
				A.bucket_trigger[ (size_t)b * A.cap.triggers + pos] = t;
			}
		}
		__syncwarp();
		if (__ballot_sync( SP_FULL, tfail)) {sp_fail( S, SP_MST_OVF_TRIGGERS); break;}
		if (lane < SP_NOF_BUCKETS) S.bucket_size[ lane] = myfill + total.get( lane);
		if (lane == 31)
		{
			uint32_t sum = 0;
			for (uint32_t b=0; b<SP_NOF_BUCKETS; ++b) sum += total.get( b);
			S.nof_triggers += sum;
			S.installed += __popc( amask);
			S.altkey_installed += __popc( replayed);
		}
		__syncwarp();
		// 4. the key definition fires on the new rule; appends in program order
		SpFireOut out; out.dispose_rule = 0; out.has_follow = 0; out.has_result = 0;
		if (live)
		{
			const SpTriggerDef def = T.trigger_defs[ keydef];
			sp_fire_core( A, S, r, def.sigtype, def.sigval, def.variable, data, false, out);
		}
		__syncwarp();
		warpCommit( A, S, lane, live, out);
		const unsigned int fired = __ballot_sync( SP_FULL, live);
		#pragma unroll
		for (int ofs=16; ofs>0; ofs>>=1) nreplayed += __shfl_down_sync( SP_FULL, nreplayed, ofs);
		if (lane == 0) S.signals += __popc( fired) + nreplayed;
	}
	__syncwarp();
}

} // namespace sp
#endif
