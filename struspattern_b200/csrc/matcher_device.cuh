// Per-document state machine of the level-2 matcher: the event loop, signal firing, program installation,
// rule expiry and result gathering, written as host+device inline functions over a flat per-warp arena.
// The CUDA kernel (matcher_kernels.cuh) runs one warp per document: the sequential steps below are executed
// by lane 0, the trigger scan is done by all 32 lanes; tests/hostsim runs the very same functions on the CPU
// against the reference while no GPU is around. The product only ever calls them from CUDA kernels.
//
// Reference behaviour restated here (file:line under the strusPattern tree):
//   EventTriggerTable add/remove/getTriggers (16 buckets, swap-remove)   src/ruleMatcherAutomaton.cpp:34-40, 114-152, 229-257
//   event data lists (append, join, reference counting)                 src/ruleMatcherAutomaton.cpp:710-770
//   fireSignal                                                           src/ruleMatcherAutomaton.cpp:772-979
//   doTransition                                                         src/ruleMatcherAutomaton.cpp:981-1064
//   defineDisposeRule / setCurrentPos (64 slot ring + heap)              src/ruleMatcherAutomaton.cpp:1066-1135
//   installProgram / replayPastEvent                                     src/ruleMatcherAutomaton.cpp:1168-1334
//   putInput / gatherResultItems / getCoveredFlags / fetchResults        src/patternMatcher.cpp:131-301
// Index values (rule, trigger, item numbers) are private to this implementation; what is kept identical to the
// reference is every ORDER that can be observed: bucket array order (append + swap-with-last removal), the
// last-in-first-out order of a rule's trigger list, of the expiry lists and of the item lists, the first-in-
// first-out follow list, and the order of the expiry heap.
#ifndef SP_MATCHER_DEVICE_CUH_INCLUDED
#define SP_MATCHER_DEVICE_CUH_INCLUDED
#include "matcher_tables.h"
#include "strus_pattern_b200.h"

#ifdef __CUDACC__
#define SP_MHD __device__ __forceinline__
#define SP_MLDG(p) __ldg(p)
#define SP_MHOSTDEV __host__ __device__ __forceinline__
#else
#define SP_MHOSTDEV inline
#define SP_MHD inline
#define SP_MLDG(p) (*(p))
#endif

/* document status values (sp_matcher_fetch: status[d]); 0 = ok */
#define SP_MST_RUNTIME    5u	/* = -SP_ERR_RUNTIME: the reference reports an error for this document */
#define SP_MST_OVERFLOW 100u	/* >= : a per-warp arena was too small; the host runs the document again with larger ones */
#define SP_MST_OVF_RULES    100u
#define SP_MST_OVF_TRIGGERS 101u
#define SP_MST_OVF_ITEMS    102u
#define SP_MST_OVF_DATAREFS 103u
#define SP_MST_OVF_RESULTS  104u
#define SP_MST_OVF_FOLLOW   105u
#define SP_MST_OVF_DISPOSE  106u
#define SP_MST_OVF_HEAP     107u
#define SP_MST_OVF_OUTPUT   108u
#define SP_MST_OVF_DEPTH    109u

#define SP_NOF_BUCKETS 16
#define SP_DISPOSE_WINDOW 64
#define SP_MAX_KEY_TRIGGER_DEFS 32
#define SP_GATHER_DEPTH 64

/* EventData, src/ruleMatcherAutomaton.hpp:200-218 */
typedef struct SpEventData
{
	uint32_t start_origseg, end_origseg, start_origpos, end_origpos, start_ordpos, end_ordpos, subdataref, formathandle;
} SpEventData;

/* Rule (:171) and its ActionSlot (:87) in one record: a rule owns exactly one slot. 64 bytes. */
typedef struct alignas(16) SpRule
{
	uint32_t value;
	uint32_t count;			/* 16 bit counter in the reference */
	uint32_t event;
	uint32_t resultHandle;
	uint32_t formatHandle;
	uint32_t start_ordpos;
	uint32_t end_ordpos;
	uint32_t start_origseg;
	uint32_t start_origpos;
	uint32_t trig_head;		/* list of live triggers (1-based trigger index, 0 = end) */
	uint32_t dataref;		/* 1-based event data reference or 0 */
	uint32_t active;		/* slot still exists (Rule::isActive) */
	uint32_t done;			/* the rule has emitted its result/follow event */
	uint32_t next;			/* expiry list link, free list link */
	uint32_t reserved[2];
} SpRule;

/* LinkedTrigger (:118) + the list element that chains the triggers of a rule. 32 bytes. */
typedef struct alignas(16) SpTrigger
{
	uint32_t link;			/* bucket << 28 | position in the bucket array */
	uint32_t rule;			/* 1-based */
	uint32_t sigtype;
	uint32_t sigval;
	uint32_t variable;
	uint32_t next;			/* next trigger of the same rule, free list link */
	uint32_t reserved[2];
} SpTrigger;

/* EventItem (:262) + list link. */
typedef struct SpItem
{
	uint32_t variable;
	SpEventData data;
	uint32_t next;
} SpItem;

typedef struct alignas(8) SpDataRef		/* EventDataReference, :241 */
{
	uint32_t items;			/* head of the item list (1-based) ; free list link */
	uint32_t refcount;
} SpDataRef;

typedef struct alignas(16) SpPendingResult		/* Result, :273 */
{
	uint32_t resultHandle, formatHandle, dataref, start_ordpos, end_ordpos, start_origseg, end_origseg, start_origpos, end_origpos;
	uint32_t eliminated;
	uint32_t out_index, item_begin;	/* place in the output of the document (filled when the document is finished) */
} SpPendingResult;

typedef struct SpFollow			/* EventStruct, :220 */
{
	SpEventData data;
	uint32_t event;
} SpFollow;

typedef struct SpStopLog		/* EventLog, :232, valid for the document `stamp` */
{
	SpEventData data;
	uint32_t timestmp;
	uint32_t stamp;
} SpStopLog;

typedef struct SpExpiry {uint32_t pos, rule;} SpExpiry;	/* DisposeEvent, :415 */

typedef struct SpArenaCaps
{
	uint32_t rules, triggers, items, datarefs, results, follow, dispose, heap;
} SpArenaCaps;

/* Mutable state of one document in flight (memory of one warp). */
typedef struct SpArena
{
	SpArenaCaps cap;
	SpRule* rules;
	SpTrigger* triggers;
	uint32_t* bucket_event;		/* [16][cap.triggers] event ids, scanned by the whole warp */
	uint32_t* bucket_trigger;	/* [16][cap.triggers] trigger index per array position */
	SpItem* items;
	SpDataRef* datarefs;
	SpPendingResult* results;
	SpFollow* follow;
	uint32_t* dispose;		/* rules to deactivate after the current event */
	SpExpiry* heap;
	SpStopLog* stoplog;		/* [nstopwords] */
	uint32_t* free_rules;		/* stacks of released record numbers, one per pool */
	uint32_t* free_triggers;
	uint32_t* free_items;
	uint32_t* free_datarefs;
	uint32_t* win_rules;		/* [64][cap.rules] expiry lists by position % 64, read backwards (last in, first out) */
	uint32_t* scratch;		/* [cap.triggers] ordered list of the triggers removed by one batch of deactivations */
} SpArena;

/* Scalars of the document in flight (shared memory of the warp). */
typedef struct SpDocState
{
	uint32_t status;
	uint32_t curpos;
	uint32_t timestmp;
	uint32_t stamp;			/* document sequence number of this warp (validates stoplog entries) */
	uint32_t nof_triggers;
	uint32_t nfollow, ndispose, nresults, nheap;
	int32_t rule_nfree, trig_nfree, item_nfree, dref_nfree;	/* fill of the free stacks */
	uint32_t rule_top, trig_top, item_top, dref_top;		/* records handed out so far (bump allocation) */
	uint32_t installed, altkey_installed, signals;
	uint64_t open_patterns;
	uint32_t bucket_size[ SP_NOF_BUCKETS];
	uint32_t win_count[ SP_DISPOSE_WINDOW];	/* fill of the expiry lists */
} SpDocState;

SP_MHD uint32_t sp_evhash( uint32_t a)
{
	a += ~(a>>5);
	a +=  (a<<3);
	a ^=  (a>>4);
	return a & (SP_NOF_BUCKETS-1);
}

SP_MHD void sp_doc_begin( SpDocState& S)
{
	uint32_t stamp = S.stamp + 1;
	S.status = 0; S.curpos = 0; S.timestmp = 0; S.stamp = stamp; S.nof_triggers = 0;
	S.nfollow = S.ndispose = S.nresults = S.nheap = 0;
	S.rule_nfree = S.trig_nfree = S.item_nfree = S.dref_nfree = 0;
	S.rule_top = S.trig_top = S.item_top = S.dref_top = 0;
	S.installed = S.altkey_installed = S.signals = 0; S.open_patterns = 0;
	for (int i=0; i<SP_NOF_BUCKETS; ++i) S.bucket_size[i] = 0;
	for (int i=0; i<SP_DISPOSE_WINDOW; ++i) S.win_count[i] = 0;
}

SP_MHD void sp_fail( SpDocState& S, uint32_t code) {if (!S.status) S.status = code;}

// ---- record pools. Record numbers are 1-based and private to this implementation (never observable). A pool hands out
// released records first (stack), then fresh ones. The counters are updated atomically so that the lanes of a warp can
// allocate concurrently in the phases that only allocate, and release concurrently in the phases that only release.
#ifdef __CUDACC__
#define SP_AADD( ptr, v) atomicAdd( (ptr), (v))
#else
template <typename T> inline T sp_host_fetch_add( T* ptr, T v) {T old = *ptr; *ptr = old + v; return old;}
#define SP_AADD( ptr, v) sp_host_fetch_add( (ptr), (v))
#endif

SP_MHD uint32_t sp_pool_alloc( int32_t* nfree, uint32_t* top, const uint32_t* stack, uint32_t cap)
{
	int32_t old = SP_AADD( nfree, (int32_t)-1);
	if (old > 0) return stack[ old-1];
	SP_AADD( nfree, (int32_t)1);
	uint32_t b = SP_AADD( top, 1u);
	if (b < cap) return b + 1;
	SP_AADD( top, (uint32_t)-1);
	return 0;
}
SP_MHD void sp_pool_free( int32_t* nfree, uint32_t* stack, uint32_t id)
{
	stack[ SP_AADD( nfree, (int32_t)1)] = id;
}

// ---- lookup of an event in the compiled tables
SP_MHD bool sp_find_event( const SpMatcherTables& T, uint32_t event, SpEventSlot& slot)
{
	uint32_t key = event + 1;
	for (uint32_t h = sp_event_slot_hash( event) & T.event_mask;; h = (h + 1) & T.event_mask)
	{
		uint32_t k = SP_MLDG( &T.event_slots[ h].key);
		if (k == key) {slot = T.event_slots[ h]; return true;}
		if (k == 0) return false;
	}
}

// ---- live triggers: EventTriggerTable
SP_MHD uint32_t sp_trigger_add( SpArena& A, SpDocState& S, uint32_t event, uint32_t rule, uint32_t sigtype, uint32_t sigval, uint32_t variable)
{
	uint32_t b = sp_evhash( event), n = S.bucket_size[ b];
	if (n >= A.cap.triggers) {sp_fail( S, SP_MST_OVF_TRIGGERS); return 0;}
	uint32_t t = sp_pool_alloc( &S.trig_nfree, &S.trig_top, A.free_triggers, A.cap.triggers);
	if (!t) {sp_fail( S, SP_MST_OVF_TRIGGERS); return 0;}
	SpTrigger& rec = A.triggers[ t-1];
	rec.link = (b << 28) | n; rec.rule = rule; rec.sigtype = sigtype; rec.sigval = sigval; rec.variable = variable; rec.next = 0;
	A.bucket_event[ (size_t)b * A.cap.triggers + n] = event;
	A.bucket_trigger[ (size_t)b * A.cap.triggers + n] = t;
	S.bucket_size[ b] = n + 1;
	S.nof_triggers += 1;
	return t;
}

SP_MHD void sp_trigger_remove( SpArena& A, SpDocState& S, uint32_t t)
{
	SpTrigger& rec = A.triggers[ t-1];
	uint32_t link = rec.link, b = link >> 28, i = link & 0x0FFFFFFFu;
	uint32_t last = S.bucket_size[ b] - 1;
	uint32_t* ev = A.bucket_event + (size_t)b * A.cap.triggers;
	uint32_t* tr = A.bucket_trigger + (size_t)b * A.cap.triggers;
	if (i != last)
	{
		uint32_t moved = tr[ last];
		ev[ i] = ev[ last];
		tr[ i] = moved;
		A.triggers[ moved-1].link = link;
	}
	S.bucket_size[ b] = last;
	S.nof_triggers -= 1;
	sp_pool_free( &S.trig_nfree, A.free_triggers, t);
}

SP_MHD uint32_t sp_trigger_event( const SpArena& A, uint32_t t)
{
	uint32_t link = A.triggers[ t-1].link;
	return A.bucket_event[ (size_t)(link >> 28) * A.cap.triggers + (link & 0x0FFFFFFFu)];
}

// ---- event data: item lists with reference counts
SP_MHD uint32_t sp_dataref_create( SpArena& A, SpDocState& S)
{
	uint32_t d = sp_pool_alloc( &S.dref_nfree, &S.dref_top, A.free_datarefs, A.cap.datarefs);
	if (!d) {sp_fail( S, SP_MST_OVF_DATAREFS); return 0;}
	A.datarefs[ d-1].items = 0; A.datarefs[ d-1].refcount = 1;
	return d;
}

SP_MHD void sp_dataref_release( SpArena& A, SpDocState& S, uint32_t d)
{
	SpDataRef& ref = A.datarefs[ d-1];
	if (ref.refcount > 1) {--ref.refcount; return;}
	if (ref.refcount == 0) {sp_fail( S, SP_MST_RUNTIME); return;}		// "illegal free of event data reference"
	ref.refcount = 0;
	// the items go back to their pool; like the reference, the sub data of the items stays referenced
	uint32_t it = ref.items;
	while (it)
	{
		uint32_t nx = A.items[ it-1].next;
		sp_pool_free( &S.item_nfree, A.free_items, it);
		it = nx;
	}
	ref.items = 0;
	sp_pool_free( &S.dref_nfree, A.free_datarefs, d);
}

SP_MHD void sp_dataref_push( SpArena& A, SpDocState& S, uint32_t d, uint32_t variable, const SpEventData& data)
{
	if (data.subdataref) SP_AADD( &A.datarefs[ data.subdataref-1].refcount, 1u);	// several lanes may take the same event
	uint32_t it = sp_pool_alloc( &S.item_nfree, &S.item_top, A.free_items, A.cap.items);
	if (!it) {sp_fail( S, SP_MST_OVF_ITEMS); return;}
	SpItem& rec = A.items[ it-1];
	rec.variable = variable; rec.data = data; rec.next = A.datarefs[ d-1].items;
	A.datarefs[ d-1].items = it;
}

// joinEventData: the source list is pushed element by element, which reverses it
SP_MHD void sp_dataref_join( SpArena& A, SpDocState& S, uint32_t dest, uint32_t src)
{
	for (uint32_t it = A.datarefs[ src-1].items; it && !S.status; it = A.items[ it-1].next)
	{
		SpItem rec = A.items[ it-1];
		sp_dataref_push( A, S, dest, rec.variable, rec.data);
	}
}

// ---- rules
SP_MHD void sp_dispose_add( SpArena& A, SpDocState& S, uint32_t rule)
{
	if (S.ndispose >= A.cap.dispose) {sp_fail( S, SP_MST_OVF_DISPOSE); return;}
	A.dispose[ S.ndispose++] = rule;
}

SP_MHD void sp_rule_deactivate( SpArena& A, SpDocState& S, uint32_t r)
{
	SpRule& rule = A.rules[ r-1];
	if (!rule.active) return;
	rule.active = 0;
	for (uint32_t t = rule.trig_head; t;)
	{
		uint32_t nx = A.triggers[ t-1].next;
		sp_trigger_remove( A, S, t);
		t = nx;
	}
	rule.trig_head = 0;
	if (rule.dataref)
	{
		sp_dataref_release( A, S, rule.dataref);
		rule.dataref = 0;
	}
}

SP_MHD void sp_rule_dispose( SpArena& A, SpDocState& S, uint32_t r)
{
	sp_rule_deactivate( A, S, r);
	sp_pool_free( &S.rule_nfree, A.free_rules, r);
}

// expiry heap: std::push_heap / std::pop_heap of libstdc++ with DisposeEvent::operator< (a < b iff a.pos > b.pos),
// restated so that rules expiring at the same position leave the heap in the same order as in the reference
SP_MHD void sp_heap_sift_up( SpExpiry* h, uint32_t hole, uint32_t top, SpExpiry value)
{
	while (hole > top)
	{
		uint32_t parent = (hole - 1) / 2;
		if (!(h[ parent].pos > value.pos)) break;
		h[ hole] = h[ parent];
		hole = parent;
	}
	h[ hole] = value;
}
SP_MHD void sp_heap_push( SpArena& A, SpDocState& S, uint32_t pos, uint32_t rule)
{
	if (S.nheap >= A.cap.heap) {sp_fail( S, SP_MST_OVF_HEAP); return;}
	SpExpiry v; v.pos = pos; v.rule = rule;
	sp_heap_sift_up( A.heap, S.nheap, 0, v);
	S.nheap += 1;
}
SP_MHD void sp_heap_pop( SpArena& A, SpDocState& S)
{
	uint32_t n = S.nheap - 1;	// new size; the old last element is re-inserted from the top
	SpExpiry* h = A.heap;
	if (n > 0)
	{
		SpExpiry value = h[ n];
		uint32_t hole = 0, child = 0;
		while (child < (n - 1) / 2)
		{
			child = 2 * (child + 1);
			if (h[ child].pos > h[ child-1].pos) --child;
			h[ hole] = h[ child];
			hole = child;
		}
		if ((n & 1) == 0 && child == (n - 2) / 2)
		{
			child = 2 * (child + 1);
			h[ hole] = h[ child-1];
			hole = child - 1;
		}
		sp_heap_sift_up( h, hole, 0, value);
	}
	S.nheap = n;
}

SP_MHD void sp_window_push( SpArena& A, SpDocState& S, uint32_t w, uint32_t r)
{
	uint32_t n = S.win_count[ w];
	if (n >= A.cap.rules) {sp_fail( S, SP_MST_OVF_RULES); return;}
	A.win_rules[ (size_t)w * A.cap.rules + n] = r;
	S.win_count[ w] = n + 1;
}

// defineDisposeRule
SP_MHD void sp_expiry_define( SpArena& A, SpDocState& S, uint32_t pos, uint32_t r)
{
	if (pos < S.curpos) {sp_fail( S, SP_MST_RUNTIME); return;}
	if (pos < S.curpos + SP_DISPOSE_WINDOW) sp_window_push( A, S, pos % SP_DISPOSE_WINDOW, r);
	else sp_heap_push( A, S, pos, r);
}

// Moves the heap entries that expire within the next 64 positions into the ring (start of a new round of the ring).
// Returns true if any was moved.
SP_MHD bool sp_heap_to_window( SpArena& A, SpDocState& S)
{
	bool moved = false;
	while (S.nheap && A.heap[ 0].pos < S.curpos + SP_DISPOSE_WINDOW && !S.status)
	{
		moved = true;
		SpExpiry top = A.heap[ 0];
		sp_window_push( A, S, top.pos % SP_DISPOSE_WINDOW, top.rule);
		sp_heap_pop( A, S);
	}
	return moved;
}

// setCurrentPos: disposes every rule whose last position lies in front of `pos`
SP_MHD void sp_set_current_pos( SpArena& A, SpDocState& S, uint32_t pos)
{
	if (S.curpos == pos) return;
	uint32_t wcnt = 0;
	for (; wcnt < SP_DISPOSE_WINDOW && S.curpos < pos; ++wcnt, ++S.curpos)
	{
		uint32_t w = S.curpos % SP_DISPOSE_WINDOW;
		if (w == 0 && sp_heap_to_window( A, S)) wcnt = 0;
		// the list is visited last in, first out
		const uint32_t* list = A.win_rules + (size_t)w * A.cap.rules;
		for (uint32_t n = S.win_count[ w]; n > 0; --n) sp_rule_dispose( A, S, list[ n-1]);
		S.win_count[ w] = 0;
	}
	if (S.curpos < pos)
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

SP_MHD void sp_follow_add( SpArena& A, SpDocState& S, const SpEventData& data, uint32_t event)
{
	if (S.nfollow >= A.cap.follow) {sp_fail( S, SP_MST_OVF_FOLLOW); return;}
	A.follow[ S.nfollow].data = data; A.follow[ S.nfollow].event = event;
	S.nfollow += 1;
}

// What one signal appends to the ordered lists of the document (dispose list, follow list, results). The slot
// update itself is private to the rule; keeping the appends apart lets the lanes of a warp fire the hits of one event
// on different rules concurrently and then commit their appends in trigger order.
typedef struct SpFireOut
{
	uint32_t dispose_rule;		/* rule to queue for deactivation or 0 */
	uint32_t has_follow, has_result;
	SpFollow follow;
	SpPendingResult result;
} SpFireOut;

// fireSignal without the appends and without the signal counter. `replay`: called from the replay of a logged stop
// word, where a follow event is an error of the rule set ("encountered past trigger with follow").
SP_MHD void sp_fire_core( SpArena& A, SpDocState& S, uint32_t r, uint32_t sigtype, uint32_t sigval, uint32_t variable, const SpEventData& data, bool replay, SpFireOut& out)
{
	SpRule& rule = A.rules[ r-1];
	bool match = false, take = false, finished = false;
	out.dispose_rule = 0; out.has_follow = 0; out.has_result = 0;
	switch (sigtype)
	{
		case SP_SIG_ANY:
			take = true;
			if (rule.count > 0)
			{
				match = true;
				rule.count -= 1;
				finished = (rule.count == 0);
				if (rule.end_ordpos < data.end_ordpos) rule.end_ordpos = data.end_ordpos;
			}
			break;
		case SP_SIG_AND:
			if (rule.count > 0)
			{
				if (!rule.value)
				{
					rule.value = data.start_ordpos;
					if (rule.end_ordpos > data.end_ordpos) rule.end_ordpos = data.end_ordpos;
				}
				if (rule.value == data.start_ordpos)
				{
					match = true;
					rule.count -= 1;
					finished = (rule.count == 0);
					take = true;
				}
			}
			break;
		case SP_SIG_SEQUENCE:
		case SP_SIG_SEQUENCE_IMM:
		case SP_SIG_WITHIN:
		{
			bool accept;
			if (sigtype == SP_SIG_WITHIN) accept = (sigval & rule.value) != 0 && rule.end_ordpos <= data.start_ordpos;
			else if (sigtype == SP_SIG_SEQUENCE) accept = sigval == rule.value && rule.end_ordpos <= data.start_ordpos;
			else accept = sigval == rule.value && rule.end_ordpos == data.start_ordpos;
			if (accept)
			{
				rule.end_ordpos = data.end_ordpos;
				if (sigtype == SP_SIG_WITHIN) rule.value &= ~sigval; else rule.value = sigval - 1;
				if (rule.count > 0)
				{
					rule.count -= 1;
					match = (rule.count == 0);
				}
				else match = true;
				finished = (rule.value == 0);
				take = true;
			}
			break;
		}
		case SP_SIG_DEL:
			rule.count = 0;
			rule.value = 0;
			out.dispose_rule = r;
			return;
		default:
			break;
	}
	if (take)
	{
		if (variable)
		{
			if (!rule.dataref) rule.dataref = sp_dataref_create( A, S);
			if (rule.dataref) sp_dataref_push( A, S, rule.dataref, variable, data);
		}
		else if (data.subdataref)
		{
			if (!rule.dataref) rule.dataref = sp_dataref_create( A, S);
			if (rule.dataref) sp_dataref_join( A, S, rule.dataref, data.subdataref);
		}
		if (rule.start_ordpos == 0)
		{
			rule.start_ordpos = data.start_ordpos;
			rule.start_origseg = data.start_origseg;
			rule.start_origpos = data.start_origpos;
		}
		else if (rule.start_ordpos > data.start_ordpos)
		{
			rule.start_ordpos = data.start_ordpos;
			if (rule.start_origseg > data.start_origseg || (rule.start_origseg == data.start_origseg && rule.start_origpos > data.start_origpos))
			{
				rule.start_origseg = data.start_origseg;
				rule.start_origpos = data.start_origpos;
			}
		}
	}
	if (match)
	{
		if (!rule.done)
		{
			if (rule.event)
			{
				if (replay) {sp_fail( S, SP_MST_RUNTIME); return;}
				SpEventData& fd = out.follow.data;
				fd.start_origseg = rule.start_origseg; fd.start_origpos = rule.start_origpos;
				fd.end_origseg = data.end_origseg; fd.end_origpos = data.end_origpos;
				fd.start_ordpos = rule.start_ordpos; fd.end_ordpos = rule.end_ordpos;
				fd.subdataref = rule.dataref; fd.formathandle = rule.formatHandle;
				out.follow.event = rule.event;
				out.has_follow = 1;
				if (rule.dataref) SP_AADD( &A.datarefs[ rule.dataref-1].refcount, 1u);
			}
			if (rule.resultHandle)
			{
				SpPendingResult& res = out.result;
				res.resultHandle = rule.resultHandle; res.formatHandle = rule.formatHandle; res.dataref = rule.dataref;
				res.start_ordpos = rule.start_ordpos; res.end_ordpos = rule.end_ordpos;
				res.start_origseg = rule.start_origseg; res.start_origpos = rule.start_origpos;
				res.end_origseg = data.end_origseg; res.end_origpos = data.end_origpos;
				res.eliminated = 0;
				out.has_result = 1;
				if (rule.dataref) SP_AADD( &A.datarefs[ rule.dataref-1].refcount, 1u);
			}
			rule.done = 1;
		}
		if (finished) out.dispose_rule = r;
	}
}

// appends of one signal in the order of the reference: follow event, result, dispose entry
SP_MHD void sp_fire_commit( SpArena& A, SpDocState& S, const SpFireOut& out)
{
	if (out.has_follow) sp_follow_add( A, S, out.follow.data, out.follow.event);
	if (out.has_result)
	{
		if (S.nresults >= A.cap.results) {sp_fail( S, SP_MST_OVF_RESULTS); return;}
		A.results[ S.nresults++] = out.result;
	}
	if (out.dispose_rule) sp_dispose_add( A, S, out.dispose_rule);
}

SP_MHD void sp_fire_signal( SpArena& A, SpDocState& S, uint32_t r, uint32_t sigtype, uint32_t sigval, uint32_t variable, const SpEventData& data, bool replay)
{
	SpFireOut out;
	S.signals += 1;
	sp_fire_core( A, S, r, sigtype, sigval, variable, data, replay, out);
	sp_fire_commit( A, S, out);
}

// replayPastEvent: an alternative-key program was installed by a later element; the logged latest occurrence of
// its original key event (a stop word) is fired on the new rule
SP_MHD void sp_replay_past_event( const SpMatcherTables& T, SpArena& A, SpDocState& S, uint32_t past_event, uint32_t r, uint32_t range)
{
	SpEventSlot es;
	if (!sp_find_event( T, past_event, es) || !es.stopword) return;
	const SpStopLog& log = A.stoplog[ es.stopword-1];
	if (log.stamp != S.stamp || log.data.start_ordpos + range < S.curpos) return;
	SpEventData data = log.data;
	uint32_t logtime = log.timestmp;
	// rules finished by the replay are deactivated at its end (the list is local to the replay in the reference)
	uint32_t dispose_base = S.ndispose;
	uint32_t delEvents[ 32]; uint32_t ndel = 0; bool delOverflow = false;
	for (uint32_t t = A.rules[ r-1].trig_head; t && !S.status; t = A.triggers[ t-1].next)
	{
		SpTrigger trg = A.triggers[ t-1];
		uint32_t tev = sp_trigger_event( A, t);
		if (trg.sigtype == SP_SIG_DEL)
		{
			if (ndel < 32) delEvents[ ndel++] = tev; else delOverflow = true;
		}
		if (tev == past_event) sp_fire_signal( A, S, r, trg.sigtype, trg.sigval, trg.variable, data, true);
	}
	if (delOverflow) {sp_fail( S, SP_MST_OVF_DEPTH); return;}
	for (uint32_t k=0; k<ndel; ++k)
	{
		SpEventSlot ds;
		if (sp_find_event( T, delEvents[ k], ds) && ds.stopword)
		{
			const SpStopLog& dl = A.stoplog[ ds.stopword-1];
			if (dl.stamp == S.stamp && dl.timestmp > logtime) {sp_rule_deactivate( A, S, r); break;}
		}
	}
	for (uint32_t k=dispose_base; k<S.ndispose; ++k) sp_rule_deactivate( A, S, A.dispose[ k]);
	S.ndispose = dispose_base;
}

// installProgram: a key event creates a new rule instance of the program
SP_MHD void sp_install_program( const SpMatcherTables& T, SpArena& A, SpDocState& S, uint32_t keyevent, const SpProgramTrigger& pt, const SpEventData& data)
{
	const SpProgram P = T.programs[ pt.program];
	if (data.start_ordpos + P.range < S.curpos) return;	// the rule cannot match anymore
	uint32_t r = sp_pool_alloc( &S.rule_nfree, &S.rule_top, A.free_rules, A.cap.rules);
	if (!r) {sp_fail( S, SP_MST_OVF_RULES); return;}
	SpRule& rule = A.rules[ r-1];
	rule.value = P.initsigval; rule.count = P.initcount & 0xFFFFu; rule.event = P.event;
	rule.resultHandle = P.resultHandle; rule.formatHandle = P.formatHandle;
	rule.start_ordpos = rule.end_ordpos = rule.start_origseg = rule.start_origpos = 0;
	rule.trig_head = 0; rule.dataref = 0; rule.active = 1; rule.done = 0; rule.next = 0;
	sp_expiry_define( A, S, data.start_ordpos + P.range, r);
	if (S.status) return;

	uint32_t keydefs[ SP_MAX_KEY_TRIGGER_DEFS]; uint32_t nkeydefs = 0;
	bool hasKeyEvent = false;
	for (uint32_t di = P.trig_begin, de = P.trig_begin + P.trig_count; di < de; ++di)
	{
		const SpTriggerDef def = T.trigger_defs[ di];
		bool install = true;
		if (def.event == keyevent)
		{
			if (nkeydefs >= SP_MAX_KEY_TRIGGER_DEFS) {sp_fail( S, SP_MST_RUNTIME); return;}	// "pattern with too many identical key events"
			keydefs[ nkeydefs++] = di;
			bool needsInstall = (def.sigtype == SP_SIG_ANY && rule.count > 1);
			if (def.is_key && !hasKeyEvent) {hasKeyEvent = true; install = needsInstall;}
			else if (def.sigtype == SP_SIG_DEL) install = needsInstall;
		}
		if (install)
		{
			uint32_t t = sp_trigger_add( A, S, def.event, r, def.sigtype, def.sigval, def.variable);
			if (!t) return;
			A.triggers[ t-1].next = rule.trig_head;
			rule.trig_head = t;
		}
	}
	S.installed += 1;
	if (pt.past_event)
	{
		S.altkey_installed += 1;
		sp_replay_past_event( T, A, S, pt.past_event, r, P.range);
		if (S.status) return;
	}
	if (nkeydefs && rule.active)
	{
		for (uint32_t k=0; k<nkeydefs && !S.status; ++k)
		{
			const SpTriggerDef def = T.trigger_defs[ keydefs[ k]];
			sp_fire_signal( A, S, r, def.sigtype, def.sigval, def.variable, data, false);
		}
	}
}

// Second half of the processing of one event of the follow list (the part behind the trigger scan):
// program installation, deferred deactivation, stop word log / release of the event's data.
SP_MHD void sp_event_finish( const SpMatcherTables& T, SpArena& A, SpDocState& S, uint32_t fi)
{
	SpFollow follow = A.follow[ fi];
	SpEventSlot es;
	bool known = sp_find_event( T, follow.event, es);
	if (known)
	{
		for (uint32_t pi = es.prog_begin, pe = es.prog_begin + es.prog_count; pi < pe && !S.status; ++pi)
		{
			sp_install_program( T, A, S, follow.event, T.program_triggers[ pi], follow.data);
		}
	}
	for (uint32_t k=0; k<S.ndispose; ++k) sp_rule_deactivate( A, S, A.dispose[ k]);
	S.ndispose = 0;
	if (known && es.stopword)
	{
		SpStopLog& log = A.stoplog[ es.stopword-1];
		log.data = follow.data; log.timestmp = ++S.timestmp; log.stamp = S.stamp;
	}
	else if (follow.data.subdataref) sp_dataref_release( A, S, follow.data.subdataref);
}

// putInput up to the start of the event loop, in three steps: order and range checks, rule expiry, first entry of the
// follow list. Returns false if the token is rejected (the document fails).
SP_MHD bool sp_input_check( SpDocState& S, const sp_lexem& tok, uint32_t origseg)
{
	if (S.curpos > tok.ordpos) {sp_fail( S, SP_MST_RUNTIME); return false;}	// "term events not fed in ascending order"
	if (S.curpos == tok.ordpos && (tok.origsize >= 0x7FFFFFFFu || origseg >= 0x7FFFFFFFu || tok.origpos >= 0x7FFFFFFFu)) {sp_fail( S, SP_MST_RUNTIME); return false;}
	if (tok.id >= (1u<<29)) {sp_fail( S, SP_MST_RUNTIME); return false;}	// "event handle out of range"
	return true;
}
SP_MHD void sp_input_queue( SpArena& A, SpDocState& S, const sp_lexem& tok, uint32_t origseg)
{
	SpEventData data;
	data.start_origseg = origseg; data.start_origpos = tok.origpos; data.end_origseg = origseg; data.end_origpos = tok.origpos + tok.origsize;
	data.start_ordpos = tok.ordpos; data.end_ordpos = tok.ordpos + 1; data.subdataref = 0; data.formathandle = 0;
	S.open_patterns += S.nof_triggers;
	S.nfollow = 0;
	sp_follow_add( A, S, data, tok.id);
}
SP_MHD void sp_input_begin( SpArena& A, SpDocState& S, const sp_lexem& tok, uint32_t origseg)
{
	// the reference advances the position before it looks at the event id (src/patternMatcher.cpp:136-156)
	if (S.curpos > tok.ordpos) {sp_fail( S, SP_MST_RUNTIME); return;}
	if (S.curpos < tok.ordpos) sp_set_current_pos( A, S, tok.ordpos);
	else if (tok.origsize >= 0x7FFFFFFFu || origseg >= 0x7FFFFFFFu || tok.origpos >= 0x7FFFFFFFu) {sp_fail( S, SP_MST_RUNTIME); return;}
	if (tok.id >= (1u<<29)) {sp_fail( S, SP_MST_RUNTIME); return;}
	sp_input_queue( A, S, tok, origseg);
}

// One signal of the trigger scan: the trigger at position i of bucket b waits for the current event.
SP_MHD void sp_fire_trigger_at( SpArena& A, SpDocState& S, uint32_t b, uint32_t i, const SpEventData& data)
{
	uint32_t t = A.bucket_trigger[ (size_t)b * A.cap.triggers + i];
	SpTrigger trg = A.triggers[ t-1];
	sp_fire_signal( A, S, trg.rule, trg.sigtype, trg.sigval, trg.variable, data, false);
}

// ---- results: fetchResults
// getCoveredFlags (option "exclusive"): results covered by another result that starts close enough are eliminated
SP_MHD void sp_mark_covered( const SpMatcherTables& T, SpArena& A, SpDocState& S)
{
	for (uint32_t ai=0; ai<S.nresults; ++ai)
	{
		const SpPendingResult a = A.results[ ai];
		for (uint32_t ni=ai; ni<S.nresults; ++ni)
		{
			const SpPendingResult f = A.results[ ni];
			if (f.start_origseg > a.end_origseg || f.start_origpos >= a.end_origpos + T.max_result_size) break;
			bool equal = f.end_origseg == a.end_origseg && f.end_origpos == a.end_origpos && f.start_origseg == a.start_origseg && f.start_origpos == a.start_origpos;
			if (f.start_origseg <= a.start_origseg && f.start_origpos <= a.start_origpos && f.end_origseg >= a.end_origseg && f.end_origpos >= a.end_origpos)
			{
				if (!equal) A.results[ ai].eliminated = 1;
			}
			if (f.start_origseg >= a.start_origseg && f.start_origpos >= a.start_origpos && f.end_origseg <= a.end_origseg && f.end_origpos <= a.end_origpos)
			{
				if (!equal) A.results[ ni].eliminated = 1;
			}
		}
	}
}

// gatherResultItems without recursion: items of `dataref` in list order, the items of a sub expression directly behind
// the item that refers to it. A sub expression without a format of its own joins the list of its parent
// (src/patternMatcher.cpp:185-188); the items of one WITH a format are what its value string is made of (:172-181): they
// follow their owner as nested records, their number in the high half of the owner's `format` (SP_ITEM_NESTED in
// include/strus_pattern_b200.h), so that the host side can map them into the value and hide them from the item list.
// out == 0: count only. Returns the number of records.
SP_MHD uint32_t sp_gather_items( const SpArena& A, SpDocState& S, uint32_t dataref, sp_item* out)
{
	uint32_t stack[ SP_GATHER_DEPTH]; uint32_t owner[ SP_GATHER_DEPTH]; int depth = 0;
	uint32_t n = 0;
	uint32_t it = A.datarefs[ dataref-1].items;
	for (;;)
	{
		if (!it)
		{
			if (depth == 0) break;
			--depth;
			it = stack[ depth];
			if (owner[ depth] != 0xFFFFFFFFu)
			{
				const uint32_t nested = n - owner[ depth] - 1;
				if (nested > 0xFFFFu) {sp_fail( S, SP_MST_OVF_DEPTH); return n;}
				if (out) out[ owner[ depth]].format |= nested << 16;
			}
			continue;
		}
		const SpItem& rec = A.items[ it-1];
		if (out)
		{
			sp_item o;
			o.variable = rec.variable; o.format = rec.data.formathandle;
			o.ordpos = rec.data.start_ordpos; o.ordend = rec.data.end_ordpos;
			o.origseg = rec.data.start_origseg; o.origpos = rec.data.start_origpos;
			o.endseg = rec.data.end_origseg; o.endpos = rec.data.end_origpos;
			out[ n] = o;
		}
		++n;
		uint32_t nx = rec.next;
		if (rec.data.subdataref)
		{
			if (depth >= SP_GATHER_DEPTH) {sp_fail( S, SP_MST_OVF_DEPTH); return n;}
			stack[ depth] = nx; owner[ depth] = rec.data.formathandle ? n - 1 : 0xFFFFFFFFu;
			++depth;
			it = A.datarefs[ rec.data.subdataref-1].items;
		}
		else it = nx;
	}
	return n;
}

SP_MHD void sp_result_record( const SpPendingResult& res, uint32_t item_begin, uint32_t item_count, sp_result& o)
{
	o.name = res.resultHandle; o.format = res.formatHandle;
	o.ordpos = res.start_ordpos; o.ordend = res.end_ordpos;
	o.origseg = res.start_origseg; o.origpos = res.start_origpos;
	o.endseg = res.end_origseg; o.endpos = res.end_origpos;
	o.item_begin = item_begin; o.item_count = item_count;
}

SP_MHD void sp_doc_stats( const SpDocState& S, uint64_t nevents, sp_matcher_stats& st)
{
	st.nof_programs_installed = S.installed;
	st.nof_altkey_programs_installed = S.altkey_installed;
	st.nof_signals_fired = S.signals;
	st.nof_open_patterns = S.open_patterns;
	st.nof_events = nevents;
}


// ---- output of one document: results that survive the "exclusive" filter and their items
// counts what sp_doc_write_output will write
SP_MHD void sp_doc_count_output( const SpMatcherTables& T, SpArena& A, SpDocState& S, uint32_t& nres, uint32_t& nitems)
{
	nres = 0; nitems = 0;
	if (T.exclusive) sp_mark_covered( T, A, S);
	for (uint32_t ri=0; ri<S.nresults && !S.status; ++ri)
	{
		const SpPendingResult& res = A.results[ ri];
		if (res.eliminated) continue;
		++nres;
		// the items of a result with a format string are consumed by the formatter (src/patternMatcher.cpp:253-262);
		// they are still delivered here (the host side layer maps them into the value string)
		if (res.dataref) nitems += sp_gather_items( A, S, res.dataref, 0);
	}
}

SP_MHD void sp_doc_write_output( SpArena& A, SpDocState& S, sp_result* out_results, sp_item* out_items)
{
	uint32_t nres = 0, nitems = 0;
	for (uint32_t ri=0; ri<S.nresults && !S.status; ++ri)
	{
		const SpPendingResult& res = A.results[ ri];
		if (res.eliminated) continue;
		uint32_t cnt = res.dataref ? sp_gather_items( A, S, res.dataref, out_items + nitems) : 0;
		sp_result_record( res, nitems, cnt, out_results[ nres]);
		nitems += cnt;
		++nres;
	}
}

// ---- memory layout of one arena (byte offsets of the components; every component 128 byte aligned)
typedef struct SpArenaLayout
{
	SpArenaCaps cap;
	uint32_t nstopwords;
	uint64_t rules, triggers, bucket_event, bucket_trigger, items, datarefs, results, follow, dispose, heap, stoplog;
	uint64_t free_rules, free_triggers, free_items, free_datarefs, win_rules, scratch;
	uint64_t bytes;
} SpArenaLayout;

SP_MHOSTDEV uint64_t sp_align128( uint64_t v) {return (v + 127) & ~(uint64_t)127;}

SP_MHOSTDEV void sp_arena_layout( const SpArenaCaps& cap, uint32_t nstopwords, SpArenaLayout& L)
{
	L.cap = cap; L.nstopwords = nstopwords;
	uint64_t o = 0;
	L.rules = o; o = sp_align128( o + (uint64_t)cap.rules * sizeof(SpRule));
	L.triggers = o; o = sp_align128( o + (uint64_t)cap.triggers * sizeof(SpTrigger));
	L.bucket_event = o; o = sp_align128( o + (uint64_t)cap.triggers * SP_NOF_BUCKETS * sizeof(uint32_t));
	L.bucket_trigger = o; o = sp_align128( o + (uint64_t)cap.triggers * SP_NOF_BUCKETS * sizeof(uint32_t));
	L.items = o; o = sp_align128( o + (uint64_t)cap.items * sizeof(SpItem));
	L.datarefs = o; o = sp_align128( o + (uint64_t)cap.datarefs * sizeof(SpDataRef));
	L.results = o; o = sp_align128( o + (uint64_t)cap.results * sizeof(SpPendingResult));
	L.follow = o; o = sp_align128( o + (uint64_t)cap.follow * sizeof(SpFollow));
	L.dispose = o; o = sp_align128( o + (uint64_t)cap.dispose * sizeof(uint32_t));
	L.heap = o; o = sp_align128( o + (uint64_t)cap.heap * sizeof(SpExpiry));
	L.stoplog = o; o = sp_align128( o + (uint64_t)(nstopwords + 1) * sizeof(SpStopLog));
	L.free_rules = o; o = sp_align128( o + (uint64_t)cap.rules * sizeof(uint32_t));
	L.free_triggers = o; o = sp_align128( o + (uint64_t)cap.triggers * sizeof(uint32_t));
	L.free_items = o; o = sp_align128( o + (uint64_t)cap.items * sizeof(uint32_t));
	L.free_datarefs = o; o = sp_align128( o + (uint64_t)cap.datarefs * sizeof(uint32_t));
	L.win_rules = o; o = sp_align128( o + (uint64_t)cap.rules * SP_DISPOSE_WINDOW * sizeof(uint32_t));
	L.scratch = o; o = sp_align128( o + (uint64_t)cap.triggers * sizeof(uint32_t));
	L.bytes = o;
}

SP_MHOSTDEV void sp_arena_bind( const SpArenaLayout& L, uint8_t* base, SpArena& A)
{
	A.cap = L.cap;
	A.rules = (SpRule*)(base + L.rules); A.triggers = (SpTrigger*)(base + L.triggers);
	A.bucket_event = (uint32_t*)(base + L.bucket_event); A.bucket_trigger = (uint32_t*)(base + L.bucket_trigger);
	A.items = (SpItem*)(base + L.items); A.datarefs = (SpDataRef*)(base + L.datarefs);
	A.results = (SpPendingResult*)(base + L.results); A.follow = (SpFollow*)(base + L.follow);
	A.dispose = (uint32_t*)(base + L.dispose); A.heap = (SpExpiry*)(base + L.heap);
	A.stoplog = (SpStopLog*)(base + L.stoplog);
	A.free_rules = (uint32_t*)(base + L.free_rules); A.free_triggers = (uint32_t*)(base + L.free_triggers);
	A.free_items = (uint32_t*)(base + L.free_items); A.free_datarefs = (uint32_t*)(base + L.free_datarefs);
	A.win_rules = (uint32_t*)(base + L.win_rules); A.scratch = (uint32_t*)(base + L.scratch);
}

// ---- one document, all steps in sequence on one thread: the definition of the result the warp kernel must
// reproduce (tests/hostsim runs this on the CPU). Returns the number of tokens consumed without error.
SP_MHD uint64_t sp_match_document_seq( const SpMatcherTables& T, SpArena& A, SpDocState& S, const sp_lexem* tokens, const uint32_t* origsegs, uint64_t tb, uint64_t te)
{
	uint64_t nevents = 0;
	sp_doc_begin( S);
	for (uint64_t ti=tb; ti<te; ++ti)
	{
		sp_input_begin( A, S, tokens[ ti], origsegs ? origsegs[ ti] : 0);
		for (uint32_t fi=0; fi<S.nfollow && !S.status; ++fi)
		{
			uint32_t ev = A.follow[ fi].event;
			if (ev)
			{
				SpEventData data = A.follow[ fi].data;
				uint32_t b = sp_evhash( ev), n = S.bucket_size[ b];
				const uint32_t* evs = A.bucket_event + (size_t)b * A.cap.triggers;
				for (uint32_t i=0; i<n && !S.status; ++i) if (evs[ i] == ev) sp_fire_trigger_at( A, S, b, i, data);
			}
			if (!S.status) sp_event_finish( T, A, S, fi);
		}
		if (S.status) break;
		++nevents;
	}
	return nevents;
}

#endif
