// Compiled rule graph of the level-2 matcher as it lives in HBM: flat, index addressed, read-only tables.
// Built by matcher_program.hpp from the pushTerm/pushExpression/definePattern calls; this is what replaces
// the ProgramTable arenas of the reference (src/ruleMatcherAutomaton.hpp:344-413: PodStackPool trigger-def
// lists, program-trigger lists and the unordered_map event -> program list).
#ifndef SP_MATCHER_TABLES_H_INCLUDED
#define SP_MATCHER_TABLES_H_INCLUDED
#include <stdint.h>

/* Trigger::SigType, src/ruleMatcherAutomaton.hpp:48 */
#define SP_SIG_ANY          0
#define SP_SIG_SEQUENCE     1
#define SP_SIG_SEQUENCE_IMM 2
#define SP_SIG_WITHIN       3
#define SP_SIG_DEL          4
#define SP_SIG_AND          5

/* Program (src/ruleMatcherAutomaton.hpp:319) with its ActionSlotDef (:291) inlined. 32 bytes. */
typedef struct alignas(16) SpProgram
{
	uint32_t initsigval;
	uint32_t initcount;
	uint32_t event;			/* follow event emitted on a match (0 = none) */
	uint32_t resultHandle;		/* pattern handle of a visible result (0 = none) */
	uint32_t formatHandle;
	uint32_t range;			/* positionRange */
	uint32_t trig_begin;		/* trigger definitions [trig_begin, trig_begin+trig_count) */
	uint32_t trig_count;
} SpProgram;

/* TriggerDef (src/ruleMatcherAutomaton.hpp:305). Stored per program in the order the reference's LIFO list
 * is iterated: last argument first (src/ruleMatcherAutomaton.cpp:276, podStackPoolBase.hpp:48-62). 16 bytes. */
typedef struct alignas(16) SpTriggerDef
{
	uint32_t event;
	uint32_t sigval;
	uint32_t variable;
	uint8_t sigtype;
	uint8_t is_key;
	uint16_t reserved;
} SpTriggerDef;

/* ProgramTrigger (src/ruleMatcherAutomaton.hpp:331), stored per key event in list iteration order. */
typedef struct alignas(8) SpProgramTrigger
{
	uint32_t program;		/* 0-based index into programs */
	uint32_t past_event;		/* original key event of a program moved to an alternative key, else 0 */
} SpProgramTrigger;

/* Open addressing hash table over every event that is a key event or a stop word. */
typedef struct alignas(16) SpEventSlot
{
	uint32_t key;			/* event id + 1; 0 = empty slot */
	uint32_t prog_begin;		/* program triggers [prog_begin, prog_begin+prog_count) */
	uint32_t prog_count;
	uint32_t stopword;		/* 1-based index into the per document stop word log, 0 = not a stop word */
} SpEventSlot;

/* ---- tables of the window evaluator (matcher_eval.cuh) ----
 * Follow events that are known at table build time: a program whose trigger definitions all have the signal type
 * "any" emits its event when it is installed (fireSignal, src/ruleMatcherAutomaton.cpp:794-806: the first signal on a
 * fresh slot with count > 0 is a match), so a term event t is always followed by the events of the "any" programs keyed
 * by t, in the order of the program list of t (installEventPrograms, :1137-1147). One record per such event. 32 bytes. */
typedef struct alignas(16) SpEvalFollow
{
	uint32_t event;			/* the follow event E */
	uint32_t qpt;			/* index into program_triggers: the entry of the emitting program Q in the list of t */
	uint32_t eslot;			/* index of the event slot of E (programs keyed by E), 0xFFFFFFFF = none */
	uint32_t format;		/* format handle of Q (EventData::formathandle of E) */
	uint32_t sub;			/* 1: E refers to the item list of its rule (the first key definition of Q for t has a variable) */
	uint32_t reserved[3];
} SpEvalFollow;

/* per event slot (parallel to event_slots): the static follow events of a term event */
typedef struct alignas(16) SpEvalSlot
{
	uint32_t follow_begin;
	uint32_t follow_count;
	uint32_t xbegin;		/* work items of a token with this event (eval_xlist range): the programs keyed by the event, */
	uint32_t xcount;		/* then the programs keyed by its follow events, in the order of the follow events */
} SpEvalSlot;

/* One work item of a term event: entry of a program list and the site it is installed at. */
typedef struct alignas(8) SpEvalItem
{
	uint32_t ptidx;			/* index into program_triggers / eval_entries / eval_vars */
	uint32_t cpos;			/* 0 = installed by the token event, c = by follow event number c-1 of the token */
} SpEvalItem;

/* Per program list entry (parallel to program_triggers): the rule instance this entry creates, compiled for the window
 * evaluator (matcher_eval.cuh: sp_eval_entry_step). The signals of the installation (key definitions fired on the fresh
 * slot, src/ruleMatcherAutomaton.cpp:1259-1269) do not depend on the document, so the slot they leave behind is stored
 * here (value0, count0, flags), next to the live triggers: event, signal type and signal value of up to 4 of them. A
 * rule is then followed through its window without reading the program or its trigger definitions. 32 bytes. */
#define SP_EVAL_NEV 4
typedef struct alignas(16) SpEvalEntry
{
	uint32_t ev[ SP_EVAL_NEV];	/* events of the live triggers, 0xFFFFFFFF = unused */
	uint32_t codes;			/* 8 bits per live trigger: signal type << 5 | (sequence: signal value - 1, within: bit number) */
	uint32_t value0;		/* slot value behind the installation */
	uint32_t meta;			/* range | ntrig << 16 | nkey << 22 | SP_EVAL_META_* */
	uint32_t meta2;			/* count0 | nitems0 << 16 | varmask << 22 | SP_EVAL_M2_* */
} SpEvalEntry;
#define SP_EVAL_META_RANGE( m)  ((m) & 0xFFFFu)
#define SP_EVAL_META_NTRIG( m)  (((m) >> 16) & 0x3Fu)
#define SP_EVAL_META_NKEY( m)   (((m) >> 22) & 0x3Fu)
#define SP_EVAL_META_GENERIC    (1u << 28)	/* not compiled: evaluated by the full fold (alternative key, more than 4 live triggers, ...) */
#define SP_EVAL_META_DEAD       (1u << 29)	/* finished by the signals of its installation */
#define SP_EVAL_META_ITEMSFULL  (1u << 30)	/* the key event is a follow event: the items of a result are counted by the full fold */
#define SP_EVAL_M2_COUNT( m)    ((m) & 0xFFFFu)
#define SP_EVAL_M2_NITEMS( m)   (((m) >> 16) & 0x3Fu)
#define SP_EVAL_M2_VARMASK( m)  (((m) >> 22) & 0xFu)	/* live triggers that have a variable */
#define SP_EVAL_M2_HASRES       (1u << 26)	/* the program has a result handle */
#define SP_EVAL_M2_DONE0        (1u << 27)	/* the rule has matched at its installation */
#define SP_EVAL_M2_END0         (1u << 28)	/* end_ordpos = position of the key event + 1 (else 0) */
#define SP_EVAL_M2_DATAREF0     (1u << 29)	/* the rule has an item list behind the installation */
#define SP_EVAL_M2_RESULT0      (1u << 30)	/* result emitted at the installation */
#define SP_EVAL_M2_DREFEMIT0    (1u << 31)	/* ... and the rule had its item list then */

/* Variables of a compiled list entry (parallel to eval_entries; read by the write pass only): of the key definitions that
 * take the key event with a variable (in firing order) and of the live triggers (same order as SpEvalEntry::ev). 32 bytes. */
#define SP_EVAL_NKEYVAR 3
typedef struct alignas(16) SpEvalVars
{
	uint32_t keyvar[ SP_EVAL_NKEYVAR];
	uint32_t nkeyvar;		/* number of key takes with a variable; > SP_EVAL_NKEYVAR: not tabulated (full fold) */
	uint32_t livevar[ SP_EVAL_NEV];
} SpEvalVars;

/* Pair index of the hash join (k_eval_count_join). More than nine rule instances out of ten never get a signal behind
 * their installation: none of the events of their live triggers occurs in their window. What such a "quiet" instance
 * adds to the counters depends on its list entry and on the number of tokens in its window only, so the quiet outcome of
 * ALL aggregated entries of an event slot is summed up at compile time (SpEvalAgg), and the count pass looks for the
 * entries that are NOT quiet from the other side: for every distinct token id X in the window of a key token with event
 * slot S, the pair (S, X) names the aggregated entries of S that have a live trigger on X (SpEvalPair -> eval_pairlist,
 * positions inside the work list of S). Those alone are loaded, probed and followed; their outcome replaces their quiet
 * share. Entries that are not aggregated (generic, a result at the installation, a range beyond the join window) are
 * listed per slot (ind_begin/ind_count -> eval_indlist) and handled one by one. */
#define SP_EVAL_AGG_RANGES 16
typedef struct alignas(16) SpEvalAgg
{
	uint32_t n;			/* aggregated entries of the slot: programs installed */
	uint32_t signals;		/* their signals at the installation */
	uint32_t ind_begin, ind_count;	/* entries handled one by one (eval_indlist range: positions inside the work list) */
	uint32_t open[ SP_EVAL_AGG_RANGES];	/* [r]: live triggers of the aggregated entries with range r that outlive their installation */
} SpEvalAgg;

typedef struct alignas(16) SpEvalPair
{
	uint32_t slot;			/* event slot of the key token, 0xFFFFFFFF = empty */
	uint32_t event;			/* token id in the window */
	uint32_t begin, count;		/* eval_pairlist range */
} SpEvalPair;

#define SP_PROGFLAG_LISTENS_FOLLOW 1u	/* some trigger definition waits for an event that is not a term event */

typedef struct SpMatcherTables
{
	const SpProgram* programs;
	const SpTriggerDef* trigger_defs;
	const SpProgramTrigger* program_triggers;
	const SpEventSlot* event_slots;
	uint32_t event_mask;		/* number of slots - 1 (power of two) */
	uint32_t nprograms;
	uint32_t ntrigger_defs;
	uint32_t nprogram_triggers;
	uint32_t nstopwords;
	uint32_t exclusive;		/* option "exclusive" (src/patternMatcher.cpp:632) */
	uint32_t max_result_size;	/* option "maxResultSize" (:628) */
	uint32_t max_trigger_defs;	/* largest trig_count */
	/* window evaluator */
	const SpEvalSlot* eval_slots;	/* [event_mask+1] */
	const SpEvalFollow* eval_follows;
	const SpEvalItem* eval_xlist;
	const uint32_t* prog_flags;	/* [nprograms] SP_PROGFLAG_* */
	const SpEvalEntry* eval_entries;	/* [nprogram_triggers] */
	const SpEvalVars* eval_vars;	/* [nprogram_triggers] */
	const SpEvalAgg* eval_agg;	/* [event_mask+1] */
	const uint32_t* eval_indlist;
	const SpEvalPair* eval_pairs;	/* [eval_pair_mask+1] open addressing, sp_eval_pair_hash */
	const uint32_t* eval_pairlist;
	uint32_t eval_pair_mask;
	uint32_t nfollows;
	uint32_t eval_join_range;	/* largest position range of a compiled list entry, at most 15: window of the hash join (k_eval_count_join) */
	uint32_t eval_capable;		/* 1: the rule set is inside what the window evaluator reproduces exactly (matcher_program.hpp: evalCapable) */
} SpMatcherTables;

#ifdef __CUDACC__
#define SP_TAB_HD __host__ __device__ __forceinline__
#else
#define SP_TAB_HD inline
#endif
SP_TAB_HD uint32_t sp_eval_pair_hash( uint32_t slot, uint32_t event)
{
	uint32_t h = (slot * 0x85EBCA6Bu) ^ (event * 0x9E3779B1u);
	return h ^ (h >> 13);
}
SP_TAB_HD uint32_t sp_event_slot_hash( uint32_t event)
{
	uint32_t h = event * 0x9E3779B1u;
	return h ^ (h >> 15);
}

#endif
