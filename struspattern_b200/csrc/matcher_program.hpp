// Host side of the level-2 matcher: the expression stack of PatternMatcherInstance and the rule graph it
// builds, the stop-word key-event optimisation, and the flattening of the graph into the device tables of
// matcher_tables.h. Pure host C++ (no CUDA), shared by the C ABI (matcher_capi.cu) and the host simulation tests.
//
// Reference behaviour restated here (file:line under the strusPattern tree):
//   eventHandle                                     src/patternMatcher.cpp:100-105
//   pushTerm / pushExpression / pushPattern         src/patternMatcher.cpp:366-520
//   attachVariable / definePattern / defineOption   src/patternMatcher.cpp:522-584, 614-644
//   createProgram / createTrigger / doneProgram     src/ruleMatcherAutomaton.cpp:268-339
//   calcEventWeight / getAltEventId / optimize      src/ruleMatcherAutomaton.cpp:341-442, 478-586
// The reference keeps trigger definitions and program lists in LIFO linked lists; here they are plain vectors
// in definition order that are read backwards where the reference iterates its lists.
#ifndef SP_MATCHER_PROGRAM_HPP_INCLUDED
#define SP_MATCHER_PROGRAM_HPP_INCLUDED
#include "matcher_tables.h"
#include "strus_pattern_b200.h"
#include "result_format.hpp"
#include <vector>
#include <string>
#include <deque>
#include <map>
#include <set>
#include <unordered_map>
#include <stdexcept>
#include <limits>
#include <algorithm>
#include <cstring>
#include <strings.h>

namespace sp {

enum EventType {TermEvent=0, ExpressionEvent=1, ReferenceEvent=2};

inline uint32_t eventHandle( EventType type, uint32_t idx)
{
	if (idx >= (1u<<29)) throw std::runtime_error( "event handle out of range");
	return idx | ((uint32_t)type << 29);
}

// string <-> dense 1-based handle in insertion order (strusBase SymbolTable as used by src/patternMatcher.cpp:45-65)
class NameTable
{
public:
	uint32_t getOrCreate( const std::string& name)
	{
		std::unordered_map<std::string,uint32_t>::const_iterator it = m_index.find( name);
		if (it != m_index.end()) return it->second;
		m_names.push_back( name);
		return m_index[ name] = (uint32_t)m_names.size();
	}
	uint32_t get( const std::string& name) const
	{
		std::unordered_map<std::string,uint32_t>::const_iterator it = m_index.find( name);
		return it == m_index.end() ? 0 : it->second;
	}
	const char* key( uint32_t handle) const {return (handle == 0 || handle > m_names.size()) ? 0 : m_names[ handle-1].c_str();}
	std::size_t size() const {return m_names.size();}
private:
	std::unordered_map<std::string,uint32_t> m_index;
	std::deque<std::string> m_names;
};

struct MatcherTablesHost
{
	std::vector<SpProgram> programs;
	std::vector<SpTriggerDef> trigger_defs;
	std::vector<SpProgramTrigger> program_triggers;
	std::vector<SpEventSlot> event_slots;
	std::vector<uint32_t> stopwords;	// event id per stop word index (0-based)
	std::vector<SpEvalSlot> eval_slots;	// window evaluator (matcher_eval.cuh): static follow events per event slot
	std::vector<SpEvalFollow> eval_follows;
	std::vector<SpEvalItem> eval_xlist;
	std::vector<uint32_t> prog_flags;
	std::vector<SpEvalEntry> eval_entries;
	std::vector<SpEvalVars> eval_vars;
	std::vector<SpEvalAgg> eval_agg;	// hash join: quiet outcome per slot, entries handled one by one, pair index
	std::vector<uint32_t> eval_indlist;
	std::vector<SpEvalPair> eval_pairs;
	std::vector<uint32_t> eval_pairlist;
	SpMatcherTables view;			// pointers are filled by the owner (host or device addresses)
};

class MatcherProgram
{
public:
	MatcherProgram()
		:m_expressionEventCnt(0),m_totalNofPrograms(0),m_stopwordOccurrenceFactor(0.01f),m_weightFactor(10.0f),m_maxRange(5)
		,m_maxResultSize(100),m_exclusive(false),m_flattened(false){}

	// ---- PatternMatcherInstanceInterface
	void defineOption( const std::string& name, double value)
	{
		const char* nam = name.c_str();
		if (0==strcasecmp( nam, "stopwordOccurrenceFactor")) m_stopwordOccurrenceFactor = (float)value;
		else if (0==strcasecmp( nam, "weightFactor")) m_weightFactor = (float)value;
		else if (0==strcasecmp( nam, "maxRange")) m_maxRange = (uint32_t)(value + std::numeric_limits<double>::epsilon());
		else if (0==strcasecmp( nam, "maxResultSize")) m_maxResultSize = (uint32_t)(value + std::numeric_limits<double>::epsilon());
		else if (0==strcasecmp( nam, "exclusive")) m_exclusive = true;
		else throw std::runtime_error( "unknown token pattern match option: '" + name + "'");
		m_flattened = false;
	}
	void defineTermFrequency( uint32_t termid, double df)
	{
		if (df <= std::numeric_limits<double>::epsilon()) throw std::runtime_error( "illegal value for df (must be positive)");
		m_frequency[ eventHandle( TermEvent, termid)] = df;
	}
	void pushTerm( uint32_t termid)
	{
		m_stack.push_back( Node( eventHandle( TermEvent, termid)));
	}
	void pushPattern( const std::string& name)
	{
		m_stack.push_back( Node( eventHandle( ReferenceEvent, m_patterns.getOrCreate( name))));
	}
	void attachVariable( const std::string& name)
	{
		if (m_stack.empty()) throw std::runtime_error( "illegal operation attach variable when no node on the stack");
		if (m_stack.back().variable) throw std::runtime_error( "more than one variable assignment to a node");
		m_stack.back().variable = m_variables.getOrCreate( name);
	}
	void pushExpression( int joinop, std::size_t argc, uint32_t range, uint32_t cardinality)
	{
		if (argc > m_stack.size() || argc > (std::size_t)std::numeric_limits<uint32_t>::max())
		{
			throw std::runtime_error( "expression references more arguments than nodes on the stack");
		}
		Rule rule;
		rule.range = range;
		rule.initsigval = 0;
		rule.initcount = cardinality ? cardinality : (uint32_t)argc;
		rule.event = eventHandle( ExpressionEvent, ++m_expressionEventCnt);
		rule.resultHandle = rule.formatHandle = 0;
		uint8_t slotSig = SP_SIG_ANY;
		switch (joinop)
		{
			case SP_OP_SEQUENCE: slotSig = SP_SIG_SEQUENCE; rule.initsigval = (uint32_t)argc; break;
			case SP_OP_SEQUENCE_IMM: slotSig = SP_SIG_SEQUENCE_IMM; rule.initsigval = (uint32_t)argc; break;
			case SP_OP_SEQUENCE_STRUCT: slotSig = SP_SIG_SEQUENCE; rule.initsigval = (uint32_t)argc - 1; --rule.initcount; break;
			case SP_OP_WITHIN:
			case SP_OP_WITHIN_STRUCT:
				slotSig = SP_SIG_WITHIN;
				if (argc > 32) throw std::runtime_error( joinop == SP_OP_WITHIN ? "operator 'within': number of arguments out of range (32)" : "operator 'within_struct': number of arguments out of range (32)");
				rule.initsigval = 0xFFFFFFFFu;
				if (joinop == SP_OP_WITHIN_STRUCT) --rule.initcount;
				break;
			case SP_OP_ANY: slotSig = SP_SIG_ANY; rule.initcount = cardinality ? cardinality : 1u; break;
			case SP_OP_AND: slotSig = SP_SIG_AND; break;
			default: throw std::runtime_error( "unknown join operation");
		}
		for (std::size_t ai=0; ai<argc; ++ai)
		{
			const Node& arg = m_stack[ m_stack.size() - argc + ai];
			SpTriggerDef def; std::memset( &def, 0, sizeof(def));
			def.event = arg.event; def.variable = arg.variable; def.sigtype = slotSig;
			switch (joinop)
			{
				case SP_OP_SEQUENCE_STRUCT:
					if (ai == 0) def.sigtype = SP_SIG_DEL;		// structure delimiter
					else {def.sigval = (uint32_t)(argc - ai); def.is_key = (ai == 1);}
					break;
				case SP_OP_WITHIN_STRUCT:
					if (ai == 0) def.sigtype = SP_SIG_DEL;
					else {def.sigval = 1u << (argc - ai); def.is_key = 1;}
					break;
				case SP_OP_SEQUENCE:
					def.sigval = (uint32_t)(argc - ai); def.is_key = (ai == 0);
					break;
				case SP_OP_SEQUENCE_IMM:
					if (ai == 0) def.sigtype = SP_SIG_SEQUENCE;	// the first element has no predecessor
					def.sigval = (uint32_t)(argc - ai); def.is_key = (ai == 0);
					break;
				case SP_OP_WITHIN:
					def.sigval = 1u << (argc - ai - 1); def.is_key = 1;
					break;
				default:
					def.is_key = 1;
					break;
			}
			rule.defs.push_back( def);
			m_eventOccurrence[ def.event] += 1;
		}
		uint32_t program = addRule( rule);
		m_stack.erase( m_stack.end() - argc, m_stack.end());
		m_stack.push_back( Node( rule.event, program));
		m_flattened = false;
	}
	// returns the format handle assigned (0 = no format string)
	uint32_t definePattern( const std::string& name, const std::string& formatstring, bool visible)
	{
		if (m_stack.empty()) throw std::runtime_error( "illegal operation close pattern when no node on the stack");
		Node& top = m_stack.back();
		uint32_t resultHandle = m_patterns.getOrCreate( name);
		uint32_t resultEvent = eventHandle( ReferenceEvent, resultHandle);
		uint32_t formatHandle = 0;
		if (!formatstring.empty())
		{
			// createResultFormat (src/patternMatcher.cpp:562-565): parsed now, its variables must be known now
			if (m_formats.size() >= 0xFFFFu) throw std::runtime_error( "too many result format strings (65535 at most)");
			ResultFormat parsed = parseResultFormat( formatstring, [this]( const std::string& vn) -> uint32_t {return m_variables.get( vn);});
			m_formats.push_back( formatstring);
			m_parsedFormats.push_back( parsed);
			formatHandle = (uint32_t)m_formats.size();
		}
		uint32_t program = top.program;
		if (!program)
		{
			// a bare term or reference is enveloped into a program of its own
			Rule rule;
			rule.range = 0; rule.initsigval = 0; rule.initcount = 1;
			rule.event = resultEvent; rule.resultHandle = resultHandle; rule.formatHandle = formatHandle;
			SpTriggerDef def; std::memset( &def, 0, sizeof(def));
			def.event = top.event; def.variable = top.variable; def.sigtype = SP_SIG_ANY; def.is_key = 1;
			rule.defs.push_back( def);
			m_eventOccurrence[ def.event] += 1;
			program = addRule( rule);
		}
		else if (top.variable)
		{
			throw std::runtime_error( "variable assignments only allowed to subexpressions of pattern");
		}
		Rule& rule = m_rules[ program-1];
		rule.event = resultEvent;
		rule.resultHandle = visible ? resultHandle : 0;
		rule.formatHandle = formatHandle;
		m_flattened = false;
		return formatHandle;
	}
	// Pair index and quiet aggregates of the hash join (matcher_tables.h: SpEvalAgg, SpEvalPair)
	static void buildJoinTables( MatcherTablesHost& T, uint32_t joinrange)
	{
		const uint32_t nslots = (uint32_t)T.event_slots.size();
		SpEvalAgg noagg; std::memset( &noagg, 0, sizeof(noagg));
		T.eval_agg.assign( nslots, noagg);
		T.eval_indlist.clear(); T.eval_pairlist.clear();
		std::map<std::pair<uint32_t,uint32_t>, std::vector<uint32_t> > pairs;
		for (uint32_t h=0; h<nslots; ++h)
		{
			const SpEvalSlot& es = T.eval_slots[ h];
			SpEvalAgg& agg = T.eval_agg[ h];
			agg.ind_begin = (uint32_t)T.eval_indlist.size();
			for (uint32_t rem=0; rem<es.xcount; ++rem)
			{
				const SpEvalEntry& en = T.eval_entries[ T.eval_xlist[ es.xbegin + rem].ptidx];
				const uint32_t range = SP_EVAL_META_RANGE( en.meta);
				const bool dead = (en.meta & SP_EVAL_META_DEAD) != 0;
				if ((en.meta & SP_EVAL_META_GENERIC) || (en.meta2 & SP_EVAL_M2_RESULT0) || (!dead && (range > joinrange || range >= SP_EVAL_AGG_RANGES)))
				{
					T.eval_indlist.push_back( rem);
					continue;
				}
				agg.n += 1;
				agg.signals += SP_EVAL_META_NKEY( en.meta);
				if (dead) continue;
				agg.open[ range] += SP_EVAL_META_NTRIG( en.meta);
				std::set<uint32_t> events;
				for (int e=0; e<SP_EVAL_NEV; ++e) if (en.ev[ e] != 0xFFFFFFFFu) events.insert( en.ev[ e]);
				for (uint32_t ev: events) pairs[ std::make_pair( h, ev)].push_back( rem);
			}
			agg.ind_count = (uint32_t)T.eval_indlist.size() - agg.ind_begin;
		}
		uint32_t size = 16;
		while (size < 2 * pairs.size() + 1) size <<= 1;
		SpEvalPair nopair; nopair.slot = 0xFFFFFFFFu; nopair.event = 0; nopair.begin = nopair.count = 0;
		T.eval_pairs.assign( size, nopair);
		for (std::map<std::pair<uint32_t,uint32_t>, std::vector<uint32_t> >::const_iterator pi = pairs.begin(); pi != pairs.end(); ++pi)
		{
			SpEvalPair p; p.slot = pi->first.first; p.event = pi->first.second;
			p.begin = (uint32_t)T.eval_pairlist.size(); p.count = (uint32_t)pi->second.size();
			T.eval_pairlist.insert( T.eval_pairlist.end(), pi->second.begin(), pi->second.end());
			uint32_t i = sp_eval_pair_hash( p.slot, p.event) & (size - 1);
			while (T.eval_pairs[ i].slot != 0xFFFFFFFFu) i = (i + 1) & (size - 1);
			T.eval_pairs[ i] = p;
		}
		if (T.eval_indlist.empty()) T.eval_indlist.push_back( 0);
		if (T.eval_pairlist.empty()) T.eval_pairlist.push_back( 0);
		T.view.eval_agg = T.eval_agg.data(); T.view.eval_indlist = T.eval_indlist.data();
		T.view.eval_pairs = T.eval_pairs.data(); T.view.eval_pairlist = T.eval_pairlist.data();
		T.view.eval_pair_mask = size - 1;
	}
	// ProgramTable::optimize
	void compile()
	{
		eliminateUnusedEvents();
		std::vector<uint32_t> eventsToMove;
		for (EventListMap::const_iterator ei = m_eventPrograms.begin(); ei != m_eventPrograms.end(); ++ei)
		{
			std::map<uint32_t,uint32_t>::const_iterator ki = m_keyOccurrence.find( ei->first);
			if (ki != m_keyOccurrence.end() && ki->second >= (float)m_totalNofPrograms * m_stopwordOccurrenceFactor)
			{
				eventsToMove.push_back( ei->first);
			}
		}
		for (uint32_t eventid: eventsToMove)
		{
			EventListMap::iterator ei = m_eventPrograms.find( eventid);
			if (ei == m_eventPrograms.end()) continue;
			const uint32_t oldlist = ei->second;
			std::vector<SpProgramTrigger> kept;
			double weight = eventWeight( eventid);
			// the list is visited last-pushed first; the kept entries are pushed onto a fresh list in that order
			for (std::size_t li = m_lists[ oldlist].size(); li > 0; --li)
			{
				SpProgramTrigger pt = m_lists[ oldlist][ li-1];
				const Rule& rule = m_rules[ pt.program-1];
				uint32_t alt = alternativeKeyEvent( eventid, rule);
				bool moved = false;
				if (alt)
				{
					double altWeight = eventWeight( alt) * m_weightFactor;
					if (!pt.past_event && m_maxRange >= rule.range && weight > altWeight)
					{
						addEventProgram( alt, pt.program, eventid);
						for (const SpTriggerDef& def: rule.defs) if (def.sigtype == SP_SIG_DEL) m_stopwords.insert( def.event);
						moved = true;
					}
				}
				if (!moved) kept.push_back( pt);
			}
			ei = m_eventPrograms.find( eventid);	// addEventProgram may have rehashed the map
			if (!kept.empty()) m_lists[ ei->second].swap( kept);
			else {m_lists[ ei->second].clear(); m_eventPrograms.erase( ei);}
		}
		m_flattened = false;
	}

	const char* patternName( uint32_t handle) const {return m_patterns.key( handle);}
	const char* variableName( uint32_t handle) const {return m_variables.key( handle);}
	const char* formatString( uint32_t handle) const {return (handle == 0 || handle > m_formats.size()) ? 0 : m_formats[ handle-1].c_str();}
	// PatternMatcherResult::value (src/patternMatcher.cpp:253-262); false: the result has no format string (value NULL)
	bool resultValue( const sp_result& r, const sp_item* docitems, std::string& out) const
	{
		if (!r.format) return false;
		mapResultFormat( m_parsedFormats, r.format, docitems + r.item_begin, r.item_count, out);
		return true;
	}
	// PatternMatcherResultItem::value (:172-181); the nested records of the item follow it
	bool itemValue( const sp_item* item, std::string& out) const
	{
		if (!SP_ITEM_FORMAT( *item)) return false;
		mapResultFormat( m_parsedFormats, SP_ITEM_FORMAT( *item), item + 1, SP_ITEM_NESTED( *item), out);
		return true;
	}
	bool exclusive() const {return m_exclusive;}
	uint32_t maxResultSize() const {return m_maxResultSize;}
	std::size_t nofPrograms() const {return m_rules.size();}
	std::size_t nofStopwords() const {return m_stopwords.size();}

	// Flat tables (pointers inside `view` address the vectors of the returned object).
	const MatcherTablesHost& tables()
	{
		if (m_flattened) return m_tables;
		MatcherTablesHost& T = m_tables;
		T = MatcherTablesHost();
		uint32_t maxdefs = 0;
		for (const Rule& rule: m_rules)
		{
			SpProgram p;
			p.initsigval = rule.initsigval; p.initcount = rule.initcount; p.event = rule.event;
			p.resultHandle = rule.resultHandle; p.formatHandle = rule.formatHandle; p.range = rule.range;
			p.trig_begin = (uint32_t)T.trigger_defs.size(); p.trig_count = (uint32_t)rule.defs.size();
			for (std::size_t di = rule.defs.size(); di > 0; --di) T.trigger_defs.push_back( rule.defs[ di-1]);
			maxdefs = std::max( maxdefs, p.trig_count);
			T.programs.push_back( p);
		}
		std::map<uint32_t,uint32_t> stopidx;
		for (uint32_t ev: m_stopwords) {T.stopwords.push_back( ev); stopidx[ ev] = (uint32_t)T.stopwords.size();}
		std::set<uint32_t> keys( m_stopwords.begin(), m_stopwords.end());
		for (EventListMap::const_iterator ei = m_eventPrograms.begin(); ei != m_eventPrograms.end(); ++ei) keys.insert( ei->first);
		uint32_t nslots = 8;
		while (nslots < 2 * keys.size() + 2) nslots <<= 1;
		SpEventSlot empty; std::memset( &empty, 0, sizeof(empty));
		T.event_slots.assign( nslots, empty);
		for (uint32_t ev: keys)
		{
			SpEventSlot slot; slot.key = ev + 1; slot.prog_begin = (uint32_t)T.program_triggers.size(); slot.prog_count = 0; slot.stopword = 0;
			EventListMap::const_iterator ei = m_eventPrograms.find( ev);
			if (ei != m_eventPrograms.end())
			{
				const std::vector<SpProgramTrigger>& list = m_lists[ ei->second];
				for (std::size_t li = list.size(); li > 0; --li)
				{
					SpProgramTrigger pt = list[ li-1];
					pt.program -= 1;
					T.program_triggers.push_back( pt);
				}
				slot.prog_count = (uint32_t)list.size();
			}
			std::map<uint32_t,uint32_t>::const_iterator si = stopidx.find( ev);
			if (si != stopidx.end()) slot.stopword = si->second;
			uint32_t h = sp_event_slot_hash( ev) & (nslots-1);
			while (T.event_slots[ h].key) h = (h + 1) & (nslots-1);
			T.event_slots[ h] = slot;
		}
		std::memset( &T.view, 0, sizeof(T.view));
		T.view.programs = T.programs.data(); T.view.trigger_defs = T.trigger_defs.data();
		T.view.program_triggers = T.program_triggers.data(); T.view.event_slots = T.event_slots.data();
		T.view.event_mask = nslots - 1;
		T.view.nprograms = (uint32_t)T.programs.size(); T.view.ntrigger_defs = (uint32_t)T.trigger_defs.size();
		T.view.nprogram_triggers = (uint32_t)T.program_triggers.size();
		T.view.nstopwords = (uint32_t)T.stopwords.size();
		T.view.exclusive = m_exclusive ? 1 : 0; T.view.max_result_size = m_maxResultSize;
		T.view.max_trigger_defs = maxdefs;
		buildEvalTables( T);
		m_flattened = true;
		return T;
	}

private:
	// Tables of the window evaluator (matcher_eval.cuh) and the decision whether it reproduces this rule set exactly.
	// It does when every event that is not a term event and that some program waits for (as key or as trigger) can only
	// come from programs that emit at their installation: all trigger definitions of type "any" on term events. Other
	// nested expressions, "and" programs (src/ruleMatcherAutomaton.cpp:807-826), the option "exclusive" (the covered
	// flags of src/patternMatcher.cpp:192-246 depend on the emission order) and programs with more than 32 trigger
	// definitions are left to the state machine.
	void buildEvalTables( MatcherTablesHost& T) const
	{
		const uint32_t nslots = (uint32_t)T.event_slots.size();
		SpEvalSlot noslot; noslot.follow_begin = noslot.follow_count = noslot.xbegin = noslot.xcount = 0;
		T.eval_slots.assign( nslots, noslot);
		T.prog_flags.assign( T.programs.size(), 0);
		bool capable = !m_exclusive;
		std::vector<char> anyEmitter( T.programs.size(), 0);
		std::set<uint32_t> listened;
		for (std::size_t pi=0; pi<T.programs.size(); ++pi)
		{
			const SpProgram& P = T.programs[ pi];
			if (P.trig_count > 32) capable = false;
			bool allAny = true, allTerm = true;
			for (uint32_t d=0; d<P.trig_count; ++d)
			{
				const SpTriggerDef& def = T.trigger_defs[ P.trig_begin + d];
				if (def.sigtype == SP_SIG_AND) capable = false;
				if (def.sigtype != SP_SIG_ANY) allAny = false;
				if ((def.event >> 29) != TermEvent) {allTerm = false; listened.insert( def.event); T.prog_flags[ pi] |= SP_PROGFLAG_LISTENS_FOLLOW;}
			}
			anyEmitter[ pi] = (allAny && allTerm && (P.initcount & 0xFFFFu) >= 1 && P.trig_count >= 1);
		}
		for (const SpEventSlot& slot: T.event_slots)
		{
			if (!slot.key || !slot.prog_count) continue;
			const uint32_t ev = slot.key - 1;
			if ((ev >> 29) != TermEvent) listened.insert( ev);
			for (uint32_t k=0; k<slot.prog_count; ++k)
			{
				const SpProgramTrigger& pt = T.program_triggers[ slot.prog_begin + k];
				if (!pt.past_event) continue;
				// alternative key programs: the replay compares the log times of structure delimiters, which the
				// evaluator does for term events
				const SpProgram& P = T.programs[ pt.program];
				for (uint32_t d=0; d<P.trig_count; ++d)
				{
					const SpTriggerDef& def = T.trigger_defs[ P.trig_begin + d];
					if (def.sigtype == SP_SIG_DEL && (def.event >> 29) != TermEvent) capable = false;
				}
				if ((pt.past_event >> 29) != TermEvent) listened.insert( pt.past_event);
			}
		}
		for (std::size_t pi=0; pi<T.programs.size(); ++pi)
		{
			if (T.programs[ pi].event && listened.count( T.programs[ pi].event) && !anyEmitter[ pi]) capable = false;
		}
		if (capable)
		{
			for (uint32_t h=0; h<nslots; ++h)
			{
				const SpEventSlot& slot = T.event_slots[ h];
				if (!slot.key || ((slot.key - 1) >> 29) != TermEvent) continue;
				T.eval_slots[ h].follow_begin = (uint32_t)T.eval_follows.size();
				for (uint32_t k=0; k<slot.prog_count; ++k)
				{
					const SpProgramTrigger& pt = T.program_triggers[ slot.prog_begin + k];
					const SpProgram& P = T.programs[ pt.program];
					if (pt.past_event || !anyEmitter[ pt.program] || !P.event || !listened.count( P.event)) continue;
					SpEvalFollow F; std::memset( &F, 0, sizeof(F));
					F.event = P.event; F.qpt = slot.prog_begin + k; F.eslot = 0xFFFFFFFFu; F.format = P.formatHandle;
					for (uint32_t d=0; d<P.trig_count; ++d)
					{
						const SpTriggerDef& def = T.trigger_defs[ P.trig_begin + d];
						if (def.event == slot.key - 1) {F.sub = def.variable ? 1 : 0; break;}	// the first signal emits
					}
					// an item that refers to F carries F's own items as nested records behind it (the value string is made of
					// them, sp_gather_items): only the state machine writes those
					if (F.format && F.sub) capable = false;
					for (uint32_t e = sp_event_slot_hash( F.event) & (nslots-1); T.event_slots[ e].key; e = (e + 1) & (nslots-1))
					{
						if (T.event_slots[ e].key == F.event + 1) {F.eslot = e; break;}
					}
					T.eval_follows.push_back( F);
				}
				T.eval_slots[ h].follow_count = (uint32_t)T.eval_follows.size() - T.eval_slots[ h].follow_begin;
			}
		}
		// list entries: live trigger events and the outcome of the installation signals
		SpEvalEntry noentry; std::memset( &noentry, 0, sizeof(noentry)); noentry.meta = SP_EVAL_META_GENERIC;
		T.eval_entries.assign( T.program_triggers.size(), noentry);
		SpEvalVars novars; std::memset( &novars, 0, sizeof(novars));
		T.eval_vars.assign( T.program_triggers.size(), novars);
		for (const SpEventSlot& slot: T.event_slots)
		{
			if (!slot.key) continue;
			for (uint32_t k=0; k<slot.prog_count; ++k)
			{
				T.eval_entries[ slot.prog_begin + k] = evalEntry( T, T.program_triggers[ slot.prog_begin + k], slot.key - 1, T.eval_vars[ slot.prog_begin + k]);
			}
		}
		uint32_t joinrange = 0;
		for (const SpEvalEntry& en: T.eval_entries)
		{
			if (!(en.meta & (SP_EVAL_META_GENERIC | SP_EVAL_META_DEAD)) && SP_EVAL_META_RANGE( en.meta) <= 15) joinrange = std::max( joinrange, SP_EVAL_META_RANGE( en.meta));
		}
		T.view.eval_join_range = joinrange;
		T.view.eval_entries = T.eval_entries.data(); T.view.eval_vars = T.eval_vars.data();
		// the work items of a term event, flattened
		for (uint32_t h=0; h<nslots; ++h)
		{
			const SpEventSlot& slot = T.event_slots[ h];
			if (!slot.key || ((slot.key - 1) >> 29) != TermEvent) continue;
			T.eval_slots[ h].xbegin = (uint32_t)T.eval_xlist.size();
			for (uint32_t k=0; k<slot.prog_count; ++k) {SpEvalItem it; it.ptidx = slot.prog_begin + k; it.cpos = 0; T.eval_xlist.push_back( it);}
			for (uint32_t c=0; c<T.eval_slots[ h].follow_count; ++c)
			{
				const SpEvalFollow& F = T.eval_follows[ T.eval_slots[ h].follow_begin + c];
				if (F.eslot == 0xFFFFFFFFu) continue;
				const SpEventSlot& fs = T.event_slots[ F.eslot];
				for (uint32_t k=0; k<fs.prog_count; ++k) {SpEvalItem it; it.ptidx = fs.prog_begin + k; it.cpos = c + 1; T.eval_xlist.push_back( it);}
			}
			T.eval_slots[ h].xcount = (uint32_t)T.eval_xlist.size() - T.eval_slots[ h].xbegin;
		}
		T.view.eval_xlist = T.eval_xlist.data();
		buildJoinTables( T, joinrange);
		T.view.eval_slots = T.eval_slots.data(); T.view.eval_follows = T.eval_follows.data(); T.view.prog_flags = T.prog_flags.data();
		T.view.nfollows = (uint32_t)T.eval_follows.size();
		T.view.eval_capable = capable ? 1 : 0;
	}

	// Installation of program pt.program by `keyevent` on an event at position S (any S >= 1) without sub data: which
	// definitions become live triggers (installProgram, src/ruleMatcherAutomaton.cpp:1204-1250) and what the direct signals
	// of the key definitions (:1259-1269, fireSignal :772-979 on a fresh slot) leave behind. With end_ordpos in {0, S+1}
	// the position tests of fireSignal do not depend on S. `termKey`: the key event is a term event (no sub data).
	static SpEvalEntry evalEntry( const MatcherTablesHost& T, const SpProgramTrigger& pt, uint32_t keyevent, SpEvalVars& vars)
	{
		std::memset( &vars, 0, sizeof(vars));
		SpEvalEntry en; std::memset( &en, 0, sizeof(en));
		for (int e=0; e<SP_EVAL_NEV; ++e) en.ev[ e] = 0xFFFFFFFFu;
		const SpProgram& P = T.programs[ pt.program];
		uint32_t count = P.initcount & 0xFFFFu, value = P.initsigval;
		bool generic = pt.past_event != 0 || P.range >= 0xFFFFu || P.trig_count > 32;
		uint32_t nev = 0, ntrig = 0, nkey = 0, varmask = 0;
		bool hasKeyEvent = false;
		std::vector<uint32_t> keydefs;
		for (uint32_t d=0; d<P.trig_count && d<32; ++d)
		{
			const SpTriggerDef& def = T.trigger_defs[ P.trig_begin + d];
			bool install = true;
			if (def.event == keyevent)
			{
				keydefs.push_back( d);
				const bool needsInstall = (def.sigtype == SP_SIG_ANY && count > 1);
				if (def.is_key && !hasKeyEvent) {hasKeyEvent = true; install = needsInstall;}
				else if (def.sigtype == SP_SIG_DEL) install = needsInstall;
			}
			if (!install) continue;
			++ntrig;
			uint32_t sig = 0;
			switch (def.sigtype)
			{
				case SP_SIG_SEQUENCE: case SP_SIG_SEQUENCE_IMM:
					if (def.sigval < 1 || def.sigval > 32) generic = true; else sig = def.sigval - 1;
					break;
				case SP_SIG_WITHIN:
					if (!def.sigval || (def.sigval & (def.sigval - 1))) generic = true; else sig = (uint32_t)__builtin_ctz( def.sigval);
					break;
				case SP_SIG_ANY: case SP_SIG_DEL: break;
				default: generic = true; break;
			}
			if ((def.event >> 29) != TermEvent || nev >= SP_EVAL_NEV) {generic = true; continue;}
			en.ev[ nev] = def.event;
			en.codes |= (((uint32_t)def.sigtype << 5) | sig) << (8 * nev);
			if (def.variable) varmask |= 1u << nev;
			vars.livevar[ nev] = def.variable;
			++nev;
		}
		bool endSet = false, done = false, disposed = false, result = false, dataref = false, drefEmit = false;
		uint32_t nitems = 0;
		for (uint32_t d: keydefs)
		{
			const SpTriggerDef& def = T.trigger_defs[ P.trig_begin + d];
			bool match = false, finished = false, take = false;
			++nkey;
			switch (def.sigtype)
			{
				case SP_SIG_ANY:
					take = true;
					if (count > 0) {match = true; --count; finished = (count == 0); endSet = true;}
					break;
				case SP_SIG_SEQUENCE: case SP_SIG_SEQUENCE_IMM: case SP_SIG_WITHIN:
				{
					bool accept;
					if (def.sigtype == SP_SIG_WITHIN) accept = (def.sigval & value) != 0 && !endSet;
					else if (def.sigtype == SP_SIG_SEQUENCE) accept = def.sigval == value && !endSet;
					else accept = false;	// end_ordpos == start_ordpos: neither 0 nor S+1 equals S
					if (accept)
					{
						endSet = true;
						if (def.sigtype == SP_SIG_WITHIN) value &= ~def.sigval; else value = def.sigval - 1;
						if (count > 0) {--count; match = (count == 0);} else match = true;
						finished = (value == 0);
						take = true;
					}
					break;
				}
				case SP_SIG_DEL: count = 0; value = 0; disposed = true; continue;
				default: generic = true; break;
			}
			if (take && def.variable)
			{
				dataref = true; ++nitems;
				if (vars.nkeyvar < SP_EVAL_NKEYVAR) vars.keyvar[ vars.nkeyvar] = def.variable;
				++vars.nkeyvar;
			}
			// a follow event taken without a variable joins the items of its rule (:911-918): left to the full fold
			if (take && !def.variable && (keyevent >> 29) != TermEvent) generic = true;
			if (match)
			{
				if (!done) {if (P.resultHandle) {result = true; drefEmit = dataref;} done = true;}
				if (finished) disposed = true;
			}
		}
		if (nkey > 63 || ntrig > 63 || nitems > 63) generic = true;
		en.value0 = value;
		en.meta = (P.range & 0xFFFFu) | ((ntrig & 0x3Fu) << 16) | ((nkey & 0x3Fu) << 22);
		en.meta2 = (count & 0xFFFFu) | ((nitems & 0x3Fu) << 16) | (varmask << 22);
		if (P.resultHandle) en.meta2 |= SP_EVAL_M2_HASRES;
		if (done) en.meta2 |= SP_EVAL_M2_DONE0;
		if (endSet) en.meta2 |= SP_EVAL_M2_END0;
		if (dataref) en.meta2 |= SP_EVAL_M2_DATAREF0;
		if (result) en.meta2 |= SP_EVAL_M2_RESULT0;
		if (drefEmit) en.meta2 |= SP_EVAL_M2_DREFEMIT0;
		if ((keyevent >> 29) != TermEvent) en.meta |= SP_EVAL_META_ITEMSFULL;
		if (generic) en.meta |= SP_EVAL_META_GENERIC;
		else if (disposed || ntrig == 0) en.meta |= SP_EVAL_META_DEAD;
		return en;
	}

	struct Node
	{
		uint32_t event, program, variable;
		explicit Node( uint32_t event_, uint32_t program_=0) :event(event_),program(program_),variable(0){}
	};
	struct Rule
	{
		uint32_t initsigval, initcount, event, resultHandle, formatHandle, range;
		std::vector<SpTriggerDef> defs;		// definition order (argument 0 first)
	};
	typedef std::unordered_map<uint32_t,uint32_t> EventListMap;	// event -> index into m_lists; same container and insertion
									// sequence as the reference so that compile() visits events in the same order

	// createProgram + createTrigger* + doneProgram; returns the 1-based program index
	uint32_t addRule( const Rule& rule)
	{
		m_rules.push_back( rule);
		uint32_t program = (uint32_t)m_rules.size();
		for (std::size_t di = rule.defs.size(); di > 0; --di)
		{
			if (rule.defs[ di-1].is_key) {addEventProgram( rule.defs[ di-1].event, program, 0); ++m_totalNofPrograms;}
		}
		return program;
	}
	void addEventProgram( uint32_t eventid, uint32_t program, uint32_t past_event)
	{
		SpProgramTrigger pt; pt.program = program; pt.past_event = past_event;
		EventListMap::iterator ei = m_eventPrograms.find( eventid);
		if (ei == m_eventPrograms.end())
		{
			m_lists.push_back( std::vector<SpProgramTrigger>( 1, pt));
			m_eventPrograms[ eventid] = (uint32_t)m_lists.size()-1;
		}
		else m_lists[ ei->second].push_back( pt);
		m_keyOccurrence[ eventid] += 1;
		if (past_event)
		{
			m_keyOccurrence[ past_event] -= 1;
			m_stopwords.insert( past_event);
		}
	}
	double eventWeight( uint32_t eventid) const
	{
		double kf = 1.0;
		std::map<uint32_t,double>::const_iterator fi = m_frequency.find( eventid);
		if (fi != m_frequency.end() && fi->second > 0.0) kf = fi->second;
		std::map<uint32_t,uint32_t>::const_iterator ki = m_keyOccurrence.find( eventid);
		if (ki != m_keyOccurrence.end() && ki->second > 0.0) kf *= ki->second;
		return kf;
	}
	// getAltEventId: the event (other than the key) whose trigger has the highest signal value; 0 = none
	uint32_t alternativeKeyEvent( uint32_t eventid, const Rule& rule) const
	{
		uint32_t selected = 0, selectedSigval = 0;
		uint8_t sigtype = SP_SIG_ANY;
		for (std::size_t di = rule.defs.size(); di > 0; --di)
		{
			const SpTriggerDef& def = rule.defs[ di-1];
			bool sequenceCase = false;
			switch (def.sigtype)
			{
				case SP_SIG_ANY: return 0;
				case SP_SIG_AND:
					if (sigtype == def.sigtype) {if (def.event != eventid) {selected = def.event; sigtype = def.sigtype;}}
					else if (sigtype == SP_SIG_ANY && def.event != eventid) {selected = def.event; sigtype = def.sigtype;}
					else return 0;
					sequenceCase = true;	// the reference falls through into the sequence case (no break, :371-391)
					break;
				case SP_SIG_SEQUENCE: case SP_SIG_SEQUENCE_IMM: case SP_SIG_WITHIN:
					sequenceCase = true;
					break;
				default: break;	// SigDel
			}
			if (!sequenceCase) continue;
			if (sigtype == def.sigtype)
			{
				if (selectedSigval < def.sigval && def.event != eventid) {selected = def.event; selectedSigval = def.sigval; sigtype = def.sigtype;}
			}
			else if (sigtype == SP_SIG_ANY && def.event != eventid) {selected = def.event; selectedSigval = def.sigval; sigtype = def.sigtype;}
			else if (sigtype == SP_SIG_SEQUENCE_IMM && def.sigtype == SP_SIG_SEQUENCE)
			{
				if (selectedSigval < def.sigval && def.event != eventid) {selected = def.event; selectedSigval = def.sigval; sigtype = def.sigtype;}
			}
			else return 0;
		}
		return selected;
	}
	void eliminateUnusedEvents()
	{
		std::set<uint32_t> used, programs;
		for (EventListMap::const_iterator ei = m_eventPrograms.begin(); ei != m_eventPrograms.end(); ++ei)
		{
			for (const SpProgramTrigger& pt: m_lists[ ei->second])
			{
				programs.insert( pt.program);
				for (const SpTriggerDef& def: m_rules[ pt.program-1].defs) used.insert( def.event);
			}
		}
		for (uint32_t program: programs)
		{
			Rule& rule = m_rules[ program-1];
			if (used.find( rule.event) == used.end()) rule.event = 0;
		}
	}

	std::vector<Node> m_stack;
	std::vector<Rule> m_rules;
	std::vector<std::vector<SpProgramTrigger> > m_lists;	// program lists in push order (1-based program indices)
	EventListMap m_eventPrograms;
	std::set<uint32_t> m_stopwords;
	std::map<uint32_t,uint32_t> m_keyOccurrence, m_eventOccurrence;
	std::map<uint32_t,double> m_frequency;
	NameTable m_patterns, m_variables;
	std::vector<std::string> m_formats;
	std::vector<ResultFormat> m_parsedFormats;
	uint32_t m_expressionEventCnt, m_totalNofPrograms;
	float m_stopwordOccurrenceFactor, m_weightFactor;
	uint32_t m_maxRange, m_maxResultSize;
	bool m_exclusive, m_flattened;
	MatcherTablesHost m_tables;
};

} // namespace sp
#endif
