// The strus C++ interfaces of the pattern module on top of the C ABI (include/strus_pattern_b200.h):
//   strus::createPatternLexer_std / createPatternMatcher_std        (reference: src/libstrus_pattern.cpp:21-47)
//   PatternLexer -> PatternLexerInstance -> PatternLexerContext       (reference: src/patternLexer.cpp:681-1183)
//   PatternMatcher -> PatternMatcherInstance -> PatternMatcherContext (reference: src/patternMatcher.cpp:107-730)
//   extern "C" strus::AnalyzerModule entryPoint                       (reference: src/modstrus_analyzer_pattern.cpp:16-61)
// Same method names, argument meaning and error behaviour: no exception crosses an interface method, errors are
// reported into the ErrorBufferInterface (reference: src/errorUtils.hpp:22-140) and the method returns 0/false/empty.
// A context is a batch of one document: match() scans one segment, putInput() buffers the lexems of the document
// and fetchResults() runs them through the device. Callers that want throughput use the batch entry points of the C ABI.
#include "strus/lib/pattern.hpp"
#include "strus/lib/error.hpp"
#include "strus/errorBufferInterface.hpp"
#include "strus/patternLexerInterface.hpp"
#include "strus/patternLexerInstanceInterface.hpp"
#include "strus/patternLexerContextInterface.hpp"
#include "strus/patternMatcherInterface.hpp"
#include "strus/patternMatcherInstanceInterface.hpp"
#include "strus/patternMatcherContextInterface.hpp"
#include "strus/analyzerModule.hpp"
#include "strus/base/dll_tags.hpp"
#include "strus_pattern_b200.h"
#include <string>
#include <vector>
#include <map>
#include <deque>
#include <stdexcept>
#include <limits>
#include <new>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <thread>

namespace {

int errorCodeOf( int rc)
{
	switch (rc)
	{
		case SP_ERR_SYNTAX: return strus::ErrorCodeSyntax;
		case SP_ERR_NOMEM: return strus::ErrorCodeOutOfMem;
		case SP_ERR_UNSUPPORTED: return strus::ErrorCodeNotImplemented;
		case SP_ERR_CUDA: return strus::ErrorCodePlatformRequirements;
		default: return strus::ErrorCodeRuntimeError;
	}
}

struct CallError :public std::runtime_error
{
	int rc;
	CallError( int rc_, const std::string& msg) :std::runtime_error(msg),rc(rc_){}
};

// exception -> error buffer, the convention of CATCH_ERROR_MAP (src/errorUtils.hpp:22-140)
#define SP_CATCH_REPORT( msg, errorhnd) \
	catch (const std::bad_alloc&) {(errorhnd)->report( strus::ErrorCodeOutOfMem, "%s: %s", msg, "memory allocation error");} \
	catch (const CallError& e) {(errorhnd)->report( errorCodeOf( e.rc), "%s: %s", msg, e.what());} \
	catch (const std::runtime_error& e) {(errorhnd)->report( strus::ErrorCodeRuntimeError, "%s: %s", msg, e.what());} \
	catch (const std::logic_error& e) {(errorhnd)->report( strus::ErrorCodeLogicError, "%s: %s", msg, e.what());} \
	catch (const std::exception& e) {(errorhnd)->report( strus::ErrorCodeUncaughtException, "%s: %s", msg, e.what());}

// ------------------------------------------------------------------------------------------------ lexer
class LexerContext :public strus::PatternLexerContextInterface
{
public:
	LexerContext( const sp_lexer* lexer_, strus::ErrorBufferInterface* errorhnd_)
		:m_errorhnd(errorhnd_),m_ctx(sp_lexer_ctx_create( lexer_, 0))
	{
		if (!m_ctx) throw CallError( SP_ERR_CUDA, sp_lexer_last_error( lexer_));
	}
	virtual ~LexerContext() {sp_lexer_ctx_destroy( m_ctx);}

	virtual std::vector<strus::analyzer::PatternLexem> match( const char* src, std::size_t srclen)
	{
		try
		{
			std::vector<strus::analyzer::PatternLexem> rt;
			if (srclen >= (std::size_t)std::numeric_limits<uint32_t>::max()) throw std::runtime_error( "size of string to scan out of range");
			uint64_t offsets[2] = {0, (uint64_t)srclen};
			int rc = sp_lexer_scan( m_ctx, (const uint8_t*)src, 0, offsets, 1);
			if (rc != SP_OK) throw CallError( rc, sp_lexer_ctx_last_error( m_ctx));
			int64_t n = sp_lexer_result_count( m_ctx);
			if (n < 0) throw CallError( (int)n, sp_lexer_ctx_last_error( m_ctx));
			std::vector<sp_lexem> lexems( (std::size_t)n);
			uint32_t status = 0;
			rc = sp_lexer_fetch( m_ctx, lexems.data(), 0, &status);
			if (rc != SP_OK) throw CallError( rc, sp_lexer_ctx_last_error( m_ctx));
			if (status) throw std::runtime_error( "size of matched term out of range");
			rt.reserve( lexems.size());
			for (const sp_lexem& lx: lexems)
			{
				rt.push_back( strus::analyzer::PatternLexem( lx.id, (int)lx.ordpos, strus::analyzer::Position( 0, (int)lx.origpos), (int)lx.origsize));
			}
			return rt;
		}
		SP_CATCH_REPORT( "failed to run pattern matching terms with regular expressions", m_errorhnd)
		return std::vector<strus::analyzer::PatternLexem>();
	}
	virtual void reset() {}		// nothing is kept between two match calls

private:
	strus::ErrorBufferInterface* m_errorhnd;
	sp_lexer_ctx* m_ctx;
};

class LexerInstance :public strus::PatternLexerInstanceInterface
{
public:
	explicit LexerInstance( strus::ErrorBufferInterface* errorhnd_)
		:m_errorhnd(errorhnd_),m_lexer(sp_lexer_create()),m_compiled(false)
	{
		if (!m_lexer) throw std::bad_alloc();
	}
	virtual ~LexerInstance() {sp_lexer_destroy( m_lexer);}

	virtual void defineOption( const std::string& name_, double value)
	{
		try {check( sp_lexer_define_option( m_lexer, name_.c_str(), value));}
		SP_CATCH_REPORT( "error in define option", m_errorhnd)
	}
	virtual void defineLexemName( unsigned int id, const std::string& name_)
	{
		try {m_lexemNames[ id] = name_;}
		SP_CATCH_REPORT( "failed to assign name to lexem", m_errorhnd)
	}
	virtual const char* getLexemName( unsigned int id) const
	{
		std::map<unsigned int,std::string>::const_iterator it = m_lexemNames.find( id);
		return it == m_lexemNames.end() ? 0 : it->second.c_str();
	}
	virtual void defineLexem( unsigned int id, const std::string& expression, unsigned int resultIndex, unsigned int level, strus::analyzer::PositionBind posbind)
	{
		try
		{
			if (m_compiled) throw std::runtime_error( "called define pattern after calling 'compile'");
			check( sp_lexer_define_lexem( m_lexer, id, expression.c_str(), expression.size(), resultIndex, level, (uint32_t)posbind));
		}
		SP_CATCH_REPORT( "failed to define term match regular expression pattern", m_errorhnd)
	}
	virtual void defineSymbol( unsigned int symbolid, unsigned int lexemid, const std::string& name_)
	{
		try
		{
			if (m_compiled) throw std::runtime_error( "called define symbol after calling 'compile'");
			check( sp_lexer_define_symbol( m_lexer, symbolid, lexemid, name_.c_str(), name_.size()));
		}
		SP_CATCH_REPORT( "failed to define term match regular expression pattern", m_errorhnd)
	}
	virtual unsigned int getSymbol( unsigned int lexemid, const std::string& name_) const
	{
		return sp_lexer_get_symbol( m_lexer, lexemid, name_.c_str(), name_.size());
	}
	virtual bool compile()
	{
		try
		{
			check( sp_lexer_compile( m_lexer));
			m_compiled = true;
			return true;
		}
		SP_CATCH_REPORT( "failed to compile regular expression patterns", m_errorhnd)
		return false;
	}
	virtual strus::PatternLexerContextInterface* createContext() const
	{
		try
		{
			if (!m_compiled) throw std::runtime_error( "called create context without calling 'compile'");
			return new LexerContext( m_lexer, m_errorhnd);
		}
		SP_CATCH_REPORT( "failed to create term match context", m_errorhnd)
		return 0;
	}
	virtual const char* name() const {return "std";}
	virtual strus::StructView view() const {return strus::StructView()( "name", name());}

private:
	void check( int rc) const {if (rc != SP_OK) throw CallError( rc, sp_lexer_last_error( m_lexer));}

	strus::ErrorBufferInterface* m_errorhnd;
	sp_lexer* m_lexer;
	bool m_compiled;
	std::map<unsigned int,std::string> m_lexemNames;
};

class Lexer :public strus::PatternLexerInterface
{
public:
	explicit Lexer( strus::ErrorBufferInterface* errorhnd_) :m_errorhnd(errorhnd_){}
	virtual std::vector<std::string> getCompileOptionNames() const
	{
		static const char* ar[] = {"CASELESS","DOTALL","MULTILINE","ALLOWEMPTY","UCP",0};	// src/patternLexer.cpp:1157
		std::vector<std::string> rt;
		for (int i=0; ar[i]; ++i) rt.push_back( ar[i]);
		return rt;
	}
	virtual strus::PatternLexerInstanceInterface* createInstance() const
	{
		try {return new LexerInstance( m_errorhnd);}
		SP_CATCH_REPORT( "failed to create term match instance", m_errorhnd)
		return 0;
	}
	virtual const char* name() const {return "std";}
	virtual strus::StructView view() const
	{
		return strus::StructView()( "name", name())( "description", "pattern lexer based on regular expression automata running on the GPU");
	}
private:
	strus::ErrorBufferInterface* m_errorhnd;
};

// ------------------------------------------------------------------------------------------------ matcher
class MatcherContext :public strus::PatternMatcherContextInterface
{
public:
	MatcherContext( sp_matcher* matcher_, strus::ErrorBufferInterface* errorhnd_)
		:m_errorhnd(errorhnd_),m_matcher(matcher_),m_ctx(sp_matcher_ctx_create( matcher_, 0)),m_curPosition(0),m_dirty(true)
	{
		if (!m_ctx) throw CallError( SP_ERR_CUDA, sp_matcher_last_error( matcher_));
		// the same automatic choice as the batch interface: a document of a few thousand tokens under a large rule
		// set takes milliseconds on the window evaluator and the better part of a second on one warp of the state machine;
		// SP_MATCHER_MODE=statemachine gives the reference's emission order
		std::memset( &m_stats, 0, sizeof(m_stats));
	}
	virtual ~MatcherContext() {sp_matcher_ctx_destroy( m_ctx);}

	// the checks of the reference's putInput (src/patternMatcher.cpp:131-162) are done here, where the error belongs
	// to the call; a rejected token is not part of the document, exactly as in the reference
	virtual void putInput( const strus::analyzer::PatternLexem& term)
	{
		try
		{
			if (m_curPosition > (unsigned int)term.ordpos())
			{
				char buf[ 128];
				std::snprintf( buf, sizeof(buf), "term events not fed in ascending order (%u > %u)", m_curPosition, (unsigned int)term.ordpos());
				throw std::runtime_error( buf);
			}
			else if (m_curPosition < (unsigned int)term.ordpos()) m_curPosition = term.ordpos();
			else if (term.origsize() >= std::numeric_limits<int32_t>::max()) throw std::runtime_error( "term event orig size out of range");
			else if (term.origpos().seg() >= std::numeric_limits<int32_t>::max()) throw std::runtime_error( "term event orig segment number out of range");
			else if (term.origpos().ofs() >= std::numeric_limits<int32_t>::max()) throw std::runtime_error( "term event orig segment byte position out of range");
			if (term.id() >= (1u<<29)) throw std::runtime_error( "event handle out of range");
			sp_lexem lx; lx.id = term.id(); lx.ordpos = (uint32_t)term.ordpos(); lx.origpos = (uint32_t)term.origpos().ofs(); lx.origsize = (uint32_t)term.origsize();
			m_tokens.push_back( lx);
			m_segs.push_back( (uint32_t)term.origpos().seg());
			m_dirty = true;
		}
		SP_CATCH_REPORT( "failed to feed input to pattern matcher", m_errorhnd)
	}
	virtual std::vector<strus::analyzer::PatternMatcherResult> fetchResults()
	{
		try
		{
			run();
			std::vector<strus::analyzer::PatternMatcherResult> rt;
			rt.reserve( m_results.size());
			m_values.clear();	// the value strings live in the context, like the reference's PatternResultFormatContext buffer
			for (const sp_result& r: m_results)
			{
				std::vector<strus::analyzer::PatternMatcherResultItem> items;
				const char* resultValue = 0;
				if (r.format)
				{
					// the items are the material of the value (src/patternMatcher.cpp:253-262)
					resultValue = valueString( sp_matcher_result_value( m_matcher, &r, m_items.data(), 0, 0), &r, 0);
				}
				else
				{
					items.reserve( r.item_count);
					for (uint32_t k=0; k<r.item_count; k += 1 + SP_ITEM_NESTED( m_items[ r.item_begin + k]))
					{
						const sp_item& it = m_items[ r.item_begin + k];
						const char* itemValue = 0;
						if (SP_ITEM_FORMAT( it)) itemValue = valueString( sp_matcher_item_value( m_matcher, &it, 0, 0), 0, &it);	// :172-181
						items.push_back( strus::analyzer::PatternMatcherResultItem(
							sp_matcher_variable_name( m_matcher, it.variable), itemValue, (int)it.ordpos, (int)it.ordend,
							strus::analyzer::Position( (int)it.origseg, (int)it.origpos), strus::analyzer::Position( (int)it.endseg, (int)it.endpos)));
					}
				}
				rt.push_back( strus::analyzer::PatternMatcherResult(
					sp_matcher_pattern_name( m_matcher, r.name), resultValue, (int)r.ordpos, (int)r.ordend,
					strus::analyzer::Position( (int)r.origseg, (int)r.origpos), strus::analyzer::Position( (int)r.endseg, (int)r.endpos), items));
			}
			return rt;
		}
		SP_CATCH_REPORT( "failed to fetch pattern match result", m_errorhnd)
		return std::vector<strus::analyzer::PatternMatcherResult>();
	}
	virtual strus::analyzer::PatternMatcherStatistics getStatistics() const
	{
		try
		{
			const_cast<MatcherContext*>( this)->run();
			strus::analyzer::PatternMatcherStatistics stats;
			stats.define( "nofProgramsInstalled", (double)m_stats.nof_programs_installed);
			stats.define( "nofAltKeyProgramsInstalled", (double)m_stats.nof_altkey_programs_installed);
			stats.define( "nofSignalsFired", (double)m_stats.nof_signals_fired);
			if (m_stats.nof_events) stats.define( "nofTriggersAvgActive", (double)m_stats.nof_open_patterns / (double)m_stats.nof_events);
			return stats;
		}
		SP_CATCH_REPORT( "failed to get pattern match statistics", m_errorhnd)
		return strus::analyzer::PatternMatcherStatistics();
	}
	virtual void reset()
	{
		m_tokens.clear(); m_segs.clear(); m_results.clear(); m_items.clear(); m_values.clear();
		std::memset( &m_stats, 0, sizeof(m_stats));
		m_curPosition = 0; m_dirty = true;
	}

private:
	void run()
	{
		if (!m_dirty) return;
		uint64_t offsets[2] = {0, (uint64_t)m_tokens.size()};
		int rc = sp_matcher_run( m_ctx, m_tokens.data(), m_segs.data(), offsets, 1, 0);
		if (rc != SP_OK) throw CallError( rc, sp_matcher_ctx_last_error( m_ctx));
		int64_t nres = sp_matcher_result_count( m_ctx), nitems = sp_matcher_item_count( m_ctx);
		if (nres < 0 || nitems < 0) throw CallError( (int)nres, sp_matcher_ctx_last_error( m_ctx));
		m_results.resize( (std::size_t)nres); m_items.resize( (std::size_t)nitems);
		uint32_t status = 0;
		rc = sp_matcher_fetch( m_ctx, m_results.data(), 0, m_items.data(), 0, &m_stats, &status);
		if (rc != SP_OK) throw CallError( rc, sp_matcher_ctx_last_error( m_ctx));
		m_dirty = false;
		if (status) throw std::runtime_error( "error in pattern matching automaton (rule set inconsistency or limit exceeded)");
	}
	// value string of a result (r) or of an item (it) of the length the first call told; kept until the next fetchResults
	const char* valueString( int64_t len, const sp_result* r, const sp_item* it)
	{
		if (len == SP_NO_VALUE) return 0;
		if (len < 0) throw CallError( (int)len, sp_matcher_last_error( m_matcher));
		m_values.push_back( std::string( (std::size_t)len, '\0'));
		std::string& val = m_values.back();
		if (len) (void)(r ? sp_matcher_result_value( m_matcher, r, m_items.data(), &val[0], (uint64_t)len) : sp_matcher_item_value( m_matcher, it, &val[0], (uint64_t)len));
		return val.c_str();
	}

	strus::ErrorBufferInterface* m_errorhnd;
	sp_matcher* m_matcher;
	sp_matcher_ctx* m_ctx;
	std::deque<std::string> m_values;
	unsigned int m_curPosition;
	bool m_dirty;
	std::vector<sp_lexem> m_tokens;
	std::vector<uint32_t> m_segs;
	std::vector<sp_result> m_results;
	std::vector<sp_item> m_items;
	sp_matcher_stats m_stats;
};

class MatcherInstance :public strus::PatternMatcherInstanceInterface
{
public:
	explicit MatcherInstance( strus::ErrorBufferInterface* errorhnd_)
		:m_errorhnd(errorhnd_),m_matcher(sp_matcher_create())
	{
		if (!m_matcher) throw std::bad_alloc();
	}
	virtual ~MatcherInstance() {sp_matcher_destroy( m_matcher);}

	virtual void defineOption( const std::string& name_, double value)
	{
		try {check( sp_matcher_define_option( m_matcher, name_.c_str(), value));}
		SP_CATCH_REPORT( "failed to define pattern matching automaton option", m_errorhnd)
	}
	virtual void defineTermFrequency( unsigned int termid, double df)
	{
		try {check( sp_matcher_define_term_frequency( m_matcher, termid, df));}
		SP_CATCH_REPORT( "failed to define term frequency", m_errorhnd)
	}
	virtual void pushTerm( unsigned int termid)
	{
		try {check( sp_matcher_push_term( m_matcher, termid));}
		SP_CATCH_REPORT( "failed to push term on the pattern match expression stack", m_errorhnd)
	}
	virtual void pushExpression( JoinOperation joinop, std::size_t argc, unsigned int range, unsigned int cardinality)
	{
		try {check( sp_matcher_push_expression( m_matcher, (int)joinop, argc, range, cardinality));}
		SP_CATCH_REPORT( "failed to push expression on the pattern match expression stack", m_errorhnd)
	}
	virtual void pushPattern( const std::string& name_)
	{
		try {check( sp_matcher_push_pattern( m_matcher, name_.c_str()));}
		SP_CATCH_REPORT( "failed to push pattern reference on the pattern match expression stack", m_errorhnd)
	}
	virtual void attachVariable( const std::string& name_)
	{
		try {check( sp_matcher_attach_variable( m_matcher, name_.c_str()));}
		SP_CATCH_REPORT( "failed to attach variable to top element of the pattern match expression stack", m_errorhnd)
	}
	virtual void definePattern( const std::string& name_, const std::string& formatstring, bool visible)
	{
		try {check( sp_matcher_define_pattern( m_matcher, name_.c_str(), formatstring.c_str(), visible ? 1 : 0));}
		SP_CATCH_REPORT( "failed to close pattern definition on the pattern match expression stack", m_errorhnd)
	}
	virtual bool compile()
	{
		try {check( sp_matcher_compile( m_matcher)); return true;}
		SP_CATCH_REPORT( "failed to compile (optimize) pattern matching automaton", m_errorhnd)
		return false;
	}
	virtual strus::PatternMatcherContextInterface* createContext() const
	{
		try {return new MatcherContext( m_matcher, m_errorhnd);}
		SP_CATCH_REPORT( "failed to create pattern match context", m_errorhnd)
		return 0;
	}
	virtual const char* name() const {return "std";}
	virtual strus::StructView view() const {return strus::StructView()( "name", name());}

private:
	void check( int rc) const {if (rc != SP_OK) throw CallError( rc, sp_matcher_last_error( m_matcher));}

	strus::ErrorBufferInterface* m_errorhnd;
	sp_matcher* m_matcher;
};

class Matcher :public strus::PatternMatcherInterface
{
public:
	explicit Matcher( strus::ErrorBufferInterface* errorhnd_) :m_errorhnd(errorhnd_){}
	virtual std::vector<std::string> getCompileOptionNames() const
	{
		static const char* ar[] = {"stopwordOccurrenceFactor","weightFactor","maxRange","exclusive","maxResultSize",0};	// src/patternMatcher.cpp:707-716
		std::vector<std::string> rt;
		for (int i=0; ar[i]; ++i) rt.push_back( ar[i]);
		return rt;
	}
	virtual strus::PatternMatcherInstanceInterface* createInstance() const
	{
		try {return new MatcherInstance( m_errorhnd);}
		SP_CATCH_REPORT( "failed to create pattern match instance", m_errorhnd)
		return 0;
	}
	virtual const char* name() const {return "std";}
	virtual strus::StructView view() const
	{
		return strus::StructView()( "name", name())( "description", "pattern matcher based on an event driven automaton running on the GPU");
	}
private:
	strus::ErrorBufferInterface* m_errorhnd;
};

// ------------------------------------------------------------------------------------------------ error buffer
// Minimal standard error buffer (strusBase createErrorBuffer_standard): one slot per thread, first error wins.
class StandardErrorBuffer :public strus::ErrorBufferInterface
{
public:
	StandardErrorBuffer( FILE* logfile_, strus::DebugTraceInterface* dbgtrace_) :m_logfile(logfile_),m_dbgtrace(dbgtrace_){}
	virtual ~StandardErrorBuffer(){}
	virtual void setLogFile( FILE* hnd) {m_logfile = hnd;}
	virtual bool setMaxNofThreads( unsigned int) {return true;}
	virtual void report( int errorcode, const char* format, ...)
	{
		char buf[ 2048];
		va_list ap;
		va_start( ap, format);
		std::vsnprintf( buf, sizeof(buf), format, ap);
		va_end( ap);
		std::lock_guard<std::mutex> lock( m_mutex);
		Slot& slot = m_slots[ std::this_thread::get_id()];
		if (m_logfile) std::fprintf( m_logfile, "[#%d] %s\n", errorcode, buf);
		if (!slot.code) {slot.code = errorcode; slot.msg = buf;}
	}
	virtual void explain( const char* format)
	{
		std::lock_guard<std::mutex> lock( m_mutex);
		Slot& slot = m_slots[ std::this_thread::get_id()];
		if (!slot.code) return;
		char buf[ 2048];
		std::snprintf( buf, sizeof(buf), format, slot.msg.c_str());
		slot.msg = buf;
	}
	virtual const char* fetchError()
	{
		std::lock_guard<std::mutex> lock( m_mutex);
		Slot& slot = m_slots[ std::this_thread::get_id()];
		if (!slot.code) return 0;
		slot.fetched = slot.msg;
		slot.code = 0; slot.msg.clear();
		return slot.fetched.c_str();
	}
	virtual bool hasError() const
	{
		std::lock_guard<std::mutex> lock( m_mutex);
		std::map<std::thread::id,Slot>::const_iterator it = m_slots.find( std::this_thread::get_id());
		return it != m_slots.end() && it->second.code != 0;
	}
	virtual strus::DebugTraceInterface* debugTrace() const {return m_dbgtrace;}
private:
	struct Slot {int code; std::string msg, fetched; Slot() :code(0){}};
	FILE* m_logfile;
	strus::DebugTraceInterface* m_dbgtrace;
	mutable std::mutex m_mutex;
	std::map<std::thread::id,Slot> m_slots;
};

const strus::PatternLexerConstructor g_lexerConstructor = {"std", strus::createPatternLexer_std};
const strus::PatternMatcherConstructor g_matcherConstructor = {"std", strus::createPatternMatcher_std};

} // namespace

namespace strus {

DLL_PUBLIC PatternLexerInterface* createPatternLexer_std( ErrorBufferInterface* errorhnd)
{
	try {return new Lexer( errorhnd);}
	SP_CATCH_REPORT( "cannot create pattern lexer", errorhnd)
	return 0;
}
DLL_PUBLIC PatternMatcherInterface* createPatternMatcher_std( ErrorBufferInterface* errorhnd)
{
	try {return new Matcher( errorhnd);}
	SP_CATCH_REPORT( "cannot create pattern matcher", errorhnd)
	return 0;
}
DLL_PUBLIC ErrorBufferInterface* createErrorBuffer_standard( FILE* logfilehandle, std::size_t, DebugTraceInterface* dbgtrace)
{
	try {return new StandardErrorBuffer( logfilehandle, dbgtrace);} catch (...) {return 0;}
}

} // namespace strus

// dlopen entry point of the strus analyzer module (src/modstrus_analyzer_pattern.cpp:59-61)
extern "C" DLL_PUBLIC strus::AnalyzerModule entryPoint;
strus::AnalyzerModule entryPoint( g_lexerConstructor, g_matcherConstructor,
	"strus_pattern_b200 0.1.0 (regex lexer and pattern automaton as CUDA kernels for sm_100a)",
	" strus_pattern_b200: no third party engine (Hyperscan and TRE are replaced by this library's own automata)\n");
