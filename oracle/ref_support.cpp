// ORACLE (test infrastructure, not product code).
// Support code needed to link the reference's own level-2 sources
// (/root/reference/src/patternMatcher.cpp, ruleMatcherAutomaton.cpp) outside of a strus
// installation: strus::runtime_error/logic_error (declared in the reference's
// src/internationalization.hpp:26-48; its implementation file is a cmake template and cannot be
// compiled as is) and a minimal ErrorBufferInterface.
#include "internationalization.hpp"
#include "strus/errorBufferInterface.hpp"
#include "strus/lib/error.hpp"
#include <cstdarg>
#include <cstdio>
#include <string>

std::runtime_error strus::runtime_error( const char* format, ...)
{
	char buffer[ 4096];
	va_list args;
	va_start( args, format);
	int n = vsnprintf( buffer, sizeof(buffer), format, args);
	va_end( args);
	if (n < 0) n = 0;
	if ((size_t)n >= sizeof(buffer)) n = sizeof(buffer)-1;
	return std::runtime_error( std::string( buffer, n));
}

std::runtime_error strus::runtime_error_ec( int apperrno, const char* format, ...)
{
	char buffer[ 4096];
	int pos = apperrno ? snprintf( buffer, sizeof(buffer), "[#%d] ", apperrno) : 0;
	va_list args;
	va_start( args, format);
	int n = vsnprintf( buffer+pos, sizeof(buffer)-pos, format, args);
	va_end( args);
	if (n < 0) n = 0;
	n += pos;
	if ((size_t)n >= sizeof(buffer)) n = sizeof(buffer)-1;
	return std::runtime_error( std::string( buffer, n));
}

std::logic_error strus::logic_error( const char* format, ...)
{
	char buffer[ 1024];
	va_list args;
	va_start( args, format);
	int n = vsnprintf( buffer, sizeof(buffer), format, args);
	va_end( args);
	if (n < 0) n = 0;
	if ((size_t)n >= sizeof(buffer)) n = sizeof(buffer)-1;
	return std::logic_error( std::string( buffer, n));
}

void strus::initMessageTextDomain(){}

namespace {
// One error slot per thread (the reference's buffer is created with a max thread count,
// tests/randomTokenPatternMatch/src/testRandomTokenPatternMatch.cpp:300).
class OracleErrorBuffer :public strus::ErrorBufferInterface
{
public:
	OracleErrorBuffer(){}
	virtual ~OracleErrorBuffer(){}
	virtual void setLogFile( FILE*){}
	virtual bool setMaxNofThreads( unsigned int){return true;}
	virtual void report( int errorcode, const char* format, ...)
	{
		char buffer[ 4096];
		va_list args;
		va_start( args, format);
		vsnprintf( buffer, sizeof(buffer), format, args);
		va_end( args);
		if (!slot().has) {slot().msg = buffer; slot().code = errorcode; slot().has = true;}
	}
	virtual void explain( const char* format)
	{
		if (slot().has)
		{
			char buffer[ 4096];
			snprintf( buffer, sizeof(buffer), format, slot().msg.c_str());
			slot().msg = buffer;
		}
	}
	virtual const char* fetchError()
	{
		if (!slot().has) return 0;
		slot().has = false;
		slot().fetched = slot().msg;
		return slot().fetched.c_str();
	}
	virtual bool hasError() const {return slot().has;}
	virtual strus::DebugTraceInterface* debugTrace() const {return 0;}
private:
	struct Slot {bool has; int code; std::string msg; std::string fetched; Slot():has(false),code(0){}};
	static Slot& slot() {static thread_local Slot s; return s;}
};
}

strus::ErrorBufferInterface* strus::createErrorBuffer_standard( FILE*, std::size_t, strus::DebugTraceInterface*)
{
	return new OracleErrorBuffer();
}
