/* Interface shim (strusBase ErrorBufferInterface): the error sink every strus interface
 * method reports into instead of throwing (reference convention: src/errorUtils.hpp:22-140). */
#ifndef _STRUS_ERROR_BUFFER_INTERFACE_HPP_INCLUDED
#define _STRUS_ERROR_BUFFER_INTERFACE_HPP_INCLUDED
#include "strus/errorCodes.hpp"
#include <cstdio>
namespace strus {
class DebugTraceInterface;
class ErrorBufferInterface
{
public:
	virtual ~ErrorBufferInterface(){}
	virtual void setLogFile( FILE* hnd)=0;
	virtual bool setMaxNofThreads( unsigned int maxNofThreads)=0;
	virtual void report( int errorcode, const char* format, ...)
#ifdef __GNUC__
		__attribute__ ((format (printf, 3, 4)))
#endif
		=0;
	virtual void explain( const char* format)=0;
	virtual const char* fetchError()=0;
	virtual bool hasError() const=0;
	virtual DebugTraceInterface* debugTrace() const=0;
};
}
#endif
