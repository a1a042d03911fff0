/* Interface shim (strusBase lib/error.hpp): standard error buffer factory. The implementation
 * shipped here (struspattern_b200/csrc/error_buffer.cpp) is a small single-slot-per-thread buffer. */
#ifndef _STRUS_LIB_ERROR_HPP_INCLUDED
#define _STRUS_LIB_ERROR_HPP_INCLUDED
#include <cstdio>
namespace strus {
class ErrorBufferInterface;
class DebugTraceInterface;
ErrorBufferInterface* createErrorBuffer_standard( FILE* logfilehandle, std::size_t maxNofThreads, DebugTraceInterface* dbgtrace);
}
#endif
