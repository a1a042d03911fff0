/* Interface shim (strusBase debug trace). */
#ifndef _STRUS_DEBUG_TRACE_INTERFACE_HPP_INCLUDED
#define _STRUS_DEBUG_TRACE_INTERFACE_HPP_INCLUDED
namespace strus {
class DebugTraceContextInterface
{
public:
	virtual ~DebugTraceContextInterface(){}
	virtual void event( const char* eventid, const char* format, ...)
#ifdef __GNUC__
		__attribute__ ((format (printf, 3, 4)))
#endif
		=0;
	virtual void open( const char* blockid, const char* name=0)=0;
	virtual void close()=0;
};
class DebugTraceInterface
{
public:
	virtual ~DebugTraceInterface(){}
	virtual DebugTraceContextInterface* createTraceContext( const char* component) const=0;
};
}
#endif
