/* Interface shim (strusBase error codes). Numeric values are placeholders; link against the
 * real strus/errorCodes.hpp in a strus installation. */
#ifndef _STRUS_ERROR_CODES_HPP_INCLUDED
#define _STRUS_ERROR_CODES_HPP_INCLUDED
namespace strus {
enum ErrorCode
{
	ErrorCodeUnknown=1,
	ErrorCodeRuntimeError=2,
	ErrorCodeLogicError=3,
	ErrorCodeUncaughtException=4,
	ErrorCodeOutOfMem=12,
	ErrorCodeInvalidArgument=22,
	ErrorCodeSyntax=30,
	ErrorCodeNotImplemented=31,
	ErrorCodeVersionMismatch=32,
	ErrorCodePlatformIncompatibility=33,
	ErrorCodePlatformRequirements=34,
	ErrorCodeBufferOverflow=35,
	ErrorCodeIOError=36
};
}
#endif
