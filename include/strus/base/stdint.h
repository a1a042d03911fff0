/* Interface shim (strusBase): fixed width integer types. */
#ifndef _STRUS_BASE_STDINT_H_INCLUDED
#define _STRUS_BASE_STDINT_H_INCLUDED
#include <stdint.h>
#endif
