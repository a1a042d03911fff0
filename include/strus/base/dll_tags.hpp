/* Interface shim (strusBase): symbol visibility tags. */
#ifndef _STRUS_BASE_DLL_TAGS_HPP_INCLUDED
#define _STRUS_BASE_DLL_TAGS_HPP_INCLUDED
#if defined(__GNUC__) || defined(__clang__)
#define DLL_PUBLIC __attribute__ ((visibility ("default")))
#define DLL_LOCAL  __attribute__ ((visibility ("hidden")))
#else
#define DLL_PUBLIC
#define DLL_LOCAL
#endif
#endif
