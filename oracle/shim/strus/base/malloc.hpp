/* Oracle build shim (strusBase): aligned allocation used by the trigger event arrays. */
#ifndef _STRUS_BASE_MALLOC_HPP_INCLUDED
#define _STRUS_BASE_MALLOC_HPP_INCLUDED
#include <cstdlib>
namespace strus {
inline void* aligned_malloc( std::size_t size, std::size_t alignment)
{
	void* rt = 0;
	if (size == 0) size = alignment;
	if (0 != posix_memalign( &rt, alignment, size)) return 0;
	return rt;
}
inline void aligned_free( void* ptr) {std::free( ptr);}
}
#endif
