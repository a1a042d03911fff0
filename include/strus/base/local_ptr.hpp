/* Interface shim (strusBase): scoped owning pointer used by strus callers. */
#ifndef _STRUS_BASE_LOCAL_PTR_HPP_INCLUDED
#define _STRUS_BASE_LOCAL_PTR_HPP_INCLUDED
#include <memory>
namespace strus {
template <class T>
class local_ptr :public std::unique_ptr<T>
{
public:
	local_ptr() {}
	explicit local_ptr( T* p) :std::unique_ptr<T>(p){}
};
}
#endif
