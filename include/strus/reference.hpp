/* Interface shim (strusBase Reference): shared ownership pointer. */
#ifndef _STRUS_REFERENCE_HPP_INCLUDED
#define _STRUS_REFERENCE_HPP_INCLUDED
#include <memory>
namespace strus {
template <class T>
class Reference :public std::shared_ptr<T>
{
public:
	Reference() {}
	Reference( T* p) :std::shared_ptr<T>(p){}
	Reference( const Reference& o) :std::shared_ptr<T>(o){}
};
}
#endif
