/* Oracle build shim (strusBase): strus::unordered_map is std::unordered_map on C++11 compilers. */
#ifndef _STRUS_BASE_UNORDERED_MAP_HPP_INCLUDED
#define _STRUS_BASE_UNORDERED_MAP_HPP_INCLUDED
#include <unordered_map>
namespace strus {
template<typename Key, typename Elem, typename Hash = std::hash<Key>, typename Pred = std::equal_to<Key> >
class unordered_map :public std::unordered_map<Key,Elem,Hash,Pred>
{
public:
	unordered_map(){}
	unordered_map( const unordered_map& o) :std::unordered_map<Key,Elem,Hash,Pred>(o){}
};
}
#endif
