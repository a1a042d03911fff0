/* Interface shim (strusBase SymbolTable): exact-match string -> dense 1-based id in
 * insertion order, and back. Same observable behaviour as the strusBase class for the
 * calls the pattern module makes (getOrCreate/get/key/size). */
#ifndef _STRUS_BASE_SYMBOL_TABLE_HPP_INCLUDED
#define _STRUS_BASE_SYMBOL_TABLE_HPP_INCLUDED
#include <string>
#include <vector>
#include <deque>
#include <unordered_map>
#include <stdint.h>
namespace strus {
class ErrorBufferInterface;
class SymbolTable
{
public:
	explicit SymbolTable( ErrorBufferInterface* errorhnd_=0) :m_errorhnd(errorhnd_){}
	virtual ~SymbolTable(){}
	uint32_t getOrCreate( const std::string& key_)
	{
		std::unordered_map<std::string,uint32_t>::const_iterator it = m_map.find( key_);
		if (it != m_map.end()) return it->second;
		m_keys.push_back( key_);
		uint32_t id = (uint32_t)m_keys.size();
		m_map[ key_] = id;
		return id;
	}
	uint32_t getOrCreate( const char* key_, std::size_t keysize) {return getOrCreate( std::string( key_, keysize));}
	uint32_t get( const std::string& key_) const
	{
		std::unordered_map<std::string,uint32_t>::const_iterator it = m_map.find( key_);
		return it == m_map.end() ? 0 : it->second;
	}
	uint32_t get( const char* key_, std::size_t keysize) const {return get( std::string( key_, keysize));}
	const char* key( uint32_t id) const {return (id == 0 || id > m_keys.size()) ? 0 : m_keys[ id-1].c_str();}
	std::size_t size() const {return m_keys.size();}
private:
	SymbolTable( const SymbolTable&);
	void operator=( const SymbolTable&);
	ErrorBufferInterface* m_errorhnd;
	std::unordered_map<std::string,uint32_t> m_map;
	std::deque<std::string> m_keys;		///< deque: c_str() pointers stay valid on growth
};
}
#endif
