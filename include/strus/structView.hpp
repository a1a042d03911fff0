/* Interface shim (strusBase StructView): introspection value, name/value pairs only. */
#ifndef _STRUS_STRUCT_VIEW_HPP_INCLUDED
#define _STRUS_STRUCT_VIEW_HPP_INCLUDED
#include <string>
#include <vector>
#include <utility>
namespace strus {
class StructView
{
public:
	StructView(){}
	StructView( int){}
	StructView( const StructView& o) :m_ar(o.m_ar){}
	StructView& operator()( const char* name, const char* value)
		{m_ar.push_back( std::pair<std::string,std::string>( name, value?value:"")); return *this;}
	StructView& operator()( const std::string& name, const std::string& value)
		{m_ar.push_back( std::pair<std::string,std::string>( name, value)); return *this;}
	const char* get( const std::string& name) const
	{
		for (std::size_t i=0; i<m_ar.size(); ++i) if (m_ar[i].first == name) return m_ar[i].second.c_str();
		return 0;
	}
	std::size_t size() const {return m_ar.size();}
private:
	std::vector<std::pair<std::string,std::string> > m_ar;
};
}
#endif
