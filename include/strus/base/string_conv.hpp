/* Interface shim (strusBase): the one string helper the pattern module uses. */
#ifndef _STRUS_BASE_STRING_CONV_HPP_INCLUDED
#define _STRUS_BASE_STRING_CONV_HPP_INCLUDED
#include <string>
#include <cctype>
namespace strus {
inline bool caseInsensitiveEquals( const std::string& a, const std::string& b)
{
	if (a.size() != b.size()) return false;
	for (std::size_t i=0; i<a.size(); ++i)
	{
		if (std::tolower((unsigned char)a[i]) != std::tolower((unsigned char)b[i])) return false;
	}
	return true;
}
}
#endif
