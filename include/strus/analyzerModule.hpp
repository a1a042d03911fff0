/* Interface shim (strusModule analyzerModule.hpp): layout of the dlopen entry point of an
 * analyzer module, reduced to what src/modstrus_analyzer_pattern.cpp:16-61 constructs.
 * The real struct (ModuleEntryPoint base with signature/version fields) comes from strusModule;
 * replace this header by the real one when building against a strus installation. */
#ifndef _STRUS_ANALYZER_MODULE_HPP_INCLUDED
#define _STRUS_ANALYZER_MODULE_HPP_INCLUDED
namespace strus {
class PatternLexerInterface;
class PatternMatcherInterface;
class ErrorBufferInterface;
struct PatternLexerConstructor
{
	typedef PatternLexerInterface* (*Create)( ErrorBufferInterface* errorhnd);
	const char* name;
	Create create;
};
struct PatternMatcherConstructor
{
	typedef PatternMatcherInterface* (*Create)( ErrorBufferInterface* errorhnd);
	const char* name;
	Create create;
};
struct AnalyzerModule
{
	AnalyzerModule( const PatternLexerConstructor& lexer_, const PatternMatcherConstructor& matcher_, const char* version_3rdparty, const char* license_3rdparty)
		:patternLexerConstructor(&lexer_),patternMatcherConstructor(&matcher_),version_3rdparty_string(version_3rdparty),license_3rdparty_string(license_3rdparty){}
	const PatternLexerConstructor* patternLexerConstructor;
	const PatternMatcherConstructor* patternMatcherConstructor;
	const char* version_3rdparty_string;
	const char* license_3rdparty_string;
};
}
#endif
