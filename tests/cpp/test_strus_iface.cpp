// Test of the strus C++ interfaces of the module (struspattern_b200/csrc/strus_iface.cpp), written the way the
// reference's own test programs drive them:
//   lexer   tests/charRegexMatch/src/testCharRegexMatch.cpp:212-262  (createPatternLexer_std, defineLexem, defineSymbol,
//           compile, createContext, match; expected lexems compared element by element in order)
//   matcher tests/simpleTokenPatternMatch/src/testSimpleTokenPatternMatch.cpp:221-305 (pushTerm, attachVariable,
//           pushExpression, definePattern, compile, putInput per token, fetchResults; compared as a set)
// The test vectors come from a line oriented case file written by tests/test_strus_iface.py out of tests/golden/*.json.
#include "strus/lib/pattern.hpp"
#include "strus/lib/error.hpp"
#include "strus/errorBufferInterface.hpp"
#include "strus/patternLexerInterface.hpp"
#include "strus/patternLexerInstanceInterface.hpp"
#include "strus/patternLexerContextInterface.hpp"
#include "strus/patternMatcherInterface.hpp"
#include "strus/patternMatcherInstanceInterface.hpp"
#include "strus/patternMatcherContextInterface.hpp"
#include "strus/analyzerModule.hpp"
#include "strus/base/local_ptr.hpp"
#include <dlfcn.h>
#include <fstream>
#include <sstream>
#include <iostream>
#include <set>
#include <string>
#include <vector>
#include <stdexcept>
#include <cstdlib>

static std::string unhex( const std::string& h)
{
	std::string rt;
	for (std::size_t i=0; i+1<h.size(); i+=2) rt.push_back( (char)std::strtoul( h.substr( i, 2).c_str(), 0, 16));
	return rt;
}

struct ExpectedLexem {unsigned int id, ordpos, origpos, origsize;};

static void checkError( strus::ErrorBufferInterface* errorhnd, const char* where)
{
	if (errorhnd->hasError()) throw std::runtime_error( std::string( where) + ": " + errorhnd->fetchError());
}

static int runLexerCase( std::istream& in, strus::ErrorBufferInterface* errorhnd)
{
	strus::local_ptr<strus::PatternLexerInterface> pt( strus::createPatternLexer_std( errorhnd));
	if (!pt.get()) throw std::runtime_error( "failed to create regular expression term matcher");
	strus::local_ptr<strus::PatternLexerInstanceInterface> ptinst( pt->createInstance());
	if (!ptinst.get()) throw std::runtime_error( "failed to create regular expression term matcher instance");
	std::string text;
	std::vector<ExpectedLexem> expected;
	std::string line;
	while (std::getline( in, line) && line != "END")
	{
		std::istringstream ls( line);
		char tag; ls >> tag;
		if (tag == 'O') {std::string name; ls >> name; ptinst->defineOption( name, 0);}
		else if (tag == 'L')
		{
			unsigned int id, resultIndex, level, posbind; std::string hex;
			ls >> id >> resultIndex >> level >> posbind >> hex;
			ptinst->defineLexem( id, unhex( hex), resultIndex, level, (strus::analyzer::PositionBind)posbind);
		}
		else if (tag == 'S') {unsigned int id, lexemid; std::string hex; ls >> id >> lexemid >> hex; ptinst->defineSymbol( id, lexemid, unhex( hex));}
		else if (tag == 'T') {std::string hex; ls >> hex; text = unhex( hex);}
		else if (tag == 'E') {ExpectedLexem e; ls >> e.id >> e.ordpos >> e.origpos >> e.origsize; expected.push_back( e);}
	}
	checkError( errorhnd, "define lexems");
	// state errors of the reference (src/patternLexer.cpp:1124-1127): no context before compile
	if (ptinst->createContext() != 0 || !errorhnd->hasError()) throw std::runtime_error( "createContext before compile must fail");
	(void)errorhnd->fetchError();
	if (!ptinst->compile()) {checkError( errorhnd, "compile"); throw std::runtime_error( "compile failed");}
	ptinst->defineLexem( 99, "x", 0, 0, strus::analyzer::BindContent);
	if (!errorhnd->hasError()) throw std::runtime_error( "defineLexem after compile must fail");
	(void)errorhnd->fetchError();
	strus::local_ptr<strus::PatternLexerContextInterface> mt( ptinst->createContext());
	checkError( errorhnd, "create context");
	std::vector<strus::analyzer::PatternLexem> result = mt->match( text.c_str(), text.size());
	checkError( errorhnd, "match");
	if (result.size() != expected.size()) throw std::runtime_error( "number of lexems differs from the expected");
	for (std::size_t i=0; i<result.size(); ++i)
	{
		const strus::analyzer::PatternLexem& r = result[i]; const ExpectedLexem& e = expected[i];
		if (r.id() != e.id || (unsigned int)r.ordpos() != e.ordpos || (unsigned int)r.origpos().ofs() != e.origpos || (unsigned int)r.origsize() != e.origsize || r.origpos().seg() != 0)
		{
			std::ostringstream msg; msg << "lexem " << i << " differs: got (" << r.id() << "," << r.ordpos() << "," << r.origpos().ofs() << "," << r.origsize() << ")";
			throw std::runtime_error( msg.str());
		}
	}
	// a second document on the same context and an empty one
	std::vector<strus::analyzer::PatternLexem> again = mt->match( text.c_str(), text.size());
	if (again.size() != result.size()) throw std::runtime_error( "context reuse changes the result");
	if (!mt->match( "", 0).empty()) throw std::runtime_error( "empty segment must give no lexems");
	checkError( errorhnd, "match again");
	return (int)result.size();
}

static int runMatcherCase( std::istream& in, strus::ErrorBufferInterface* errorhnd)
{
	strus::local_ptr<strus::PatternMatcherInterface> pt( strus::createPatternMatcher_std( errorhnd));
	if (!pt.get()) throw std::runtime_error( "failed to create pattern matcher");
	strus::local_ptr<strus::PatternMatcherInstanceInterface> ptinst( pt->createInstance());
	if (!ptinst.get()) throw std::runtime_error( "failed to create pattern matcher instance");
	std::vector<strus::analyzer::PatternLexem> doc;
	std::set<std::string> expected;
	std::string line;
	while (std::getline( in, line) && line != "END")
	{
		std::istringstream ls( line);
		char tag; ls >> tag;
		if (tag == 't') {unsigned int termid; std::string var; ls >> termid >> var; ptinst->pushTerm( termid); if (var != "-") ptinst->attachVariable( var);}
		else if (tag == 'x') {int op; unsigned int argc, range, card; ls >> op >> argc >> range >> card; ptinst->pushExpression( (strus::PatternMatcherInstanceInterface::JoinOperation)op, argc, range, card);}
		else if (tag == 'p') {std::string name; int visible; ls >> name >> visible; ptinst->definePattern( name, "", visible != 0);}
		else if (tag == 'd') {unsigned int id, ordpos, origpos, origsize; ls >> id >> ordpos >> origpos >> origsize; doc.push_back( strus::analyzer::PatternLexem( id, ordpos, strus::analyzer::Position( 0, origpos), origsize));}
		else if (tag == 'e') {std::string name; unsigned int pos; ls >> name >> pos; std::ostringstream key; key << name << "@" << pos; expected.insert( key.str());}
	}
	checkError( errorhnd, "define patterns");
	if (!ptinst->compile()) {checkError( errorhnd, "compile"); throw std::runtime_error( "compile failed");}
	strus::local_ptr<strus::PatternMatcherContextInterface> mt( ptinst->createContext());
	checkError( errorhnd, "create context");
	for (const strus::analyzer::PatternLexem& tok: doc) mt->putInput( tok);
	checkError( errorhnd, "put input");
	std::vector<strus::analyzer::PatternMatcherResult> results = mt->fetchResults();
	checkError( errorhnd, "fetch results");
	std::set<std::string> got;
	for (const strus::analyzer::PatternMatcherResult& r: results)
	{
		std::ostringstream key; key << r.name() << "@" << r.ordpos(); got.insert( key.str());
		// items come out last bound first (doc/webpage/introduction_struspattern.htm:233-236)
		if (r.items().size() != 2 || std::string( r.items()[0].name()) != "A2" || std::string( r.items()[1].name()) != "A1") throw std::runtime_error( "unexpected result items");
	}
	if (got != expected) throw std::runtime_error( "result set differs from the expected");
	strus::analyzer::PatternMatcherStatistics stats = mt->getStatistics();
	if (stats.items().size() != 4 || stats.items()[0].value() <= 0) throw std::runtime_error( "unexpected statistics");
	// descending position is an error of the call, the token is not part of the document
	mt->putInput( strus::analyzer::PatternLexem( 1, 1, strus::analyzer::Position( 0, 0), 1));
	if (!errorhnd->hasError()) throw std::runtime_error( "putInput with a descending position must fail");
	(void)errorhnd->fetchError();
	mt->reset();
	if (!mt->fetchResults().empty()) throw std::runtime_error( "reset must clear the context");
	checkError( errorhnd, "after reset");
	return (int)results.size();
}

// Result format strings through the strus classes: the money amount example of the article
// (doc/webpage/introduction_struspattern.htm:148-161): Currency = CURRENCY_CHF ["CHF"]; Currency = CURRENCY_EUR ["EUR"];
// MoneyAmount = sequence_imm( value=AMOUNT, currency=Currency) ["{currency} {value}"].
static int runFormatCase( strus::ErrorBufferInterface* errorhnd)
{
	enum {AMOUNT=1, CHF=2, EUR=3, WORD=4};
	strus::local_ptr<strus::PatternMatcherInterface> pt( strus::createPatternMatcher_std( errorhnd));
	strus::local_ptr<strus::PatternMatcherInstanceInterface> ptinst( pt->createInstance());
	ptinst->pushTerm( CHF); ptinst->definePattern( "Currency", "CHF", true);
	ptinst->pushTerm( EUR); ptinst->definePattern( "Currency", "EUR", true);
	ptinst->pushTerm( AMOUNT); ptinst->attachVariable( "value");
	ptinst->pushPattern( "Currency"); ptinst->attachVariable( "currency");
	ptinst->pushExpression( strus::PatternMatcherInstanceInterface::OpSequenceImm, 2, 1, 0);
	ptinst->definePattern( "MoneyAmount", "{currency} {value}", true);
	ptinst->pushTerm( AMOUNT); ptinst->attachVariable( "value");
	ptinst->pushPattern( "Currency"); ptinst->attachVariable( "currency");
	ptinst->pushExpression( strus::PatternMatcherInstanceInterface::OpSequenceImm, 2, 1, 0);
	ptinst->definePattern( "PlainAmount", "", true);
	checkError( errorhnd, "define patterns with format strings");
	ptinst->pushTerm( AMOUNT);
	ptinst->definePattern( "Broken", "{nosuchvariable}", true);
	if (!errorhnd->hasError()) throw std::runtime_error( "a format string with an unknown variable must be refused");
	(void)errorhnd->fetchError();
	if (!ptinst->compile()) {checkError( errorhnd, "compile"); throw std::runtime_error( "compile failed");}
	strus::local_ptr<strus::PatternMatcherContextInterface> mt( ptinst->createContext());
	// "pay 3'645 Eur or 12 sFr. now"
	static const unsigned int toks[][3] = {{WORD,0,3},{AMOUNT,4,5},{EUR,10,3},{WORD,14,2},{AMOUNT,17,2},{CHF,20,4},{WORD,25,3}};
	for (unsigned int i=0; i<7; ++i) mt->putInput( strus::analyzer::PatternLexem( toks[i][0], i+1, strus::analyzer::Position( 0, toks[i][1]), toks[i][2]));
	std::vector<strus::analyzer::PatternMatcherResult> results = mt->fetchResults();
	checkError( errorhnd, "fetch results with format strings");
	std::set<std::string> got;
	for (const strus::analyzer::PatternMatcherResult& r: results)
	{
		std::string key = std::string( r.name()) + "=" + (r.value() ? r.value() : "(null)");
		if (std::string( r.name()) == "PlainAmount")
		{
			// no format string on the result: items visible, the item that refers to Currency carries Currency's value
			if (r.value() || r.items().size() != 2) throw std::runtime_error( "unexpected items of a result without format string");
			for (const strus::analyzer::PatternMatcherResultItem& it: r.items())
			{
				key += std::string( " ") + it.name() + "=" + (it.value() ? it.value() : "(null)");
			}
		}
		else if (!r.items().empty()) throw std::runtime_error( "a result with a format string must not have items");
		got.insert( key);
	}
	std::set<std::string> expected;
	expected.insert( "Currency=EUR"); expected.insert( "Currency=CHF");
	expected.insert( "MoneyAmount=EUR \x01" "0,4,0,9\x02"); expected.insert( "MoneyAmount=CHF \x01" "0,17,0,19\x02");
	expected.insert( "PlainAmount=(null) currency=EUR value=(null)"); expected.insert( "PlainAmount=(null) currency=CHF value=(null)");
	if (got != expected)
	{
		std::string msg( "values differ from the expected:");
		for (const std::string& g: got) msg += " [" + g + "]";
		throw std::runtime_error( msg);
	}
	return (int)results.size();
}

int main( int argc, const char** argv)
{
	if (argc < 2) {std::cerr << "usage: test_strus_iface <casefile> [<module.so>]" << std::endl; return 2;}
	strus::local_ptr<strus::ErrorBufferInterface> errorhnd( strus::createErrorBuffer_standard( 0, 1, 0));
	try
	{
		std::ifstream in( argv[1]);
		std::string line;
		int ncases = 0;
		while (std::getline( in, line))
		{
			if (line == "LEXER") {std::cout << "lexer case: " << runLexerCase( in, errorhnd.get()) << " lexems OK" << std::endl; ++ncases;}
			else if (line == "FORMATS") {std::cout << "format case: " << runFormatCase( errorhnd.get()) << " results OK" << std::endl; ++ncases;}
			else if (line == "MATCHER") {std::cout << "matcher case: " << runMatcherCase( in, errorhnd.get()) << " results OK" << std::endl; ++ncases;}
		}
		if (argc > 2)
		{
			// the module entry point as strusModule's loader reads it (src/modstrus_analyzer_pattern.cpp:59-61)
			void* hnd = dlopen( argv[2], RTLD_NOW);
			if (!hnd) throw std::runtime_error( std::string( "dlopen failed: ") + dlerror());
			strus::AnalyzerModule* mod = (strus::AnalyzerModule*)dlsym( hnd, "entryPoint");
			if (!mod || std::string( mod->patternLexerConstructor->name) != "std" || std::string( mod->patternMatcherConstructor->name) != "std") throw std::runtime_error( "bad module entry point");
			strus::local_ptr<strus::PatternMatcherInterface> pm( mod->patternMatcherConstructor->create( errorhnd.get()));
			if (!pm.get() || std::string( pm->name()) != "std") throw std::runtime_error( "module constructor failed");
			std::cout << "module entry point OK" << std::endl;
		}
		std::cout << "done " << ncases << " cases" << std::endl;
		return ncases ? 0 : 1;
	}
	catch (const std::exception& e)
	{
		std::cerr << "ERROR " << e.what() << std::endl;
		return 1;
	}
}
