// C ABI of the level-1 lexer (include/strus_pattern_b200.h): instance = compiled tables in HBM,
// context = scratch and output buffers of one in-flight batch. There is no CPU path: every scan launches the
// sm_100a kernels of lexer_kernels.cuh and fails with SP_ERR_CUDA if no device is usable.
#include "strus_pattern_b200.h"
#include "lexer_compile.hpp"
#include "lexer_kernels.cuh"
#include "lexer_islands.cuh"
#include "device_util.cuh"
#include <vector>
#include <string>
#include <cstring>
#include <cstdlib>

using namespace sp;

struct sp_lexer
{
	LexerProgram prog;
	std::string lasterror;
	bool uploaded;
	int device;
	SpLexerTables dev;		// device view of the tables
	std::vector<void*> allocations;

	sp_lexer() :uploaded(false),device(-1) {std::memset( &dev, 0, sizeof(dev));}
	~sp_lexer() {for (void* p: allocations) cudaFree( p);}

	template <typename T> const T* put( const std::vector<T>& v)
	{
		void* p = 0;
		std::size_t bytes = (v.size() ? v.size() : 1) * sizeof(T);
		SP_CUDA( cudaMalloc( &p, bytes + 16));
		allocations.push_back( p);
		if (!v.empty()) SP_CUDA( cudaMemcpy( p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice));
		return static_cast<const T*>( p);
	}
	void upload()
	{
		SP_CUDA( cudaGetDevice( &device));
		const LexerTablesHost& H = prog.tables();
		dev = H.view;
		dev.fwd_next = put( H.fwd_next); dev.fwd_class = put( H.fwd_class);
		dev.fwd_acc_off = put( H.fwd_acc_off); dev.fwd_acc_span = put( H.fwd_acc_span); dev.fwd_acc_pat = put( H.fwd_acc_pat);
		dev.rev_next = put( H.rev_next); dev.rev_classes = put( H.rev_classes);
		dev.cap_next = put( H.cap_next); dev.cap_classes = put( H.cap_classes); dev.cap = put( H.cap);
		dev.pat = put( H.pat); dev.sym_slots = put( H.sym_slots); dev.sym_pool = put( H.sym_pool);
		uploaded = true;
	}
};

static int lexerFail( std::string& lasterror, const std::exception& e)
{
	lasterror = e.what();
	if (dynamic_cast<const UnsupportedError*>( &e)) {lasterror = "unsupported: " + lasterror; return SP_ERR_UNSUPPORTED;}
	if (dynamic_cast<const SyntaxError*>( &e)) return SP_ERR_SYNTAX;
	if (dynamic_cast<const CudaError*>( &e)) return SP_ERR_CUDA;
	if (dynamic_cast<const std::bad_alloc*>( &e)) return SP_ERR_NOMEM;
	return SP_ERR_ARGUMENT;
}
#define SP_LEXER_TRY( obj, ...) try {__VA_ARGS__; return SP_OK;} catch (const std::exception& e) {return lexerFail( (obj)->lasterror, e);}

extern "C" {

sp_lexer* sp_lexer_create(void) {try {return new sp_lexer();} catch (...) {return 0;}}
void sp_lexer_destroy( sp_lexer* lx) {delete lx;}
const char* sp_lexer_last_error( const sp_lexer* lx) {return lx->lasterror.c_str();}

int sp_lexer_define_option( sp_lexer* lx, const char* name, double)
	{SP_LEXER_TRY( lx, lx->prog.defineOption( name))}
int sp_lexer_define_lexem( sp_lexer* lx, uint32_t id, const char* expression, size_t expression_len, uint32_t result_index, uint32_t level, uint32_t posbind)
	{SP_LEXER_TRY( lx, lx->prog.defineLexem( id, std::string( expression, expression_len), result_index, level, posbind))}
int sp_lexer_define_symbol( sp_lexer* lx, uint32_t symbolid, uint32_t lexemid, const char* name, size_t name_len)
	{SP_LEXER_TRY( lx, lx->prog.defineSymbol( symbolid, lexemid, std::string( name, name_len)))}
uint32_t sp_lexer_get_symbol( const sp_lexer* lx, uint32_t lexemid, const char* name, size_t name_len)
{
	try {return lx->prog.getSymbol( lexemid, std::string( name, name_len));} catch (...) {return 0;}
}
int sp_lexer_compile( sp_lexer* lx)
{
	SP_LEXER_TRY( lx, {
		if (lx->prog.compiled()) throw std::runtime_error( "lexer compiled twice");
		lx->prog.compile();
		lx->upload();
	})
}
int64_t sp_lexer_serialize( const sp_lexer* lx, void* image, uint64_t capacity)
{
	try
	{
		std::vector<uint8_t> bytes;
		lx->prog.serialize( bytes);
		if (image && capacity >= bytes.size()) std::memcpy( image, bytes.data(), bytes.size());
		return (int64_t)bytes.size();
	}
	catch (const std::exception& e) {return lexerFail( const_cast<sp_lexer*>( lx)->lasterror, e);}
}
int sp_lexer_deserialize( sp_lexer* lx, const void* image, uint64_t size)
{
	SP_LEXER_TRY( lx, {
		lx->prog.deserialize( static_cast<const uint8_t*>( image), (std::size_t)size);
		lx->upload();
	})
}
int64_t sp_lexer_table_info( const sp_lexer* lx, const char* key)
{
	std::string k( key);
	const LexerTablesHost& H = lx->prog.tables();
	if (k == "fwd_states") return lx->prog.forwardStates();
	if (k == "fwd_stride") return H.view.fwd_stride;
	if (k == "fwd_table_bytes") return (int64_t)(H.fwd_next.size() * sizeof(uint16_t));
	if (k == "fwd_accept_states") return (int64_t)H.view.fwd_states - (int64_t)H.view.fwd_acc_base;
	if (k == "rev_states") return lx->prog.reverseStates();
	if (k == "rev_table_bytes") return (int64_t)(H.rev_next.size() * sizeof(uint32_t));
	if (k == "definitions") return (int64_t)H.pat.size();
	if (k == "symbols") return (int64_t)H.view.has_symbols;
	if (k == "compiled") return lx->prog.compiled() ? 1 : 0;
	return -1;
}

} // extern "C"

#include "lexer_context.inl"
