/*
 * strus_pattern_b200.h -- C ABI of the B200-native strusPattern hot path.
 *
 * This is the drop-in boundary: plain C, plain pointers and sizes, int return codes
 * (0 = ok, <0 = error, message via sp_*_last_error). The C++ strus interface classes
 * (PatternLexer*Interface / PatternMatcher*Interface, struspattern_b200/csrc/strus_iface.cpp)
 * and the python harness (struspattern_b200/capi.py) are thin layers over these calls.
 * Every entry point names the reference interface (file:line under the strusPattern
 * source tree) it replaces.
 *
 * There is no CPU fallback behind this ABI: every scan/run call launches sm_100a kernels
 * and fails with SP_ERR_CUDA when no device is usable.
 */
#ifndef STRUS_PATTERN_B200_H_INCLUDED
#define STRUS_PATTERN_B200_H_INCLUDED
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SP_OK            0
#define SP_ERR_ARGUMENT -1   /* bad argument / call in wrong state (reference: runtime_error -> report) */
#define SP_ERR_SYNTAX   -2   /* regex not compilable (reference: hyperscanErrorCode(HS_COMPILER_ERROR)) */
#define SP_ERR_NOMEM    -3
#define SP_ERR_CUDA     -4   /* no device, launch failure */
#define SP_ERR_RUNTIME  -5   /* per-document runtime error (e.g. token >= 65535 bytes, arena overflow) */
#define SP_ERR_UNSUPPORTED -6

/* analyzer::PositionBind (strus/analyzer/positionBind.hpp) */
#define SP_BIND_CONTENT     0
#define SP_BIND_SUCCESSOR   1
#define SP_BIND_PREDECESSOR 2
#define SP_BIND_UNIQUE      3

/* PatternMatcherInstanceInterface::JoinOperation */
#define SP_OP_SEQUENCE        0
#define SP_OP_SEQUENCE_IMM    1
#define SP_OP_SEQUENCE_STRUCT 2
#define SP_OP_WITHIN          3
#define SP_OP_WITHIN_STRUCT   4
#define SP_OP_ANY             5
#define SP_OP_AND             6

/* One lexem = analyzer::PatternLexem(id, ordpos, Position(seg, origpos), origsize).
 * The lexer always emits segment 0 (src/patternLexer.cpp:908); callers that feed several
 * segments to the matcher pass the segment numbers in a parallel array. 16 bytes, so a
 * warp reads/writes lexem records with 16-byte vector accesses. */
typedef struct sp_lexem
{
	uint32_t id;
	uint32_t ordpos;
	uint32_t origpos;
	uint32_t origsize;
} sp_lexem;

/* One result = analyzer::PatternMatcherResult (src/patternMatcher.cpp:248-269). */
typedef struct sp_result
{
	uint32_t name;        /* pattern handle: sp_matcher_pattern_name(m, name) */
	uint32_t format;      /* format handle (0 = none) */
	uint32_t ordpos;
	uint32_t ordend;
	uint32_t origseg;
	uint32_t origpos;
	uint32_t endseg;
	uint32_t endpos;
	uint32_t item_begin;  /* index of the first item, relative to the document's item range */
	uint32_t item_count;
} sp_result;

/* One result item = analyzer::PatternMatcherResultItem (src/patternMatcher.cpp:164-190).
 * `format`: low half = format handle of the sub expression the item refers to (0 = none: the value of the item is the
 * piece of the source it spans), high half = number of NESTED records that follow the item: the items of that sub
 * expression, which the reference hands to the formatter instead of the caller (:172-181). A consumer that lists the
 * items of a result skips them (i += 1 + SP_ITEM_NESTED(item)); sp_matcher_item_value maps them into the value string. */
#define SP_ITEM_FORMAT( it)  ((it).format & 0xFFFFu)
#define SP_ITEM_NESTED( it)  ((it).format >> 16)
typedef struct sp_item
{
	uint32_t variable;    /* variable handle: sp_matcher_variable_name(m, variable) */
	uint32_t format;      /* SP_ITEM_FORMAT | SP_ITEM_NESTED << 16 */
	uint32_t ordpos;
	uint32_t ordend;
	uint32_t origseg;
	uint32_t origpos;
	uint32_t endseg;
	uint32_t endpos;
} sp_item;

/* Per-document statistics = PatternMatcherContext::getStatistics (src/patternMatcher.cpp:303-318). */
typedef struct sp_matcher_stats
{
	uint64_t nof_programs_installed;
	uint64_t nof_altkey_programs_installed;
	uint64_t nof_signals_fired;
	uint64_t nof_open_patterns;   /* sum over events of live triggers (reference: double accumulator) */
	uint64_t nof_events;          /* number of putInput calls */
} sp_matcher_stats;

/* ---------------------------------------------------------------------------------------
 * Level 1: regex lexer.  Replaces PatternLexerInstance / PatternLexerContext
 * (src/patternLexer.cpp:961-1151, 681-959) and the Hyperscan database behind them.
 * ------------------------------------------------------------------------------------- */
typedef struct sp_lexer sp_lexer;          /* = PatternLexerInstanceInterface */
typedef struct sp_lexer_ctx sp_lexer_ctx;  /* = PatternLexerContextInterface, batched */

/* PatternLexer::createInstance, src/patternLexer.cpp:1165 */
sp_lexer* sp_lexer_create(void);
void sp_lexer_destroy( sp_lexer* lx);
const char* sp_lexer_last_error( const sp_lexer* lx);
/* PatternLexerInstance::defineOption, src/patternLexer.cpp:1031 (CASELESS DOTALL MULTILINE ALLOWEMPTY UCP BYTECHAR) */
int sp_lexer_define_option( sp_lexer* lx, const char* name, double value);
/* PatternLexerInstance::defineLexem, src/patternLexer.cpp:990 -> PatternTable::definePattern :245 */
int sp_lexer_define_lexem( sp_lexer* lx, uint32_t id, const char* expression, size_t expression_len,
			   uint32_t result_index, uint32_t level, uint32_t posbind);
/* PatternLexerInstance::defineSymbol, src/patternLexer.cpp:1008 -> PatternTable::defineSymbol :264 */
int sp_lexer_define_symbol( sp_lexer* lx, uint32_t symbolid, uint32_t lexemid, const char* name, size_t name_len);
/* PatternLexerInstance::getSymbol, src/patternLexer.cpp:1020 (0 = not defined) */
uint32_t sp_lexer_get_symbol( const sp_lexer* lx, uint32_t lexemid, const char* name, size_t name_len);
/* PatternLexerInstance::compile, src/patternLexer.cpp:1068 (replaces hs_compile_ext_multi :1083);
 * builds the transition tables on the host and uploads them to the current CUDA device. */
int sp_lexer_compile( sp_lexer* lx);
/* Introspection of the compiled tables (DESIGN.md "lexer tables"); values by key name, -1 if unknown. */
int64_t sp_lexer_table_info( const sp_lexer* lx, const char* key);
/* Compiled lexer <-> flat byte image (no counterpart in the reference, which calls
 * hs_compile_ext_multi on every start, src/patternLexer.cpp:1083, and never hs_serialize_database).
 * sp_lexer_serialize returns the image size (>= 0) and fills `image` if `capacity` is large enough (call with
 * image = NULL to ask for the size), or a negative SP_ERR_*. sp_lexer_deserialize loads an image into a lexer that
 * has no definitions yet and uploads the tables, like sp_lexer_compile does; images of another table layout or
 * damaged images are rejected with SP_ERR_ARGUMENT. */
int64_t sp_lexer_serialize( const sp_lexer* lx, void* image, uint64_t capacity);
int sp_lexer_deserialize( sp_lexer* lx, const void* image, uint64_t size);

/* PatternLexerInstance::createContext, src/patternLexer.cpp:1120: scratch + output buffers of one
 * in-flight batch. cuda_stream is a cudaStream_t (NULL = default stream). */
sp_lexer_ctx* sp_lexer_ctx_create( const sp_lexer* lx, void* cuda_stream);
void sp_lexer_ctx_destroy( sp_lexer_ctx* ctx);
const char* sp_lexer_ctx_last_error( const sp_lexer_ctx* ctx);

/* PatternLexerContext::match (src/patternLexer.cpp:858) for a batch of documents laid out
 * back to back: document d is text[doc_offsets[d] .. doc_offsets[d+1]). doc_offsets is a host
 * array of ndocs+1 entries. text_on_device != 0: text is a device pointer (inputs resident in
 * HBM); else a host pointer that is copied to the device inside the call.
 * Asynchronous on the context's stream; results stay on the device until fetched. */
int sp_lexer_scan( sp_lexer_ctx* ctx, const uint8_t* text, int text_on_device,
		   const uint64_t* doc_offsets, uint32_t ndocs);
/* Blocks until the scan finished; returns total number of lexems (or <0 on error). */
int64_t sp_lexer_result_count( sp_lexer_ctx* ctx);
/* Device views of the last scan's output (valid until the next scan on this context). */
const sp_lexem* sp_lexer_result_lexems_device( sp_lexer_ctx* ctx);
const uint64_t* sp_lexer_result_offsets_device( sp_lexer_ctx* ctx);   /* ndocs+1 */
const uint32_t* sp_lexer_result_status_device( sp_lexer_ctx* ctx);    /* ndocs, 0 = ok else SP_ERR_* (positive) */
/* Copy results to host arrays: lexems[count], offsets[ndocs+1], status[ndocs] (each may be NULL). */
int sp_lexer_fetch( sp_lexer_ctx* ctx, sp_lexem* lexems, uint64_t* offsets, uint32_t* status);
/* The same for the documents [first_doc, first_doc + ndocs) of the last scan: offsets[ndocs+1] are the stored (absolute)
 * offsets, lexems receives offsets[ndocs] - offsets[0] records (capacity = room in `lexems`). Pages through a large batch. */
int sp_lexer_fetch_range( sp_lexer_ctx* ctx, uint32_t first_doc, uint32_t ndocs, sp_lexem* lexems, uint64_t capacity,
			  uint64_t* offsets, uint32_t* status);
/* Counters of the last scan (for roofline accounting): "raw_events", "lexems", "chunks",
 * "fixup_chunks", "text_bytes". */
int64_t sp_lexer_ctx_counter( sp_lexer_ctx* ctx, const char* key);

/* ---------------------------------------------------------------------------------------
 * Level 2: token pattern matcher.  Replaces PatternMatcherInstance / PatternMatcherContext
 * (src/patternMatcher.cpp:345-704, 107-341) and ProgramTable / StateMachine
 * (src/ruleMatcherAutomaton.cpp:259-586, 589-1334).
 * ------------------------------------------------------------------------------------- */
typedef struct sp_matcher sp_matcher;          /* = PatternMatcherInstanceInterface */
typedef struct sp_matcher_ctx sp_matcher_ctx;  /* = PatternMatcherContextInterface, batched */

sp_matcher* sp_matcher_create(void);                                          /* src/patternMatcher.cpp:718 */
void sp_matcher_destroy( sp_matcher* m);
const char* sp_matcher_last_error( const sp_matcher* m);
int sp_matcher_define_option( sp_matcher* m, const char* name, double value);  /* :614 */
int sp_matcher_define_term_frequency( sp_matcher* m, uint32_t termid, double df); /* :361 */
int sp_matcher_push_term( sp_matcher* m, uint32_t termid);                    /* :366 */
int sp_matcher_push_expression( sp_matcher* m, int joinop, size_t argc, uint32_t range, uint32_t cardinality); /* :377 */
int sp_matcher_push_pattern( sp_matcher* m, const char* name);                /* :510 */
int sp_matcher_attach_variable( sp_matcher* m, const char* name);             /* :522 */
int sp_matcher_define_pattern( sp_matcher* m, const char* name, const char* formatstring, int visible); /* :545 */
/* PatternMatcherInstance::compile (:646): ProgramTable::optimize (src/ruleMatcherAutomaton.cpp:512). */
int sp_matcher_compile( sp_matcher* m);
/* Flatten the rule graph into device tables and upload them to the current device. Called
 * implicitly by sp_matcher_ctx_create; explicit call allowed (idempotent until the next define). */
int sp_matcher_upload( sp_matcher* m);
const char* sp_matcher_pattern_name( const sp_matcher* m, uint32_t handle);    /* patternMap.key, :250 */
const char* sp_matcher_variable_name( const sp_matcher* m, uint32_t handle);   /* variableMap.key, :170 */
/* format string of a format handle (sp_result.format / sp_item.format), resultFormatHandles, :560-566 */
const char* sp_matcher_format_string( const sp_matcher* m, uint32_t handle);
/* Value strings = PatternMatcherResult::value / PatternMatcherResultItem::value (src/patternMatcher.cpp:253-262, :172-181;
 * the formatter is strusAnalyzer's PatternResultFormat, restated in csrc/result_format.hpp). `result` and `doc_items`
 * are records of one document as sp_matcher_fetch delivers them (doc_items = items + item_offsets[d]); `item` points
 * into such an array (its nested records follow it). Returns the length of the value (the first `capacity` bytes are
 * written, no terminator: call again with a larger buffer if the length exceeds it), SP_NO_VALUE if the record has no
 * format string (the reference's value is NULL then), or a negative SP_ERR_*. A result WITH a format string has no
 * visible items (the reference passes them to the formatter): all its records are material of the value.
 * The matcher does not see the source text: a span of the source is written into the value as a reference
 * "\x01" seg "," pos "," endseg "," endpos "\x02" (decimal numbers); sp_resolve_value replaces the references of
 * segment 0 by the bytes of `source` (same return convention). */
#define SP_NO_VALUE -100
int64_t sp_matcher_result_value( const sp_matcher* m, const sp_result* result, const sp_item* doc_items, char* buf, uint64_t capacity);
int64_t sp_matcher_item_value( const sp_matcher* m, const sp_item* item, char* buf, uint64_t capacity);
int64_t sp_resolve_value( const char* value, uint64_t value_len, const uint8_t* source, uint64_t source_len, char* buf, uint64_t capacity);
int64_t sp_matcher_table_info( const sp_matcher* m, const char* key);

sp_matcher_ctx* sp_matcher_ctx_create( sp_matcher* m, void* cuda_stream);      /* :586 */
void sp_matcher_ctx_destroy( sp_matcher_ctx* ctx);
const char* sp_matcher_ctx_last_error( const sp_matcher_ctx* ctx);
/* putInput for every token of every document + fetchResults (:131, :271). Document d owns
 * tokens[tok_offsets[d] .. tok_offsets[d+1]); tokens must be in non-decreasing ordpos order per
 * document (violations give status SP_ERR_RUNTIME for that document, like the reference's
 * "term events not fed in ascending order"). origsegs may be NULL (all zero).
 * on_device != 0: tokens/origsegs/tok_offsets are device pointers, else host pointers. */
int sp_matcher_run( sp_matcher_ctx* ctx, const sp_lexem* tokens, const uint32_t* origsegs,
		    const uint64_t* tok_offsets, uint32_t ndocs, int on_device);
/* Blocks; total results / items of the last run. */
int64_t sp_matcher_result_count( sp_matcher_ctx* ctx);
int64_t sp_matcher_item_count( sp_matcher_ctx* ctx);
const sp_result* sp_matcher_results_device( sp_matcher_ctx* ctx);
const sp_item* sp_matcher_items_device( sp_matcher_ctx* ctx);
/* Copy to host: results[nres], result_offsets[ndocs+1], items[nitems], item_offsets[ndocs+1],
 * stats[ndocs], status[ndocs]; each may be NULL. */
int sp_matcher_fetch( sp_matcher_ctx* ctx, sp_result* results, uint64_t* result_offsets,
		      sp_item* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status);

/* The same for the documents [first_doc, first_doc + ndocs) of the last run (fetchResults of a slice of the batch):
 * result_offsets / item_offsets (required) receive the stored (absolute) offsets, results / items the records between
 * offsets[0] and offsets[ndocs]. */
int sp_matcher_fetch_range( sp_matcher_ctx* ctx, uint32_t first_doc, uint32_t ndocs, sp_result* results, uint64_t result_capacity,
			    uint64_t* result_offsets, sp_item* items, uint64_t item_capacity, uint64_t* item_offsets,
			    sp_matcher_stats* stats, uint32_t* status);

/* Compact wire format for host-bound output (the result and item records are 37 bytes per byte of text on the C3
 * workload: the device-to-host link bounds the end-to-end rate). The records of the last run are packed on the device
 * (16 bytes per result, 12 per item instead of 40 and 32), copied, and unpacked on the host by sp_unpack_results into
 * the same sp_result / sp_item records sp_matcher_fetch delivers. Packable are runs whose records lie in segment 0 (what
 * the lexer emits), with ordinal spans < 256 (results) / < 16 (items), source spans < 65536 bytes, at most 255 items per
 * result, pattern handles < 2^18, format handles < 2^14 on results and none on items, variable handles < 4096;
 * anything else makes sp_matcher_fetch_packed fail with SP_ERR_UNSUPPORTED and the caller uses sp_matcher_fetch.
 * No counterpart in the reference (its results are C++ objects in host memory). */
typedef struct sp_result_packed
{
	uint32_t name_format;  /* name | format << 18 */
	uint32_t ordpos;
	uint32_t origpos;
	uint32_t lens;         /* (ordend - ordpos) | item_count << 8 | (endpos - origpos) << 16 */
} sp_result_packed;
typedef struct sp_item_packed
{
	uint32_t var_lens;     /* variable | (ordend - ordpos) << 12 | (endpos - origpos) << 16 */
	uint32_t ordpos;
	uint32_t origpos;
} sp_item_packed;
/* Like sp_matcher_fetch with packed records: results[nres], result_offsets[ndocs+1], items[nitems],
 * item_offsets[ndocs+1], stats[ndocs], status[ndocs]; results, items and the offsets are required. */
int sp_matcher_fetch_packed( sp_matcher_ctx* ctx, sp_result_packed* results, uint64_t* result_offsets,
			     sp_item_packed* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status);
/* Host side: the documents [0, ndocs) described by result_offsets / item_offsets (as fetched; absolute) unpacked into
 * out_results[result_offsets[ndocs] - result_offsets[0]] and out_items[item_offsets[ndocs] - item_offsets[0]]. */
void sp_unpack_results( const sp_result_packed* results, const uint64_t* result_offsets, const sp_item_packed* items,
			const uint64_t* item_offsets, uint32_t ndocs, sp_result* out_results, sp_item* out_items);

/* Counters of the last run ("results", "items", "tokens", "arena_retries", "pool_retries", "kernel_launches", ...)
 * and tuning knobs ("arena=N", "warps_per_sm=N"); -1 if unknown. */
int64_t sp_matcher_ctx_counter( sp_matcher_ctx* ctx, const char* key);

/* Fused pipeline: lexems of a finished sp_lexer_scan are fed to the matcher without leaving the device. */
int sp_matcher_run_lexer_output( sp_matcher_ctx* ctx, sp_lexer_ctx* lexer_ctx);

/* Count pass of the level-2 window evaluator WITHOUT any allocation. No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving
 * loops, harnesses that forbid it). It answers "how many results and items will every document of this batch give"
 * (what a caller needs to size its own output buffers) with the same kernels as sp_matcher_run (k_eval_prepare,
 * k_eval_count_join, k_eval_count_full, k_eval_docs and the prefix sums between them), their scratch arrays carved
 * from one static device array of SP_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until exit.
 * The call does not call cudaMalloc / cudaFree or any allocating routine; the rule tables must already be on the
 * device (sp_matcher_upload or an earlier context). All pointers are DEVICE pointers of the caller:
 *   in:  tokens[ntokens], tok_offsets[ndocs+1] (as for sp_matcher_run with on_device != 0, origsegs all zero)
 *   out: doc_results[ndocs], doc_items[ndocs] (= the differences of sp_matcher_fetch's result_offsets / item_offsets),
 *        doc_stats[ndocs] (= sp_matcher_fetch's stats), doc_flags[ndocs] (1 = the document is order dependent or
 *        irregular: sp_matcher_run hands it to the state machine; its counts here are 0 and its stats undefined).
 * Limits (beyond them the call fails with SP_ERR_ARGUMENT and a telling last error, nothing is computed): ntokens <=
 * SP_NOALLOC_MAX_TOKENS, ndocs <= SP_NOALLOC_MAX_DOCS, rule instances of the batch <= SP_NOALLOC_MAX_RULES, rules that
 * need the full fold <= SP_NOALLOC_QUEUE; rule sets the evaluator does not cover (sp_matcher_table_info
 * "eval_capable" == 0) are refused. The outputs are complete when the call returns (it waits for the stream).
 * Every call on a device shares the one scratch array, which keeps the contents of the last call (the call clears what
 * has to start at zero): calls that overlap on the GPU (two streams, two host threads) race - serialise them. */
#define SP_NOALLOC_MAX_TOKENS    262144u
#define SP_NOALLOC_MAX_DOCS      65536u
#define SP_NOALLOC_MAX_RULES     (1u << 24)
#define SP_NOALLOC_QUEUE         262144u
#define SP_NOALLOC_SCRATCH_BYTES (16u << 20)
int sp_matcher_count_noalloc( sp_matcher* m, const sp_lexem* tokens_device, const uint64_t* tok_offsets_device, uint32_t ndocs, uint64_t ntokens,
			      uint32_t* doc_results_device, uint32_t* doc_items_device, sp_matcher_stats* doc_stats_device,
			      uint32_t* doc_flags_device, void* cuda_stream);

/* Library identification. */
const char* sp_version(void);
int sp_device_count(void);

#ifdef __cplusplus
}
#endif
#endif
