// Compiled lexer tables as they live in HBM (plain pointers; the same struct describes the host staging copy).
// Built by lexer_compile.hpp from the lexem definitions; this is what replaces the opaque hs_database_t of
// the reference (src/patternLexer.cpp:1083-1086, 1114).
#ifndef SP_LEXER_TABLES_H_INCLUDED
#define SP_LEXER_TABLES_H_INCLUDED
#include <stdint.h>

#define SP_REV_ACCEPT 0x80000000u	/* reverse transition: a match start lies in front of the byte just consumed */
#define SP_REV_STOP   0x40000000u	/* reverse transition: no further start can follow */
#define SP_REV_ROWMASK 0x3FFFFFFFu
#define SP_NO_POS 0xFFFFFFFFu

/* One lexem definition = PatternDef of the reference (src/patternLexer.cpp:45-177). */
typedef struct SpPatternInfo
{
	uint32_t id;			/* lexem id reported */
	uint8_t level;
	uint8_t posbind;
	uint8_t symtab;			/* 1-based symbol table or 0 */
	uint8_t resultidx;		/* selected sub expression or 0 */
	uint32_t rev_init[4];		/* reverse automaton: initial row by kind of the byte behind the match */
	uint32_t rev_stride;		/* reverse automaton: row length; last column = begin of text */
	uint32_t rev_class;		/* offset of this definition's 256 entry byte class map in rev_classes */
	uint32_t cap_first;		/* index of the five capture selection automata in `cap` (resultidx != 0) */
} SpPatternInfo;

/* Anchored forward automaton used by the capture selection (whole, A, GB, G, B of "A(G)B"). */
typedef struct SpAnchoredDfa
{
	uint32_t init[4];		/* initial row by kind of the previous byte */
	uint32_t stride;
	uint32_t classes;		/* offset into cap_classes */
	uint32_t accept_on_entry;	/* nonzero if the empty string matches in every context (never used: see cap_next flags) */
} SpAnchoredDfa;

typedef struct SpSymbolSlot
{
	uint32_t hash;			/* 0 = empty slot */
	uint32_t symid;
	uint32_t str;			/* offset into symbol_pool */
	uint16_t len;
	uint16_t symtab;
} SpSymbolSlot;

/* the events of one accepting state: fwd_acc_pat[off .. off+n) (one 8-byte load instead of two loads of fwd_acc_off) */
typedef struct alignas(8) SpAccSpan
{
	uint32_t off, n;
} SpAccSpan;

typedef struct SpLexerTables
{
	/* forward automaton over all definitions */
	const uint16_t* fwd_next;	/* [fwd_states][fwd_stride] state ids; column fwd_stride-1 = end of text */
	const uint8_t* fwd_class;	/* [256] byte -> column */
	const uint32_t* fwd_acc_off;	/* [fwd_states - fwd_acc_base + 1] */
	const SpAccSpan* fwd_acc_span;	/* [fwd_states - fwd_acc_base] the same as (offset, count) pairs */
	const uint16_t* fwd_acc_pat;	/* definition indices (0-based), ascending per state */
	uint32_t fwd_states;
	uint32_t fwd_stride;
	uint32_t fwd_init;
	uint32_t fwd_acc_base;		/* states >= fwd_acc_base report matches */
	uint32_t fwd_cut_lo;		/* states in [fwd_cut_lo, fwd_cut_lo+fwd_cut_n): every match still pending started at the byte just consumed or later */
	uint32_t fwd_cut_n;		/* 0 = automaton built without cut states (fused kernel not usable) */
	uint32_t fwd_max_acc;		/* largest number of definitions one state reports */
	/* reverse automata, one per definition */
	const uint32_t* rev_next;	/* entries: absolute row offset | SP_REV_ACCEPT | SP_REV_STOP */
	const uint8_t* rev_classes;
	/* capture selection automata */
	const uint32_t* cap_next;	/* same entry format as rev_next */
	const uint8_t* cap_classes;
	const SpAnchoredDfa* cap;
	/* definitions and symbols */
	const SpPatternInfo* pat;
	uint32_t npat;
	const SpSymbolSlot* sym_slots;
	uint32_t sym_mask;		/* slots-1 (power of two) or 0 if there are no symbols */
	const uint8_t* sym_pool;
	uint32_t has_symbols;
} SpLexerTables;

#endif
