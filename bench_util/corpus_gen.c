/* BENCH/TEST INFRASTRUCTURE: deterministic synthetic corpus of the shape SURVEY.md 8(d) C2 names.
 * Counter based RNG (splitmix64 of seed, document, draw) so that every document is generated independently
 * (threads) and identically on every machine. Valid UTF-8 only, no NUL, tokens far below 65535 bytes.
 *   token mix by draw: 78% vocabulary words (Zipf S=1; ASCII share capitalised 12%), 8% numbers, 2% dates,
 *   1% e-mails, 1% URLs, 10% punctuation; single space separators, '\n' about every 80 bytes.
 * Vocabulary: rank 0..nfixed-1 are the words given by the caller (phrase parts, abbreviations), the rest
 * are generated: `nonascii` share of them from Latin-1 supplement / Greek / Cyrillic (2 byte), CJK (3 byte)
 * and 0.1% from U+1F600.. (4 byte). */
#include <stdint.h>
#include <stdlib.h>
#include <stdio.h>
#include <string.h>
#include <math.h>
#include <pthread.h>

static uint64_t mix64( uint64_t x)
{
	x += 0x9E3779B97F4A7C15ull;
	x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
	x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
	return x ^ (x >> 31);
}
typedef struct {uint64_t key, ctr;} rng_t;
static uint64_t rnd( rng_t* r) {return mix64( r->key ^ mix64( r->ctr++));}
static double rndu( rng_t* r) {return (double)(rnd( r) >> 11) * (1.0 / 9007199254740992.0);}
static uint32_t rndn( rng_t* r, uint32_t n) {return (uint32_t)(rnd( r) % n);}

typedef struct
{
	uint64_t seed;
	const char* vocab; const uint32_t* vocab_off; uint32_t nvocab;	/* concatenated words */
	const double* cdf;						/* Zipf cumulative, nvocab entries */
	uint8_t* text; const uint64_t* doc_off; uint32_t ndocs;
	uint32_t tid, nthreads;
} job_t;

static int put_utf8( char* o, uint32_t cp)
{
	if (cp < 0x80) {o[0] = (char)cp; return 1;}
	if (cp < 0x800) {o[0] = (char)(0xC0 | (cp >> 6)); o[1] = (char)(0x80 | (cp & 0x3F)); return 2;}
	if (cp < 0x10000) {o[0] = (char)(0xE0 | (cp >> 12)); o[1] = (char)(0x80 | ((cp >> 6) & 0x3F)); o[2] = (char)(0x80 | (cp & 0x3F)); return 3;}
	o[0] = (char)(0xF0 | (cp >> 18)); o[1] = (char)(0x80 | ((cp >> 12) & 0x3F)); o[2] = (char)(0x80 | ((cp >> 6) & 0x3F)); o[3] = (char)(0x80 | (cp & 0x3F));
	return 4;
}

/* generated vocabulary word of rank `rank` into o (max 64 bytes); returns its length */
int corpus_vocab_word( uint64_t seed, uint32_t rank, double nonascii, char* o)
{
	rng_t r = {mix64( seed ^ 0x766F636162ull) ^ rank, 0};
	int n = 0, len = 2 + (int)rndn( &r, 9);
	double u = rndu( &r);
	if (u >= nonascii)
	{
		for (int i=0; i<len; ++i) o[n++] = (char)('a' + rndn( &r, 26));
		return n;
	}
	double v = rndu( &r);
	int cap = rndu( &r) < 0.12;
	for (int i=0; i<len; ++i)
	{
		uint32_t cp;
		if (v < 0.001) cp = 0x1F600 + rndn( &r, 64);
		else if (v < 0.30) cp = (cap && i == 0) ? 0xC0 + rndn( &r, 23) : 0xE0 + rndn( &r, 23);		/* Latin-1: À..Ö / à..ö */
		else if (v < 0.55) cp = (cap && i == 0) ? 0x391 + rndn( &r, 17) : 0x3B1 + rndn( &r, 17);	/* Greek Α..Ρ / α..ρ */
		else if (v < 0.80) cp = (cap && i == 0) ? 0x410 + rndn( &r, 32) : 0x430 + rndn( &r, 32);	/* Cyrillic */
		else {cp = 0x4E00 + rndn( &r, 0x5000); if (i >= 4) break;}					/* CJK ideographs */
		n += put_utf8( o + n, cp);
	}
	return n;
}

static uint32_t zipf_rank( const double* cdf, uint32_t n, double u)
{
	double x = u * cdf[ n-1];
	uint32_t lo = 0, hi = n-1;
	while (lo < hi) {uint32_t mid = (lo + hi) >> 1; if (cdf[ mid] < x) lo = mid+1; else hi = mid;}
	return lo;
}

static int gen_token( const job_t* J, rng_t* r, char* o)
{
	static const char* tlds[] = {"com","org","net","edu","gov","de","ch","uk","fr","io"};
	static const char punct[] = ".,.;:.!?,\"'().,[]";
	double u = rndu( r);
	int n = 0;
	if (u < 0.78)
	{
		uint32_t k = zipf_rank( J->cdf, J->nvocab, rndu( r));
		uint32_t b = J->vocab_off[ k], e = J->vocab_off[ k+1];
		memcpy( o, J->vocab + b, e - b);
		n = (int)(e - b);
		if ((unsigned char)o[0] >= 'a' && (unsigned char)o[0] <= 'z' && rndu( r) < 0.12) o[0] = (char)(o[0] - 32);
	}
	else if (u < 0.86)
	{
		int digits = 1 + (int)rndn( r, 6);
		for (int i=0; i<digits; ++i) o[n++] = (char)('0' + rndn( r, 10));
		double w = rndu( r);
		if (w < 0.2) {o[n++] = rndn( r, 2) ? '.' : ','; o[n++] = (char)('0' + rndn( r, 10)); o[n++] = (char)('0' + rndn( r, 10));}
		else if (w < 0.25) {memmove( o+1, o, n); o[0] = '$'; n++;}
		else if (w < 0.30) {o[n++] = 't'; o[n++] = 'h';}
	}
	else if (u < 0.88)
	{
		if (rndn( r, 2)) n = sprintf( o, "%04u-%02u-%02u", 1900 + rndn( r, 200), 1 + rndn( r, 12), 1 + rndn( r, 28));
		else n = sprintf( o, "%u.%u.%04u", 1 + rndn( r, 28), 1 + rndn( r, 12), 1900 + rndn( r, 200));
	}
	else if (u < 0.89)
	{
		int a = 3 + (int)rndn( r, 8), b = 3 + (int)rndn( r, 8);
		for (int i=0; i<a; ++i) o[n++] = (char)('a' + rndn( r, 26));
		o[n++] = '@';
		for (int i=0; i<b; ++i) o[n++] = (char)('a' + rndn( r, 26));
		o[n++] = '.';
		n += sprintf( o + n, "%s", tlds[ rndn( r, 10)]);
	}
	else if (u < 0.90)
	{
		int b = 3 + (int)rndn( r, 10);
		n = sprintf( o, rndn( r, 2) ? "http://" : (rndn( r, 2) ? "https://" : "www."));
		for (int i=0; i<b; ++i) o[n++] = (char)('a' + rndn( r, 26));
		o[n++] = '.';
		n += sprintf( o + n, "%s", tlds[ rndn( r, 10)]);
		if (o[0] == 'h' && rndn( r, 2)) {o[n++] = '/'; for (int i=0; i<5; ++i) o[n++] = (char)('a' + rndn( r, 26));}
	}
	else o[n++] = punct[ rndn( r, sizeof(punct)-1)];
	return n;
}

static void* worker( void* arg)
{
	const job_t* J = (const job_t*)arg;
	char tok[ 160];
	for (uint32_t d=J->tid; d<J->ndocs; d+=J->nthreads)
	{
		rng_t r = {mix64( J->seed ^ mix64( 0x646F63ull + d)), 0};
		uint8_t* o = J->text + J->doc_off[ d];
		uint64_t len = J->doc_off[ d+1] - J->doc_off[ d], n = 0, col = 0;
		while (n < len)
		{
			int tn = gen_token( J, &r, tok);
			int punct = (tn == 1 && !((tok[0] | 0x20) >= 'a' && (tok[0] | 0x20) <= 'z') && !(tok[0] >= '0' && tok[0] <= '9'));
			if (n + tn + 1 > len) {memset( o + n, ' ', len - n); break;}
			if (punct && n && o[ n-1] == ' ' && tok[0] != '(' && tok[0] != '[' && tok[0] != '"') --n;	/* "word." not "word ." */
			memcpy( o + n, tok, tn); n += tn;
			col += tn + 1;
			if (col > 80) {o[ n++] = '\n'; col = 0;} else o[ n++] = ' ';
		}
	}
	return 0;
}

/* document sizes log-uniform in [min_doc,max_doc], total = target_bytes exactly; returns the number of documents */
uint32_t corpus_layout( uint64_t seed, uint64_t target_bytes, uint32_t min_doc, uint32_t max_doc, uint64_t* doc_off, uint32_t max_docs)
{
	rng_t r = {mix64( seed ^ 0x6C61796F7574ull), 0};
	uint64_t total = 0; uint32_t nd = 0;
	doc_off[0] = 0;
	while (total < target_bytes && nd < max_docs)
	{
		double l = log( (double)min_doc) + rndu( &r) * (log( (double)max_doc) - log( (double)min_doc));
		uint64_t len = (uint64_t)exp( l);
		if (total + len > target_bytes || nd + 1 == max_docs) len = target_bytes - total;
		total += len;
		doc_off[ ++nd] = total;
	}
	return nd;
}

int corpus_fill( uint64_t seed, const char* vocab, const uint32_t* vocab_off, uint32_t nvocab, double zipf_s,
		 uint8_t* text, const uint64_t* doc_off, uint32_t ndocs, uint32_t nthreads)
{
	double* cdf = (double*)malloc( sizeof(double) * nvocab);
	double acc = 0.0;
	for (uint32_t k=0; k<nvocab; ++k) {acc += 1.0 / pow( (double)(k+1), zipf_s); cdf[ k] = acc;}
	if (nthreads < 1) nthreads = 1;
	if (nthreads > 256) nthreads = 256;
	pthread_t th[ 256]; job_t jobs[ 256];
	for (uint32_t t=0; t<nthreads; ++t)
	{
		job_t j = {seed, vocab, vocab_off, nvocab, cdf, text, doc_off, ndocs, t, nthreads};
		jobs[ t] = j;
		pthread_create( &th[ t], 0, worker, &jobs[ t]);
	}
	for (uint32_t t=0; t<nthreads; ++t) pthread_join( th[ t], 0);
	free( cdf);
	return 0;
}
