// Batched PatternLexerContext (included by lexer_capi.cu): owns the per-batch device buffers and drives the
// kernel pipeline count -> verify/fix-up -> scan -> emit -> finish -> resolve -> scan -> write.

struct sp_lexer_ctx
{
	const sp_lexer* lx;
	cudaStream_t stream;
	std::string lasterror;
	uint32_t chunk, warmup;
	// device buffers (grow only)
	DeviceBuffer text, doc_off, start_state, end_state, ev_count, known, redo, scalars, ev_off, events, doc_ev_begin,
		     ar, doc_kept, doc_nlex, lex_off, lexems, status, scantmp,
		     ev_doc, isl_flags, tile_nisl, tile_isl_off, isl_start, isl_kept, isl_off, elem, elem_doc, doc_has_content,
		     doc_elem_begin, aggmax, aggmin, aggord, slots, doc_first_content, fix_list, doc_base,
		     f_region, f_kept, f_agg, f_lead, f_inctail, f_docstart, f_prefix, f_doc_chunk, f_doc_local, f_succ, head;
	bool hybrid;		// SP_LEXER_HYBRID=0: never resolve per chunk over the completed events (island/slot pipeline instead)
	int last_path;		// second half of the last batch: 2 island/slot pipeline, 3 fused scan kernel, 4 chunk resolution over completed events (default)
	bool fused;		// SP_LEXER_FUSED=0 switches the fused segment kernel off (island/slot pipeline for every batch)
	uint64_t fused_batches, fused_fallbacks;
	bool last_fused, fusedGaveUp;
	bool tma;		// SP_LEXER_TMA=1: the scans read their chunks from TMA-staged shared memory tiles instead of global memory
	uint64_t nisl, nelem;
	// last batch
	uint32_t ndocs;
	uint64_t total, nchunks, nev, nlex, fixups, rounds, readbacks;	// readbacks: k_read_back launches of the last scan
	bool valid;
	uint64_t* pinned;	// [4] host staging for scalar read backs
	bool blockingSync;	// SP_BLOCKING_SYNC=1 or counter "sync:blocking": host threads sleep in waits instead of spinning (many contexts per host)
	cudaEvent_t syncEvent;
	// optional per kernel timing with CUDA events on the context's stream (SP_LEXER_PROFILE=1 or counter "profile:on")
	enum KernelId {KCount, KVerify, KEmit, KFinish, KResolve, KWrite, KScan, KIslFlags, KIslResolve, KOrdpos, KIslSlots, KSlotScan, KFused, KFusedScan, KFusedWrite, KResolveChunks, NofKernelIds};
	struct Span {int kernel; cudaEvent_t a, b;};
	bool profile;
	std::vector<Span> spans;
	std::size_t nspans;
	double kernel_ms[ NofKernelIds];
	uint64_t kernel_calls[ NofKernelIds];

	void spanBegin( int kernel)
	{
		if (!profile) return;
		if (nspans == spans.size())
		{
			Span sp; sp.kernel = kernel;
			SP_CUDA( cudaEventCreate( &sp.a)); SP_CUDA( cudaEventCreate( &sp.b));
			spans.push_back( sp);
		}
		spans[ nspans].kernel = kernel;
		SP_CUDA( cudaEventRecord( spans[ nspans].a, stream));
	}
	void spanEnd()
	{
		if (!profile) return;
		SP_CUDA( cudaEventRecord( spans[ nspans].b, stream));
		++nspans;
	}
	void collectSpans()
	{
		for (std::size_t i=0; i<nspans; ++i)
		{
			float ms = 0.0f;
			if (cudaEventElapsedTime( &ms, spans[i].a, spans[i].b) == cudaSuccess) {kernel_ms[ spans[i].kernel] += ms; kernel_calls[ spans[i].kernel] += 1;}
		}
		nspans = 0;
	}

	sp_lexer_ctx( const sp_lexer* lx_, cudaStream_t stream_)
		:lx(lx_),stream(stream_),chunk(512),warmup(96),ndocs(0),total(0),nchunks(0),nev(0),nlex(0),fixups(0),rounds(0),readbacks(0),valid(false),pinned(0),blockingSync(false),syncEvent(0),profile(false),nspans(0)
	{
		std::memset( kernel_ms, 0, sizeof(kernel_ms)); std::memset( kernel_calls, 0, sizeof(kernel_calls));
		if (std::getenv( "SP_LEXER_PROFILE")) profile = true;
		if (const char* be = std::getenv( "SP_BLOCKING_SYNC")) blockingSync = be[0] != '0';
		{
			const char* te = std::getenv( "SP_LEXER_TMA");
			tma = te && te[0] == '1';	// off by default: measured slower, 66 KB of tile per block cost the scan its occupancy
			// 66 KB of dynamic shared memory per block: above the default limit of 48 KB
			if (cudaFuncSetAttribute( k_lex_count<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LexTileBytes) != cudaSuccess
			||  cudaFuncSetAttribute( k_lex_emit<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LexTileBytes) != cudaSuccess) {cudaGetLastError(); tma = false;}
		}
		const char* fe = std::getenv( "SP_LEXER_FUSED");
		fused = (fe && fe[0] == '1');	// the fully fused scan kernel is correct but slower than emit + finish + chunk resolution
		const char* he = std::getenv( "SP_LEXER_HYBRID");
		hybrid = !(he && he[0] == '0');
		last_path = -1;
		fused_batches = fused_fallbacks = 0; last_fused = false;
		nisl = nelem = 0;
		const char* e;
		if ((e = std::getenv( "SP_LEXER_CHUNK")) != 0 && std::atoi( e) >= 16) chunk = (uint32_t)std::atoi( e);
		if ((e = std::getenv( "SP_LEXER_WARMUP")) != 0 && std::atoi( e) >= 0) warmup = (uint32_t)std::atoi( e);
		SP_CUDA( cudaMallocHost( (void**)&pinned, 4 * sizeof(uint64_t)));
	}
	~sp_lexer_ctx()
	{
		if (pinned) cudaFreeHost( pinned);
		for (Span& sp: spans) {cudaEventDestroy( sp.a); cudaEventDestroy( sp.b);}
		if (syncEvent) cudaEventDestroy( syncEvent);
	}

	cudaError_t waitStream()
	{
		if (!blockingSync) return cudaStreamSynchronize( stream);
		if (!syncEvent)
		{
			cudaError_t e = cudaEventCreateWithFlags( &syncEvent, cudaEventBlockingSync | cudaEventDisableTiming);
			if (e != cudaSuccess) return e;
		}
		cudaError_t e = cudaEventRecord( syncEvent, stream);
		return e != cudaSuccess ? e : cudaEventSynchronize( syncEvent);
	}

	uint64_t readScalar( const void* devptr, std::size_t bytes)
	{
		pinned[0] = 0;
		readBackAsync( pinned, devptr, bytes, stream);
		++readbacks;
		SP_CUDA( waitStream());
		return pinned[0];
	}



	// Island discovery shared by both island pipelines: flags, per tile counts, island start list. Returns nisl.
	uint64_t findIslands( IslandParams& Q)
	{
		uint64_t ntiles = (nev + TsTile - 1) / TsTile;
		Q.flags = isl_flags.as<uint8_t>( nev);
		Q.tile_nisl = tile_nisl.as<uint32_t>( ntiles);
		uint64_t* d_tile_isl_off = tile_isl_off.as<uint64_t>( ntiles + 1);
		Q.tile_isl_off = d_tile_isl_off;
		uint64_t* d_aggmax = aggmax.as<uint64_t>( ntiles); uint64_t* d_aggmin = aggmin.as<uint64_t>( ntiles);
		spanBegin( KIslFlags);
		k_isl_reduce<<<(unsigned int)ntiles, TsThreads, 0, stream>>>( Q, d_aggmax, d_aggmin);
		k_tile_carry<OpMax><<<1, 1024, 0, stream>>>( d_aggmax, ntiles, 0, (uint64_t*)0);
		k_tile_carry<OpMin><<<1, 1024, 0, stream>>>( d_aggmin, ntiles, 1, (uint64_t*)0);
		k_isl_flags<<<(unsigned int)ntiles, TsThreads, 0, stream>>>( Q, d_aggmax, d_aggmin);
		exclusiveScan( Q.tile_nisl, ntiles, d_tile_isl_off, scantmp, stream);
		spanEnd();
		uint64_t n = readScalar( d_tile_isl_off + ntiles, sizeof(uint64_t));
		Q.nisl = n;
		Q.isl_start = isl_start.as<uint32_t>( n + 1);
		spanBegin( KIslFlags); k_isl_compact<<<(unsigned int)ntiles, TsThreads, 0, stream>>>( Q); spanEnd();
		return n;
	}

	// fused segment pipeline (lexer_fused.cuh). Returns false if the batch needs the island pipeline.
	bool resolveFused( const SpScanParams& P, uint64_t* d_lex_off, bool overEvents=false)
	{
		const uint32_t M = P.ar_mult;
		// positions advanced and lexems emitted are packed as (low, high) halves of one 64 bit word per chunk and summed together
		if (nev * M >= 0xFFFFFFFFull) throw std::runtime_error( "batch too large: more than 2^32 event slots (split the batch)");
		SpFusedParams F; std::memset( &F, 0, sizeof(F));
		F.T = P.T; F.text = P.text; F.doc_off = P.doc_off; F.ndocs = ndocs; F.total = total; F.chunk = chunk; F.nchunks = nchunks;
		F.start_state = P.start_state; F.ev_off = P.ev_off; F.ar_mult = M;
		F.slots = slots.as<SpSlot>( nev * M + 1);
		F.region_begin = f_region.as<uint64_t>( nchunks); F.kept = f_kept.as<uint32_t>( nchunks); F.agg = f_agg.as<uint64_t>( nchunks);
		F.lead_pred = f_lead.as<uint32_t>( nchunks); F.inc_tail = f_inctail.as<uint32_t>( nchunks); F.has_docstart = f_docstart.as<uint8_t>( nchunks);
		uint64_t* d_prefix = f_prefix.as<uint64_t>( nchunks + 1);
		F.prefix = d_prefix;
		F.doc_chunk = f_doc_chunk.as<uint32_t>( (std::size_t)ndocs + 1); F.doc_local = f_doc_local.as<uint64_t>( (std::size_t)ndocs + 1);
		F.doc_has_content = doc_has_content.as<uint8_t>( (std::size_t)ndocs + 1); F.doc_succ_early = f_succ.as<uint8_t>( (std::size_t)ndocs + 1);
		F.doc_base = doc_base.as<uint64_t>( (std::size_t)ndocs + 1); F.lex_off = d_lex_off;
		uint64_t* d_scalars = (uint64_t*)scalars.ptr();
		F.need_legacy = reinterpret_cast<uint32_t*>( d_scalars + 2);
		F.events = P.events; F.ev_doc = P.ev_doc; F.head = P.head; F.nev = nev;
		SP_CUDA( cudaMemsetAsync( F.doc_chunk, overEvents ? 0xFF : 0, ((std::size_t)ndocs + 1) * sizeof(uint32_t), stream));
		SP_CUDA( cudaMemsetAsync( F.doc_local, 0, ((std::size_t)ndocs + 1) * sizeof(uint64_t), stream));
		SP_CUDA( cudaMemsetAsync( F.doc_has_content, 0, (std::size_t)ndocs + 1, stream));
		SP_CUDA( cudaMemsetAsync( F.doc_succ_early, 0, (std::size_t)ndocs + 1, stream));
		SP_CUDA( cudaMemsetAsync( d_scalars + 1, 0, 2 * sizeof(uint64_t), stream));
		unsigned int grid = (unsigned int)((nchunks + LexFusedThreads - 1) / LexFusedThreads);
		if (overEvents)
		{
			spanBegin( KResolveChunks); k_lex_resolve_chunks<<<grid, LexFusedThreads, 0, stream>>>( F); spanEnd();
		}
		else
		{
			static bool smemSet = false;
			if (!smemSet) {SP_CUDA( cudaFuncSetAttribute( k_lex_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)LexFusedSmem)); smemSet = true;}
			spanBegin( KFused); k_lex_fused<<<grid, LexFusedThreads, LexFusedSmem, stream>>>( F); spanEnd();
		}
		spanBegin( KFusedScan);
		k_fused_fix<<<(unsigned int)((nchunks + 255) / 256), 256, 0, stream>>>( F);
		AggScan f; f.agg = F.agg; f.prefix = d_prefix;
		tileScan( f, nchunks, aggord.as<uint64_t>( (nchunks + TsTile - 1) / TsTile), stream, d_scalars + 1);
		k_fused_docs<<<(ndocs + 1 + 255) / 256, 256, 0, stream>>>( F, d_scalars + 1);
		spanEnd();
		readBackAsync( pinned, d_scalars + 1, 2 * sizeof(uint64_t), stream);
		++readbacks;
		SP_CUDA( waitStream());
		if ((uint32_t)pinned[1]) {++fused_fallbacks; return false;}
		nlex = pinned[0] >> 32;
		F.lexems = lexems.as<sp_lexem>( nlex);
		SP_CUDA( cudaMemsetAsync( P.status, 0, (std::size_t)ndocs * sizeof(uint32_t) + 4, stream));
		spanBegin( KFusedWrite);
		k_fused_write<<<(unsigned int)((nchunks + LexWriteThreads/32 - 1) / (LexWriteThreads/32)), LexWriteThreads, 0, stream>>>( F);
		spanEnd();
		SP_CUDA( cudaGetLastError());
		nisl = nelem = 0;
		++fused_batches;
		last_path = overEvents ? 4 : 3;
		return true;
	}

	// slot pipeline (lexer_islands.cuh): one resolution pass per island + one sum scan over the slots
	void resolveSlots( const SpScanParams& P, uint64_t* d_lex_off)
	{
		nisl = nelem = nlex = 0;
		SP_CUDA( cudaMemsetAsync( P.status, 0, (std::size_t)ndocs * sizeof(uint32_t) + 4, stream));
		if (!nev)
		{
			SP_CUDA( cudaMemsetAsync( d_lex_off, 0, ((std::size_t)ndocs + 1) * sizeof(uint64_t), stream));
			(void)lexems.as<sp_lexem>( 1);
			return;
		}
		const uint64_t nslots = nev * P.ar_mult;
		if (nslots >= 0xFFFFFFFFull) throw std::runtime_error( "batch too large: more than 2^32 event slots (split the batch)");
		IslandParams Q; std::memset( &Q, 0, sizeof(Q));
		Q.T = P.T; Q.events = P.events; Q.ev_doc = P.ev_doc; Q.nev = nev; Q.ndocs = ndocs; Q.ar_mult = P.ar_mult;
		nisl = findIslands( Q);
		SpSlotParams S; std::memset( &S, 0, sizeof(S));
		S.T = P.T; S.events = P.events; S.ev_doc = P.ev_doc; S.nev = nev; S.ar_mult = P.ar_mult;
		S.slots = slots.as<SpSlot>( nslots);
		S.isl_start = Q.isl_start; S.nisl = nisl;
		S.doc_status = P.status;
		S.doc_has_content = doc_has_content.as<uint8_t>( ndocs);
		S.doc_first_content = doc_first_content.as<uint32_t>( ndocs);
		S.fix_list = fix_list.as<uint32_t>( nisl);
		uint64_t* d_scalars = (uint64_t*)scalars.ptr();
		S.fix_count = reinterpret_cast<uint32_t*>( d_scalars + 2);
		S.doc_ev_begin = P.doc_ev_begin; S.ndocs = ndocs;
		S.doc_base = doc_base.as<uint64_t>( (std::size_t)ndocs + 1);
		S.lex_off = d_lex_off;
		SP_CUDA( cudaMemsetAsync( S.doc_has_content, 0, (std::size_t)ndocs + 1, stream));
		SP_CUDA( cudaMemsetAsync( S.doc_first_content, 0xFF, ((std::size_t)ndocs + 1) * sizeof(uint32_t), stream));
		SP_CUDA( cudaMemsetAsync( d_scalars + 1, 0, 2 * sizeof(uint64_t), stream));
		spanBegin( KIslSlots);
		k_isl_to_slots<<<(unsigned int)((nisl + 127) / 128), 128, 0, stream>>>( S);
		k_isl_fix<<<64, 128, 0, stream>>>( S);
		spanEnd();
		// the output holds at most one lexem per slot; its exact size is known after the carry step, so the lexem
		// buffer is sized by the number of valid slots' upper bound (nslots) only when it has to grow
		uint64_t ntiles = (nslots + TsTile - 1) / TsTile;
		uint64_t* d_agg = aggord.as<uint64_t>( ntiles);
		SlotScan f; f.P = S;
		spanBegin( KSlotScan);
		k_tile_reduce<SlotScan><<<(unsigned int)ntiles, TsThreads, 0, stream>>>( f, nslots, d_agg);
		k_tile_carry<OpSum><<<1, 1024, 0, stream>>>( d_agg, ntiles, 0, d_scalars + 1);
		spanEnd();
		nlex = readScalar( d_scalars + 1, sizeof(uint64_t)) >> 32;
		S.lexems = lexems.as<sp_lexem>( nlex);
		f.P = S;
		spanBegin( KSlotScan);
		k_doc_bases<<<(ndocs + 1 + 7) / 8, 256, 0, stream>>>( S, d_agg, d_scalars + 1);
		k_tile_apply<SlotScan><<<(unsigned int)ntiles, TsThreads, 0, stream>>>( f, nslots, d_agg);
		spanEnd();
		SP_CUDA( cudaGetLastError());
	}

	// island pipeline: raw events through HBM (emit, finish), then the second half: chunk resolution (default) or the island/slot pipeline
	void resolveLegacy( SpScanParams& P, uint64_t* d_ev_off, uint64_t* d_doc_ev_begin, uint64_t* d_lex_off)
	{
		P.events = events.as<SpRawEvent>( nev);
		// (not tried again behind a fused scan that gave up: it would give up for the same reason)
		const bool chunkResolution = hybrid && nchunks && nev && P.T.fwd_cut_n && !fusedGaveUp;
		P.head = chunkResolution ? head.as<uint32_t>( nchunks) : 0;
		SP_CUDA( cudaMemsetAsync( d_doc_ev_begin, 0, ((std::size_t)ndocs + 1) * sizeof(uint64_t), stream));
		if (ndocs) SP_CUDA( cudaMemcpyAsync( d_doc_ev_begin + ndocs, d_ev_off + nchunks, sizeof(uint64_t), cudaMemcpyDeviceToDevice, stream));
		if (nchunks)
		{
			unsigned int grid = (unsigned int)((nchunks + LexScanThreads - 1) / LexScanThreads);
			spanBegin( KEmit);
			if (useTma( P)) k_lex_emit<true><<<grid, LexScanThreads, LexTileBytes, stream>>>( P);
			else k_lex_emit<false><<<grid, LexScanThreads, 0, stream>>>( P);
			spanEnd();
		}
		P.ev_doc = ev_doc.as<uint32_t>( nev);
		if (nev)
		{
			unsigned int grid = (unsigned int)((nev + LexEventThreads - 1) / LexEventThreads);
			spanBegin( KFinish); k_lex_finish<<<grid, LexEventThreads, 0, stream>>>( P, nev); spanEnd();
		}
		if (chunkResolution && resolveFused( P, d_lex_off, true)) {}
		else
		{
			if (chunkResolution && nev) k_evdoc_clear<<<(unsigned int)((nev + 255) / 256), 256, 0, stream>>>( P.ev_doc, nev);
			resolveSlots( P, d_lex_off); last_path = 2;
		}
	}

	// bulk copies need 16 byte aligned sources: the text pointer (chunks are 512 bytes)
	bool useTma( const SpScanParams& P) const {return tma && chunk == 512 && (reinterpret_cast<uintptr_t>( P.text) & 15) == 0;}

	void scan( const uint8_t* text_in, int text_on_device, const uint64_t* doc_offsets, uint32_t ndocs_)
	{
		valid = false;
		if (!lx->uploaded) throw std::runtime_error( "called scan without calling 'compile'");
		int dev = -1;
		SP_CUDA( cudaGetDevice( &dev));
		if (dev != lx->device) throw std::runtime_error( "lexer tables live on another device than the current one");
		for (uint32_t d=0; d<ndocs_; ++d) if (doc_offsets[d+1] < doc_offsets[d]) throw std::runtime_error( "document offsets are not ascending");
		if (ndocs_ && doc_offsets[0] != 0) throw std::runtime_error( "document offsets must start at 0");
		ndocs = ndocs_;
		total = ndocs ? doc_offsets[ ndocs] : 0;
		nchunks = (total + chunk - 1) / chunk;
		nev = nlex = fixups = rounds = readbacks = 0;

		SpScanParams P; std::memset( &P, 0, sizeof(P));
		P.T = lx->dev;
		if (text_on_device) P.text = text_in;
		else
		{
			uint8_t* t = text.as<uint8_t>( total + 16);
			if (total) SP_CUDA( cudaMemcpyAsync( t, text_in, total, cudaMemcpyHostToDevice, stream));
			P.text = t;
		}
		uint64_t* d_doc_off = doc_off.as<uint64_t>( (std::size_t)ndocs + 1);
		if (ndocs) SP_CUDA( cudaMemcpyAsync( d_doc_off, doc_offsets, ((std::size_t)ndocs + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, stream));
		else SP_CUDA( cudaMemsetAsync( d_doc_off, 0, sizeof(uint64_t), stream));
		P.doc_off = d_doc_off; P.ndocs = ndocs; P.total = total; P.chunk = chunk; P.warmup = warmup; P.nchunks = nchunks;
		P.start_state = start_state.as<uint32_t>( nchunks); P.end_state = end_state.as<uint32_t>( nchunks);
		P.ev_count = ev_count.as<uint32_t>( nchunks);
		P.known = known.as<uint8_t>( nchunks); P.redo = redo.as<uint8_t>( nchunks);
		uint64_t* d_scalars = scalars.as<uint64_t>( 4);
		P.nbad = reinterpret_cast<uint32_t*>( d_scalars);
		uint64_t* d_ev_off = ev_off.as<uint64_t>( nchunks + 1);
		P.ev_off = d_ev_off;
		uint64_t* d_doc_ev_begin = doc_ev_begin.as<uint64_t>( (std::size_t)ndocs + 1);
		P.doc_ev_begin = d_doc_ev_begin;
		P.doc_kept = doc_kept.as<uint32_t>( ndocs); P.doc_nlex = doc_nlex.as<uint32_t>( ndocs);
		uint64_t* d_lex_off = lex_off.as<uint64_t>( (std::size_t)ndocs + 1);
		P.lex_off = d_lex_off;
		P.status = status.as<uint32_t>( ndocs);
		P.ar_mult = lx->dev.has_symbols ? 2 : 1;

		if (nchunks)
		{
			unsigned int grid = (unsigned int)((nchunks + LexScanThreads - 1) / LexScanThreads);
			unsigned int vgrid = (unsigned int)((nchunks + 255) / 256);
			spanBegin( KCount);
			if (useTma( P)) k_lex_count<true><<<grid, LexScanThreads, LexTileBytes, stream>>>( P, 0);
			else k_lex_count<false><<<grid, LexScanThreads, 0, stream>>>( P, 0);
			spanEnd();
			for (;;)
			{
				SP_CUDA( cudaMemsetAsync( P.nbad, 0, sizeof(uint32_t), stream));
				spanBegin( KVerify); k_lex_verify<<<vgrid, 256, 0, stream>>>( P); spanEnd();
				SP_CUDA( cudaGetLastError());
				uint32_t nbad = (uint32_t)readScalar( P.nbad, sizeof(uint32_t));
				if (!nbad) break;
				fixups += nbad; ++rounds;
				spanBegin( KCount); k_lex_count<false><<<grid, LexScanThreads, 0, stream>>>( P, 1); spanEnd();
			}
		}
		spanBegin( KScan); exclusiveScan( P.ev_count, nchunks, d_ev_off, scantmp, stream); spanEnd();
		nev = readScalar( d_ev_off + nchunks, sizeof(uint64_t));
		last_fused = false;
		const uint64_t fallbacksBefore = fused_fallbacks;
		if (fused && nchunks && nev && P.T.fwd_cut_n && P.T.npat <= SP_FUSED_MAXPAT && P.T.fwd_max_acc + 1 <= SP_FUSED_BUF) last_fused = resolveFused( P, d_lex_off);
		fusedGaveUp = !last_fused && fused_fallbacks != fallbacksBefore;
		if (!last_fused) resolveLegacy( P, d_ev_off, d_doc_ev_begin, d_lex_off);
		SP_CUDA( cudaGetLastError());
		valid = true;
	}
};

extern "C" {

sp_lexer_ctx* sp_lexer_ctx_create( const sp_lexer* lx, void* cuda_stream)
{
	try {return new sp_lexer_ctx( lx, (cudaStream_t)cuda_stream);}
	catch (const std::exception& e) {const_cast<sp_lexer*>( lx)->lasterror = e.what(); return 0;}
}
void sp_lexer_ctx_destroy( sp_lexer_ctx* ctx) {delete ctx;}
const char* sp_lexer_ctx_last_error( const sp_lexer_ctx* ctx) {return ctx->lasterror.c_str();}

int sp_lexer_scan( sp_lexer_ctx* ctx, const uint8_t* text, int text_on_device, const uint64_t* doc_offsets, uint32_t ndocs)
	{SP_LEXER_TRY( ctx, ctx->scan( text, text_on_device, doc_offsets, ndocs))}

int64_t sp_lexer_result_count( sp_lexer_ctx* ctx)
{
	if (!ctx->valid) {ctx->lasterror = "no finished scan"; return SP_ERR_ARGUMENT;}
	if (ctx->waitStream() != cudaSuccess) {ctx->lasterror = "kernel failure"; return SP_ERR_CUDA;}
	return (int64_t)ctx->nlex;
}
const sp_lexem* sp_lexer_result_lexems_device( sp_lexer_ctx* ctx) {return ctx->valid ? (const sp_lexem*)ctx->lexems.ptr() : 0;}
const uint64_t* sp_lexer_result_offsets_device( sp_lexer_ctx* ctx) {return ctx->valid ? (const uint64_t*)ctx->lex_off.ptr() : 0;}
const uint32_t* sp_lexer_result_status_device( sp_lexer_ctx* ctx) {return ctx->valid ? (const uint32_t*)ctx->status.ptr() : 0;}

int sp_lexer_fetch( sp_lexer_ctx* ctx, sp_lexem* lexems, uint64_t* offsets, uint32_t* status)
{
	SP_LEXER_TRY( ctx, {
		if (!ctx->valid) throw std::runtime_error( "no finished scan");
		if (lexems && ctx->nlex) SP_CUDA( cudaMemcpyAsync( lexems, ctx->lexems.ptr(), ctx->nlex * sizeof(sp_lexem), cudaMemcpyDeviceToHost, ctx->stream));
		if (offsets) SP_CUDA( cudaMemcpyAsync( offsets, ctx->lex_off.ptr(), ((std::size_t)ctx->ndocs + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
		if (status && ctx->ndocs) SP_CUDA( cudaMemcpyAsync( status, ctx->status.ptr(), (std::size_t)ctx->ndocs * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		SP_CUDA( ctx->waitStream());
	})
}

int sp_lexer_fetch_range( sp_lexer_ctx* ctx, uint32_t first_doc, uint32_t ndocs, sp_lexem* lexems, uint64_t capacity, uint64_t* offsets, uint32_t* status)
{
	SP_LEXER_TRY( ctx, {
		if (!ctx->valid) throw std::runtime_error( "no finished scan");
		if ((uint64_t)first_doc + ndocs > ctx->ndocs || !offsets) throw std::runtime_error( "document range out of bounds");
		SP_CUDA( cudaMemcpyAsync( offsets, static_cast<const uint64_t*>( ctx->lex_off.ptr()) + first_doc, ((std::size_t)ndocs + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
		SP_CUDA( ctx->waitStream());
		const uint64_t b = offsets[ 0], n = offsets[ ndocs] - b;
		if (lexems)
		{
			if (n > capacity) throw std::runtime_error( "lexem buffer too small for the document range");
			if (n) SP_CUDA( cudaMemcpyAsync( lexems, static_cast<const sp_lexem*>( ctx->lexems.ptr()) + b, n * sizeof(sp_lexem), cudaMemcpyDeviceToHost, ctx->stream));
		}
		if (status && ndocs) SP_CUDA( cudaMemcpyAsync( status, static_cast<const uint32_t*>( ctx->status.ptr()) + first_doc, (std::size_t)ndocs * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
		SP_CUDA( ctx->waitStream());
	})
}

int64_t sp_lexer_ctx_counter( sp_lexer_ctx* ctx, const char* key)
{
	std::string k( key);
	if (k == "raw_events") return (int64_t)ctx->nev;
	if (k == "tma") return ctx->tma ? 1 : 0;
	if (k == "tma=0") {ctx->tma = false; return 0;}
	if (k == "tma=1") {ctx->tma = true; return 1;}
	if (k == "lexems") return (int64_t)ctx->nlex;
	if (k == "chunks") return (int64_t)ctx->nchunks;
	if (k == "fixup_chunks") return (int64_t)ctx->fixups;
	if (k == "fixup_rounds") return (int64_t)ctx->rounds;
	if (k == "islands") return (int64_t)ctx->nisl;
	if (k == "fused") return ctx->last_fused ? 1 : 0;
	if (k == "second_half") return ctx->last_path;
	if (k == "fused_batches") return (int64_t)ctx->fused_batches;
	if (k == "fused_fallbacks") return (int64_t)ctx->fused_fallbacks;
	if (k == "elements") return (int64_t)ctx->nelem;
	if (k == "text_bytes") return (int64_t)ctx->total;
	if (k == "ndocs") return ctx->valid ? (int64_t)ctx->ndocs : -1;
	if (k == "chunk_bytes") return (int64_t)ctx->chunk;
	if (k == "profile:on") {ctx->profile = true; return 1;}
	if (k == "sync:blocking") {ctx->blockingSync = true; return 1;}
	if (k == "sync:spin") {ctx->blockingSync = false; return 1;}
	if (k == "profile:off") {ctx->profile = false; return 0;}
	if (k == "profile:reset") {std::memset( ctx->kernel_ms, 0, sizeof(ctx->kernel_ms)); std::memset( ctx->kernel_calls, 0, sizeof(ctx->kernel_calls)); return 0;}
	if (k.compare( 0, 8, "time_us:") == 0 || k.compare( 0, 6, "calls:") == 0)
	{
		static const char* names[] = {"k_lex_count","k_lex_verify","k_lex_emit","k_lex_finish","k_lex_resolve","k_lex_write","k_scan","k_isl_flags","k_isl_resolve","k_ordpos","k_isl_slots","k_slot_scan","k_lex_fused","k_fused_scan","k_fused_write","k_lex_resolve_chunks"};
		static_assert( sizeof(names)/sizeof(names[0]) == sp_lexer_ctx::NofKernelIds, "kernel name table out of date");
		if (ctx->waitStream() != cudaSuccess) return -1;
		try {ctx->collectSpans();} catch (...) {return -1;}
		bool calls = k[0] == 'c';
		std::string name = k.substr( calls ? 6 : 8);
		for (int i=0; i<sp_lexer_ctx::NofKernelIds; ++i)
		{
			if (name == names[i]) return calls ? (int64_t)ctx->kernel_calls[i] : (int64_t)(ctx->kernel_ms[i] * 1000.0);
		}
		return -1;
	}
	if (k == "kernel_launches")
	{
		// count, verify (+2 per repair round), 3 scan, emit, finish; then the chosen second half
		// fused: count, verify (+2 per repair round), 3 scan, fused, fix, 3 scan, docs, write
		// (+ the scalar read backs, one k_read_back launch each)
		if (ctx->last_path == 3) return 1 + 1 + 2*(int64_t)ctx->rounds + 3 + 1 + 1 + 3 + 1 + 1 + (int64_t)ctx->readbacks;
		// chunk resolution: count, verify, 3 scan, emit, finish, resolve, fix, 3 scan, docs, write
		if (ctx->last_path == 4) return 1 + 1 + 2*(int64_t)ctx->rounds + 3 + 1 + 1 + 1 + 1 + 3 + 1 + 1 + (int64_t)ctx->readbacks;
		int64_t first = 1 + 1 + 2*(int64_t)ctx->rounds + 3 + 1 + 1;
		return first + 8 + 2 + 2 + 2 + (int64_t)ctx->readbacks;
	}
	return -1;
}

const char* sp_version(void) {return "strus_pattern_b200 0.1.0";}
int sp_device_count(void) {int n = 0; return cudaGetDeviceCount( &n) == cudaSuccess ? n : 0;}

} // extern "C"
