// Batched PatternMatcherContext (included by matcher_capi.cu): owns the per-warp arenas, the output pools and the
// result buffers, and drives k_match_documents -> k_match_collect (-> another round for overflowed documents)
// -> prefix sums -> k_match_compact.

struct sp_matcher_ctx
{
	sp_matcher* m;
	cudaStream_t stream;
	std::string lasterror;
	// configuration
	SpArenaCaps cap;		// arena capacities of the first round (grown by 4 for documents that overflow)
	uint32_t warps_per_sm;
	bool warps_fixed;	// warps per SM set by the caller (SP_MATCHER_WARPS_PER_SM / counter "warps_per_sm="): no automatic choice
	uint32_t cap_growths;
	uint64_t arena_budget;		// bytes of HBM the arenas may take
	int nsm;
	// device buffers (grow only)
	DeviceBuffer tokens, origsegs, tok_off, arenas, res_pool, item_pool, scalars, doc_res_begin, doc_res_count, doc_item_begin,
		     doc_item_count, stats, status, list_a, list_b, res_off, item_off, results, items, scantmp,
		     tok_nres, tok_nitems, tok_res_off, tok_item_off, flags, flag_list, tok_slot, tok_work, tok_docs, tok_work_off, resbits, queue,
		     packed_results, packed_items;
	uint64_t queue_cap, eval_queued, eval_items;
	int write_blocks;	// blocks per SM the write pass is compiled for: 5 (96 registers, default; measured per 64 MiB of C3 text: 3 blocks = 168 registers 13.8 ms, 4 = 128 registers 11.4 ms, 5 = 96 registers 10.2 ms, 6 = 80 registers 10.9 ms); SP_EVAL_WRITE_BLOCKS=3|4|6 for A/B runs
	bool count_by_steps;	// SP_EVAL_COUNT=step: the count pass that follows every rule token by token (k_eval_count_step), for A/B runs
	int mode;			// 0 = automatic (window evaluator if the rule set allows it), 1 = state machine (emission order), 2 = evaluator required
	uint64_t eval_docs, eval_flagged, eval_runs;
	double eval_count_ms, eval_write_ms;
	double res_per_token, items_per_token;	// output pool sizing, learned from the last run
	// last batch
	uint32_t ndocs;
	uint64_t ntokens, nres, nitems, retries, pool_retries, launches;
	bool valid;
	uint64_t* pinned;		// host staging for scalar read backs [8]
	cudaEvent_t ev_a, ev_b;
	double kernel_ms; uint64_t kernel_calls;
	bool profile, trace;	// SP_MATCHER_TRACE=1: one line per round of the run loop on stderr

	sp_matcher_ctx( sp_matcher* m_, cudaStream_t stream_)
		:m(m_),stream(stream_),warps_per_sm(16),warps_fixed(false),cap_growths(0),arena_budget(0),nsm(0),res_per_token(0.5),items_per_token(1.0)
		,ndocs(0),ntokens(0),nres(0),nitems(0),retries(0),pool_retries(0),launches(0),valid(false),pinned(0),ev_a(0),ev_b(0)
		,kernel_ms(0.0),kernel_calls(0),profile(false),mode(0),eval_docs(0),eval_flagged(0),eval_runs(0),eval_count_ms(0.0),eval_write_ms(0.0),queue_cap(0),eval_queued(0),eval_items(0),count_by_steps(false)
	{
		cap.rules = 2048; cap.triggers = 2048; cap.items = 8192; cap.datarefs = 4096; cap.results = 4096;
		cap.follow = 256; cap.dispose = 2048; cap.heap = 1024;
		const char* e;
		if ((e = std::getenv( "SP_MATCHER_ARENA")) != 0 && std::atoi( e) >= 4)
		{
			uint32_t n = (uint32_t)std::atoi( e);
			cap.rules = cap.triggers = cap.datarefs = cap.results = cap.dispose = n; cap.items = 4*n; cap.follow = n < 64 ? n : n/8; cap.heap = n/2 + 1;
		}
		if ((e = std::getenv( "SP_MATCHER_WARPS_PER_SM")) != 0 && std::atoi( e) >= 1) {warps_per_sm = (uint32_t)std::atoi( e); warps_fixed = true;}
		if (std::getenv( "SP_MATCHER_PROFILE")) profile = true;
		if ((e = std::getenv( "SP_EVAL_COUNT")) != 0) count_by_steps = !std::strcmp( e, "step");
		write_blocks = ((e = std::getenv( "SP_EVAL_WRITE_BLOCKS")) != 0 && std::atoi( e) >= 3 && std::atoi( e) <= 6) ? std::atoi( e) : 5;
		if ((e = std::getenv( "SP_MATCHER_MODE")) != 0) mode = !std::strcmp( e, "statemachine") ? 1 : !std::strcmp( e, "eval") ? 2 : 0;
		trace = std::getenv( "SP_MATCHER_TRACE") != 0;
		int dev = 0;
		SP_CUDA( cudaGetDevice( &dev));
		SP_CUDA( cudaDeviceGetAttribute( &nsm, cudaDevAttrMultiProcessorCount, dev));
		std::size_t freeb = 0, totalb = 0;
		SP_CUDA( cudaMemGetInfo( &freeb, &totalb));
		arena_budget = freeb / 2;	// the lexer context of a pipeline and the result pools need the rest
		SP_CUDA( cudaMallocHost( (void**)&pinned, 8 * sizeof(uint64_t)));
		SP_CUDA( cudaEventCreate( &ev_a)); SP_CUDA( cudaEventCreate( &ev_b));
	}
	~sp_matcher_ctx()
	{
		if (pinned) cudaFreeHost( pinned);
		if (ev_a) cudaEventDestroy( ev_a);
		if (ev_b) cudaEventDestroy( ev_b);
	}

	void readScalars( const void* devptr, std::size_t bytes)
	{
		readBackAsync( pinned, devptr, bytes, stream);
		++launches;	// k_read_back
		SP_CUDA( cudaStreamSynchronize( stream));
	}

	// The per-document state machine (k_match_documents) over all documents (init_list == 0) or over the listed ones:
	// fills the pools and the per document descriptors (doc_res_begin/count, doc_item_begin/count, stats, status).
	void runStateMachine( MatchParams& P, uint32_t* init_list, uint32_t init_nlist)
	{
		uint64_t* d_scalars = scalars.as<uint64_t>( 8);
		P.cursors = reinterpret_cast<unsigned long long*>( d_scalars);
		uint32_t* d_words = reinterpret_cast<uint32_t*>( d_scalars + 2);	// [0] next_doc [1] list count [2..4] overflow kinds
		P.next_doc = d_words;
		const uint64_t ntok_run = init_list ? std::max<uint64_t>( 1, ntokens * init_nlist / std::max<uint32_t>( ndocs, 1)) : ntokens;
		uint64_t res_cap = (uint64_t)(res_per_token * 1.25 * (double)ntok_run) + 4096;
		uint64_t item_cap = (uint64_t)(items_per_token * 1.25 * (double)ntok_run) + 8192;
		SpArenaCaps curcap = cap;
		uint32_t* d_list = init_list; uint32_t nlist = init_nlist;
		bool flip = false, fresh = true;
		uint32_t retry_docs = 0;
		for (int round=0;; ++round)
		{
			if (round > 40) throw std::runtime_error( "matcher arenas keep overflowing (rule set too large for the device memory)");
			sp_arena_layout( curcap, P.T.nstopwords, P.layout);
			uint32_t ntodo = d_list ? nlist : ndocs;
			uint64_t maxwarps = arena_budget / P.layout.bytes;
			if (maxwarps < 1) throw std::runtime_error( "matcher arena does not fit the device memory");
			const bool wide = !warps_fixed && (uint64_t)ntodo > (uint64_t)nsm * 16 && maxwarps > (uint64_t)nsm * 16;
			const uint32_t wps = wide ? 32 : warps_per_sm;
			uint64_t nwarps = std::min<uint64_t>( std::min<uint64_t>( ntodo, (uint64_t)nsm * wps), maxwarps);
			uint32_t nblocks = (uint32_t)((nwarps + MatchWarpsPerBlock - 1) / MatchWarpsPerBlock);
			P.arena_base = arenas.as<uint8_t>( (std::size_t)nblocks * MatchWarpsPerBlock * P.layout.bytes);
			// the stop word log is validated by a per-warp document stamp that restarts with every launch
			SP_CUDA( cudaMemset2DAsync( P.arena_base + P.layout.stoplog, P.layout.bytes, 0, ((std::size_t)P.T.nstopwords + 1) * sizeof(SpStopLog),
						    (std::size_t)nblocks * MatchWarpsPerBlock, stream));
			P.res_pool = res_pool.as<sp_result>( res_cap); P.item_pool = item_pool.as<sp_item>( item_cap);
			P.res_cap = res_cap; P.item_cap = item_cap;
			P.doc_list = d_list; P.nlist = nlist;
			if (fresh) SP_CUDA( cudaMemsetAsync( d_scalars, 0, 8 * sizeof(uint64_t), stream));
			else SP_CUDA( cudaMemsetAsync( d_words, 0, 5 * sizeof(uint32_t), stream));
			fresh = false;
			uint32_t* d_newlist = (flip ? list_b : list_a).as<uint32_t>( ntodo);
			if (profile || trace) SP_CUDA( cudaEventRecord( ev_a, stream));
			if (wide || warps_per_sm > 16) k_match_documents<8><<<nblocks, MatchThreads, 0, stream>>>( P);
			else k_match_documents<4><<<nblocks, MatchThreads, 0, stream>>>( P);
			if (profile || trace) SP_CUDA( cudaEventRecord( ev_b, stream));
			k_match_collect<<<(ndocs + 255) / 256, 256, 0, stream>>>( P.status, ndocs, d_newlist, d_words + 1, d_words + 2);
			SP_CUDA( cudaGetLastError());
			launches += 2;
			readScalars( d_scalars, 8 * sizeof(uint64_t));
			if (profile)
			{
				float ms = 0.0f;
				if (cudaEventElapsedTime( &ms, ev_a, ev_b) == cudaSuccess) {kernel_ms += ms; kernel_calls += 1;}
			}
			if (trace)
			{
				float ms = 0.0f; cudaEventElapsedTime( &ms, ev_a, ev_b);
				const uint32_t* w = reinterpret_cast<const uint32_t*>( pinned + 2);
				fprintf( stderr, "[sp_matcher] round %d: docs %u, warps %llu, arena %llu KB (rules %u triggers %u items %u results %u follow %u dispose %u heap %u), kernel %.1f ms, overflowed docs %u (arena %u output %u depth %u), results %llu items %llu\n",
					round, ntodo, (unsigned long long)nwarps, (unsigned long long)(P.layout.bytes >> 10), curcap.rules, curcap.triggers, curcap.items, curcap.results,
					curcap.follow, curcap.dispose, curcap.heap, ms, w[1], w[2], w[3], w[4], (unsigned long long)pinned[0], (unsigned long long)pinned[1]);
			}
			uint64_t want_res = pinned[0], want_items = pinned[1];
			const uint32_t* words = reinterpret_cast<const uint32_t*>( pinned + 2);
			uint32_t listcount = words[1];
			bool ovfArena = words[2] != 0, ovfOutput = words[3] != 0, ovfDepth = words[4] != 0;
			if (ovfDepth) throw std::runtime_error( "nesting of result items too deep");
			if (ovfOutput)
			{
				// output pools too small: size them from what this round asked for and run everything again
				++pool_retries;
				res_cap = std::max<uint64_t>( res_cap * 2, want_res + want_res / 4 + 4096);
				item_cap = std::max<uint64_t>( item_cap * 2, want_items + want_items / 4 + 8192);
				d_list = init_list; nlist = init_nlist; fresh = true;
				continue;
			}
			if (!ovfArena) break;
			++retries;
			// only the pools that overflowed grow (a document stops at its first overflow, so another pool may follow in
			// the next round): the arena stays small enough for a full grid of warps (growing all pools together ended
			// at 400 MB per warp and 113 warps in flight on the C3 rule set)
			{
				const uint32_t kinds = words[2];
				if (kinds & (1u << (SP_MST_OVF_RULES - SP_MST_OVERFLOW))) curcap.rules *= 4;
				if (kinds & (1u << (SP_MST_OVF_TRIGGERS - SP_MST_OVERFLOW))) curcap.triggers *= 4;
				if (kinds & (1u << (SP_MST_OVF_ITEMS - SP_MST_OVERFLOW))) curcap.items *= 4;
				if (kinds & (1u << (SP_MST_OVF_DATAREFS - SP_MST_OVERFLOW))) curcap.datarefs *= 4;
				if (kinds & (1u << (SP_MST_OVF_RESULTS - SP_MST_OVERFLOW))) curcap.results *= 4;
				if (kinds & (1u << (SP_MST_OVF_FOLLOW - SP_MST_OVERFLOW))) curcap.follow *= 4;
				if (kinds & (1u << (SP_MST_OVF_DISPOSE - SP_MST_OVERFLOW))) curcap.dispose *= 4;
				if (kinds & (1u << (SP_MST_OVF_HEAP - SP_MST_OVERFLOW))) curcap.heap *= 4;
			}
			d_list = d_newlist; nlist = listcount; flip = !flip;
			retry_docs = std::max( retry_docs, listcount);
		}
		// documents of the next batch will look like these: start with the arenas that were needed (two growth steps at most)
		if (std::memcmp( &curcap, &cap, sizeof(cap)) != 0 && cap_growths < 4 && (uint64_t)retry_docs * 50 >= (init_list ? init_nlist : ndocs)) {cap = curcap; ++cap_growths;}
	}

	// tokens / origsegs / tok_offsets already on the device; ntokens known
	void runDevice( const sp_lexem* d_tokens, const uint32_t* d_origsegs, const uint64_t* d_tok_off, uint32_t ndocs_, uint64_t ntokens_)
	{
		valid = false;
		m->upload();
		int dev = -1;
		SP_CUDA( cudaGetDevice( &dev));
		if (dev != m->device) throw std::runtime_error( "matcher tables live on another device than the current one");
		ndocs = ndocs_; ntokens = ntokens_;
		nres = nitems = 0;

		MatchParams P; std::memset( &P, 0, sizeof(P));
		P.T = m->dev;
		P.tokens = d_tokens; P.origsegs = d_origsegs; P.tok_off = d_tok_off; P.ndocs = ndocs;
		P.doc_res_begin = doc_res_begin.as<uint64_t>( ndocs); P.doc_res_count = doc_res_count.as<uint32_t>( ndocs);
		P.doc_item_begin = doc_item_begin.as<uint64_t>( ndocs); P.doc_item_count = doc_item_count.as<uint32_t>( ndocs);
		P.stats = stats.as<sp_matcher_stats>( ndocs); P.status = status.as<uint32_t>( ndocs);
		uint64_t* d_res_off = res_off.as<uint64_t>( (std::size_t)ndocs + 1);
		uint64_t* d_item_off = item_off.as<uint64_t>( (std::size_t)ndocs + 1);
		if (mode == 2 && !P.T.eval_capable) throw std::runtime_error( "the rule set is outside of what the window evaluator reproduces (mode eval)");
		const bool useEval = mode != 1 && P.T.eval_capable && ndocs && ntokens;

		if (!useEval)
		{
			if (ndocs) runStateMachine( P, 0, 0);
			// document order: offsets by prefix sum, then one warp per document copies its records
			exclusiveScan( P.doc_res_count, ndocs, d_res_off, scantmp, stream);
			exclusiveScan( P.doc_item_count, ndocs, d_item_off, scantmp, stream);
			launches += 6;
			readScalars( d_res_off + ndocs, sizeof(uint64_t)); nres = pinned[0];
			readScalars( d_item_off + ndocs, sizeof(uint64_t)); nitems = pinned[0];
			sp_result* d_results = results.as<sp_result>( nres);
			sp_item* d_items = items.as<sp_item>( nitems);
			if (ndocs)
			{
				k_match_compact<<<(ndocs + 7) / 8, 256, 0, stream>>>( P, d_res_off, d_item_off, d_results, d_items);
				SP_CUDA( cudaGetLastError());
				launches += 1;
			}
		}
		else
		{
			// window evaluator: work items per token, count pass, prefix sums, write pass; flagged documents go through
			// the state machine
			EvalParams E; std::memset( &E, 0, sizeof(E));
			E.T = P.T; E.tokens = d_tokens; E.origsegs = d_origsegs; E.tok_off = d_tok_off; E.ndocs = ndocs; E.ntokens = ntokens;
			E.tok_slot = tok_slot.as<uint32_t>( ntokens); E.tok_work = tok_work.as<uint32_t>( ntokens); E.tok_doc = tok_docs.as<uint32_t>( ntokens);
			uint64_t* d_tok_work_off = tok_work_off.as<uint64_t>( ntokens + 1);
			E.tok_work_off = d_tok_work_off;
			E.tok_nres = tok_nres.as<uint32_t>( ntokens); E.tok_nitems = tok_nitems.as<uint32_t>( ntokens);
			uint64_t* d_tok_res_off = tok_res_off.as<uint64_t>( ntokens + 1);
			uint64_t* d_tok_item_off = tok_item_off.as<uint64_t>( ntokens + 1);
			E.tok_res_off = d_tok_res_off; E.tok_item_off = d_tok_item_off;
			E.stats = P.stats; E.flags = flags.as<uint32_t>( ndocs);
			uint32_t* d_flag_list = flag_list.as<uint32_t>( ndocs);
			uint64_t* d_escalars = scalars.as<uint64_t>( 16) + 8;	// [0] results [1] items of evaluator documents [2] flagged documents [3] queued items
			E.queue_count = reinterpret_cast<unsigned long long*>( d_escalars + 3);
			SP_CUDA( cudaMemsetAsync( E.flags, 0, (std::size_t)ndocs * sizeof(uint32_t), stream));
			const uint32_t nblocks = (uint32_t)(((ntokens + EvalRun - 1) / EvalRun + EvalWarpsPerBlock - 1) / EvalWarpsPerBlock);
			if (profile) SP_CUDA( cudaEventRecord( ev_a, stream));
			k_eval_prepare<<<(unsigned int)((ntokens + 255) / 256), 256, 0, stream>>>( E);
			exclusiveScan( E.tok_work, ntokens, d_tok_work_off, scantmp, stream);
			SP_CUDA( cudaGetLastError());
			launches += 4;
			readScalars( d_tok_work_off + ntokens, sizeof(uint64_t));
			const uint64_t nwork = pinned[0];
			E.resbits = resbits.as<uint32_t>( nwork / 32 + 2);
			for (int attempt=0;; ++attempt)
			{
				if (queue_cap < 65536) queue_cap = 65536;
				E.queue = reinterpret_cast<unsigned long long*>( queue.as<uint64_t>( queue_cap)); E.queue_cap = queue_cap;
				SP_CUDA( cudaMemsetAsync( P.stats, 0, (std::size_t)ndocs * sizeof(sp_matcher_stats), stream));
				SP_CUDA( cudaMemsetAsync( d_escalars, 0, 4 * sizeof(uint64_t), stream));
				SP_CUDA( cudaMemsetAsync( E.resbits, 0, (nwork / 32 + 2) * sizeof(uint32_t), stream));
				if (count_by_steps) k_eval_count_step<<<nblocks, EvalThreads, 0, stream>>>( E);
				else
				{
					const uint64_t ntiles = (ntokens + EvalTile - 1) / EvalTile;
					const unsigned int jgrid = (unsigned int)((ntiles + EvalWarpsPerBlock - 1) / EvalWarpsPerBlock);
					// 8 blocks per SM = 64 registers; builds for 10 / 12 blocks (48 / 40 registers, 0.5 / 0.8 KB of spills) measured slower: 26.4 / 27.7 against 24.7 ms per 64 MiB
					k_eval_count_join<8><<<jgrid, EvalThreads, 0, stream>>>( E);
				}
				k_eval_count_full<<<(unsigned int)nsm * 8, 128, 0, stream>>>( E);
				if (profile) SP_CUDA( cudaEventRecord( ev_b, stream));
				exclusiveScan( E.tok_nres, ntokens, d_tok_res_off, scantmp, stream);
				exclusiveScan( E.tok_nitems, ntokens, d_tok_item_off, scantmp, stream);
				k_eval_docs<<<(ndocs + 255) / 256, 256, 0, stream>>>( E, P.doc_res_count, P.doc_item_count, P.status, d_flag_list,
										     reinterpret_cast<uint32_t*>( d_escalars + 2), reinterpret_cast<unsigned long long*>( d_escalars));
				SP_CUDA( cudaGetLastError());
				launches += 9;
				readScalars( d_escalars, 4 * sizeof(uint64_t));
				if (pinned[3] <= queue_cap) break;
				if (attempt) throw std::runtime_error( "work queue of the evaluator keeps overflowing");
				queue_cap = pinned[3] + pinned[3] / 4;		// the flags of this attempt are kept: they can only be set again
			}
			eval_queued += pinned[3];
			if (profile)
			{
				float ms = 0.0f;
				if (cudaEventElapsedTime( &ms, ev_a, ev_b) == cudaSuccess) {eval_count_ms += ms;}
			}
			const uint32_t nflagged = (uint32_t)(pinned[2] & 0xFFFFFFFFu);
			++eval_runs; eval_docs += ndocs; eval_flagged += nflagged; eval_items += nwork;
			if (trace) fprintf( stderr, "[sp_matcher] evaluator: docs %u, tokens %llu, rules %llu (queued for the full fold %llu), results %llu items %llu, flagged docs %u\n", ndocs,
					     (unsigned long long)ntokens, (unsigned long long)nwork, (unsigned long long)pinned[3], (unsigned long long)pinned[0], (unsigned long long)pinned[1], nflagged);
			if (nflagged) runStateMachine( P, d_flag_list, nflagged);
			exclusiveScan( P.doc_res_count, ndocs, d_res_off, scantmp, stream);
			exclusiveScan( P.doc_item_count, ndocs, d_item_off, scantmp, stream);
			launches += 6;
			readScalars( d_res_off + ndocs, sizeof(uint64_t)); nres = pinned[0];
			readScalars( d_item_off + ndocs, sizeof(uint64_t)); nitems = pinned[0];
			E.doc_res_off = d_res_off; E.doc_item_off = d_item_off;
			E.results = results.as<sp_result>( nres); E.items = items.as<sp_item>( nitems);
			if (profile) SP_CUDA( cudaEventRecord( ev_a, stream));
			if (write_blocks == 3) k_eval_write<3><<<nblocks, EvalThreads, 0, stream>>>( E);
			else if (write_blocks == 4) k_eval_write<4><<<nblocks, EvalThreads, 0, stream>>>( E);
			else if (write_blocks == 6) k_eval_write<6><<<nblocks, EvalThreads, 0, stream>>>( E);
			else k_eval_write<5><<<nblocks, EvalThreads, 0, stream>>>( E);
			if (profile) SP_CUDA( cudaEventRecord( ev_b, stream));
			if (nflagged) k_eval_place<<<(nflagged + 7) / 8, 256, 0, stream>>>( d_flag_list, nflagged, P, d_res_off, d_item_off, E.results, E.items);
			SP_CUDA( cudaGetLastError());
			launches += nflagged ? 2 : 1;
			if (profile)
			{
				SP_CUDA( cudaStreamSynchronize( stream));
				float ms = 0.0f;
				if (cudaEventElapsedTime( &ms, ev_a, ev_b) == cudaSuccess) {eval_write_ms += ms;}
			}
		}
		if (ntokens)
		{
			res_per_token = std::max( 0.01, (double)nres / (double)ntokens);
			items_per_token = std::max( 0.01, (double)nitems / (double)ntokens);
		}
		valid = true;
	}

	void run( const sp_lexem* tokens_in, const uint32_t* origsegs_in, const uint64_t* tok_offsets, uint32_t ndocs_, int on_device)
	{
		valid = false;
		if (on_device)
		{
			uint64_t total = 0;
			if (ndocs_) {readScalars( tok_offsets + ndocs_, sizeof(uint64_t)); total = pinned[0];}
			runDevice( tokens_in, origsegs_in, tok_offsets, ndocs_, total);
			return;
		}
		for (uint32_t d=0; d<ndocs_; ++d) if (tok_offsets[d+1] < tok_offsets[d]) throw std::runtime_error( "token offsets are not ascending");
		uint64_t total = ndocs_ ? tok_offsets[ ndocs_] : 0;
		sp_lexem* d_tokens = tokens.as<sp_lexem>( total);
		uint32_t* d_segs = 0;
		uint64_t* d_off = tok_off.as<uint64_t>( (std::size_t)ndocs_ + 1);
		if (total) SP_CUDA( cudaMemcpyAsync( d_tokens, tokens_in, total * sizeof(sp_lexem), cudaMemcpyHostToDevice, stream));
		if (origsegs_in)
		{
			d_segs = origsegs.as<uint32_t>( total);
			if (total) SP_CUDA( cudaMemcpyAsync( d_segs, origsegs_in, total * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
		}
		if (ndocs_) SP_CUDA( cudaMemcpyAsync( d_off, tok_offsets, ((std::size_t)ndocs_ + 1) * sizeof(uint64_t), cudaMemcpyHostToDevice, stream));
		runDevice( d_tokens, d_segs, d_off, ndocs_, total);
	}
};

extern "C" {

sp_matcher_ctx* sp_matcher_ctx_create( sp_matcher* m, void* cuda_stream)
{
	try {m->upload(); return new sp_matcher_ctx( m, (cudaStream_t)cuda_stream);}
	catch (const std::exception& e) {m->lasterror = e.what(); return 0;}
}
void sp_matcher_ctx_destroy( sp_matcher_ctx* ctx) {delete ctx;}
const char* sp_matcher_ctx_last_error( const sp_matcher_ctx* ctx) {return ctx->lasterror.c_str();}

int sp_matcher_run( sp_matcher_ctx* ctx, const sp_lexem* tokens, const uint32_t* origsegs, const uint64_t* tok_offsets, uint32_t ndocs, int on_device)
	{SP_MATCHER_TRY( ctx, ctx->run( tokens, origsegs, tok_offsets, ndocs, on_device))}

int sp_matcher_run_lexer_output( sp_matcher_ctx* ctx, sp_lexer_ctx* lexer_ctx)
{
	SP_MATCHER_TRY( ctx, {
		int64_t nlex = sp_lexer_result_count( lexer_ctx);
		int64_t nd = sp_lexer_ctx_counter( lexer_ctx, "ndocs");
		if (nlex < 0 || nd < 0) throw std::runtime_error( "lexer context holds no finished scan");
		ctx->runDevice( sp_lexer_result_lexems_device( lexer_ctx), 0, sp_lexer_result_offsets_device( lexer_ctx), (uint32_t)nd, (uint64_t)nlex);
	})
}

int64_t sp_matcher_result_count( sp_matcher_ctx* ctx)
{
	if (!ctx->valid) {ctx->lasterror = "no finished run"; return SP_ERR_ARGUMENT;}
	if (cudaStreamSynchronize( ctx->stream) != cudaSuccess) {ctx->lasterror = "kernel failure"; return SP_ERR_CUDA;}
	return (int64_t)ctx->nres;
}
int64_t sp_matcher_item_count( sp_matcher_ctx* ctx)
{
	if (!ctx->valid) {ctx->lasterror = "no finished run"; return SP_ERR_ARGUMENT;}
	return (int64_t)ctx->nitems;
}
const sp_result* sp_matcher_results_device( sp_matcher_ctx* ctx) {return ctx->valid ? (const sp_result*)ctx->results.ptr() : 0;}
const sp_item* sp_matcher_items_device( sp_matcher_ctx* ctx) {return ctx->valid ? (const sp_item*)ctx->items.ptr() : 0;}

int sp_matcher_fetch( sp_matcher_ctx* ctx, sp_result* results, uint64_t* result_offsets, sp_item* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status)
{
	SP_MATCHER_TRY( ctx, {
		if (!ctx->valid) throw std::runtime_error( "no finished run");
		std::size_t nd = ctx->ndocs;
		cudaStream_t st = ctx->stream;
		if (results && ctx->nres) SP_CUDA( cudaMemcpyAsync( results, ctx->results.ptr(), ctx->nres * sizeof(sp_result), cudaMemcpyDeviceToHost, st));
		if (result_offsets) SP_CUDA( cudaMemcpyAsync( result_offsets, ctx->res_off.ptr(), (nd + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		if (items && ctx->nitems) SP_CUDA( cudaMemcpyAsync( items, ctx->items.ptr(), ctx->nitems * sizeof(sp_item), cudaMemcpyDeviceToHost, st));
		if (item_offsets) SP_CUDA( cudaMemcpyAsync( item_offsets, ctx->item_off.ptr(), (nd + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		if (stats && nd) SP_CUDA( cudaMemcpyAsync( stats, ctx->stats.ptr(), nd * sizeof(sp_matcher_stats), cudaMemcpyDeviceToHost, st));
		if (status && nd) SP_CUDA( cudaMemcpyAsync( status, ctx->status.ptr(), nd * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaStreamSynchronize( st));
	})
}

int sp_matcher_fetch_packed( sp_matcher_ctx* ctx, sp_result_packed* results, uint64_t* result_offsets, sp_item_packed* items, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status)
{
	SP_MATCHER_TRY( ctx, {
		if (!ctx->valid) throw std::runtime_error( "no finished run");
		if (!results || !result_offsets || !items || !item_offsets) throw std::runtime_error( "packed fetch needs buffers for results, items and their offsets");
		std::size_t nd = ctx->ndocs;
		cudaStream_t st = ctx->stream;
		sp_result_packed* d_res = ctx->packed_results.as<sp_result_packed>( ctx->nres);
		sp_item_packed* d_items = ctx->packed_items.as<sp_item_packed>( ctx->nitems);
		uint32_t* d_flag = reinterpret_cast<uint32_t*>( ctx->scalars.as<uint64_t>( 32) + 24);
		SP_CUDA( cudaMemsetAsync( d_flag, 0, sizeof(uint32_t), st));
		if (ctx->nres) k_pack_results<<<(unsigned int)((ctx->nres + 255) / 256), 256, 0, st>>>( static_cast<const sp_result*>( ctx->results.ptr()), ctx->nres, d_res, d_flag);
		if (ctx->nitems) k_pack_items<<<(unsigned int)((ctx->nitems + 255) / 256), 256, 0, st>>>( static_cast<const sp_item*>( ctx->items.ptr()), ctx->nitems, d_items, d_flag);
		SP_CUDA( cudaGetLastError());
		ctx->readScalars( d_flag, sizeof(uint32_t));
		if ((uint32_t)ctx->pinned[ 0]) throw UnpackableError();
		if (ctx->nres) SP_CUDA( cudaMemcpyAsync( results, d_res, ctx->nres * sizeof(sp_result_packed), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaMemcpyAsync( result_offsets, ctx->res_off.ptr(), (nd + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		if (ctx->nitems) SP_CUDA( cudaMemcpyAsync( items, d_items, ctx->nitems * sizeof(sp_item_packed), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaMemcpyAsync( item_offsets, ctx->item_off.ptr(), (nd + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		if (stats && nd) SP_CUDA( cudaMemcpyAsync( stats, ctx->stats.ptr(), nd * sizeof(sp_matcher_stats), cudaMemcpyDeviceToHost, st));
		if (status && nd) SP_CUDA( cudaMemcpyAsync( status, ctx->status.ptr(), nd * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaStreamSynchronize( st));
	})
}

void sp_unpack_results( const sp_result_packed* results, const uint64_t* result_offsets, const sp_item_packed* items, const uint64_t* item_offsets, uint32_t ndocs,
			sp_result* out_results, sp_item* out_items)
{
	const uint64_t rb = result_offsets[ 0], ib = item_offsets[ 0];
	for (uint32_t d=0; d<ndocs; ++d)
	{
		uint32_t item_begin = 0;
		for (uint64_t i=result_offsets[ d]-rb, e=result_offsets[ d+1]-rb; i<e; ++i)
		{
			const sp_result_packed p = results[ i];
			sp_result r;
			r.name = p.name_format & 0x3FFFFu; r.format = p.name_format >> 18;
			r.ordpos = p.ordpos; r.ordend = p.ordpos + (p.lens & 0xFFu);
			r.origseg = 0; r.origpos = p.origpos; r.endseg = 0; r.endpos = p.origpos + (p.lens >> 16);
			r.item_begin = item_begin; r.item_count = (p.lens >> 8) & 0xFFu;
			item_begin += r.item_count;
			out_results[ i] = r;
		}
	}
	for (uint64_t i=0, e=item_offsets[ ndocs]-ib; i<e; ++i)
	{
		const sp_item_packed p = items[ i];
		sp_item it;
		it.variable = p.var_lens & 0xFFFu; it.format = 0;
		it.ordpos = p.ordpos; it.ordend = p.ordpos + ((p.var_lens >> 12) & 0xFu);
		it.origseg = 0; it.origpos = p.origpos; it.endseg = 0; it.endpos = p.origpos + (p.var_lens >> 16);
		out_items[ i] = it;
	}
}

int sp_matcher_fetch_range( sp_matcher_ctx* ctx, uint32_t first_doc, uint32_t ndocs, sp_result* results, uint64_t result_capacity, uint64_t* result_offsets,
			    sp_item* items, uint64_t item_capacity, uint64_t* item_offsets, sp_matcher_stats* stats, uint32_t* status)
{
	SP_MATCHER_TRY( ctx, {
		if (!ctx->valid) throw std::runtime_error( "no finished run");
		if ((uint64_t)first_doc + ndocs > ctx->ndocs || !result_offsets || !item_offsets) throw std::runtime_error( "document range out of bounds");
		cudaStream_t st = ctx->stream;
		SP_CUDA( cudaMemcpyAsync( result_offsets, static_cast<const uint64_t*>( ctx->res_off.ptr()) + first_doc, ((std::size_t)ndocs + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaMemcpyAsync( item_offsets, static_cast<const uint64_t*>( ctx->item_off.ptr()) + first_doc, ((std::size_t)ndocs + 1) * sizeof(uint64_t), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaStreamSynchronize( st));
		const uint64_t rb = result_offsets[ 0], rn = result_offsets[ ndocs] - rb, ib = item_offsets[ 0], in = item_offsets[ ndocs] - ib;
		if ((results && rn > result_capacity) || (items && in > item_capacity)) throw std::runtime_error( "result buffers too small for the document range");
		if (results && rn) SP_CUDA( cudaMemcpyAsync( results, static_cast<const sp_result*>( ctx->results.ptr()) + rb, rn * sizeof(sp_result), cudaMemcpyDeviceToHost, st));
		if (items && in) SP_CUDA( cudaMemcpyAsync( items, static_cast<const sp_item*>( ctx->items.ptr()) + ib, in * sizeof(sp_item), cudaMemcpyDeviceToHost, st));
		if (stats && ndocs) SP_CUDA( cudaMemcpyAsync( stats, static_cast<const sp_matcher_stats*>( ctx->stats.ptr()) + first_doc, (std::size_t)ndocs * sizeof(sp_matcher_stats), cudaMemcpyDeviceToHost, st));
		if (status && ndocs) SP_CUDA( cudaMemcpyAsync( status, static_cast<const uint32_t*>( ctx->status.ptr()) + first_doc, (std::size_t)ndocs * sizeof(uint32_t), cudaMemcpyDeviceToHost, st));
		SP_CUDA( cudaStreamSynchronize( st));
	})
}

int64_t sp_matcher_ctx_counter( sp_matcher_ctx* ctx, const char* key)
{
	std::string k( key);
	if (k == "results") return (int64_t)ctx->nres;
	if (k == "items") return (int64_t)ctx->nitems;
	if (k == "tokens") return (int64_t)ctx->ntokens;
	if (k == "ndocs") return (int64_t)ctx->ndocs;
	if (k == "arena_retries") return (int64_t)ctx->retries;
	if (k == "pool_retries") return (int64_t)ctx->pool_retries;
	if (k == "kernel_launches") return (int64_t)ctx->launches;
	if (k == "launches:reset") {ctx->launches = 0; return 0;}
	if (k == "profile:on") {ctx->profile = true; return 1;}
	if (k == "profile:off") {ctx->profile = false; return 0;}
	if (k == "profile:reset") {ctx->kernel_ms = 0.0; ctx->kernel_calls = 0; ctx->eval_count_ms = ctx->eval_write_ms = 0.0; return 0;}
	if (k == "time_us:k_eval_count") return (int64_t)(ctx->eval_count_ms * 1000.0);
	if (k == "time_us:k_eval_write") return (int64_t)(ctx->eval_write_ms * 1000.0);
	if (k == "eval_docs") return (int64_t)ctx->eval_docs;
	if (k == "eval_flagged") return (int64_t)ctx->eval_flagged;
	if (k == "eval_runs") return (int64_t)ctx->eval_runs;
	if (k == "eval_queued") return (int64_t)ctx->eval_queued;
	if (k == "eval_rules") return (int64_t)ctx->eval_items;
	if (k == "eval_capable") return (int64_t)ctx->m->dev.eval_capable;
	if (k == "mode") return ctx->mode;
	if (k == "mode=auto") return ctx->mode = 0;
	if (k == "mode=statemachine") return ctx->mode = 1;
	if (k == "mode=eval") return ctx->mode = 2;
	if (k == "count=step") {ctx->count_by_steps = true; return 1;}
	if (k == "count=join") {ctx->count_by_steps = false; return 0;}
	if (k == "time_us:k_match_documents") return (int64_t)(ctx->kernel_ms * 1000.0);
	if (k == "calls:k_match_documents") return (int64_t)ctx->kernel_calls;
	if (k == "warps_per_sm") return (int64_t)ctx->warps_per_sm;
	if (k.compare( 0, 13, "warps_per_sm=") == 0) {int v = std::atoi( k.c_str() + 13); if (v >= 1) {ctx->warps_per_sm = (uint32_t)v; ctx->warps_fixed = true;} return ctx->warps_per_sm;}
	if (k.compare( 0, 6, "arena=") == 0)
	{
		int n = std::atoi( k.c_str() + 6);
		if (n >= 4)
		{
			SpArenaCaps& c = ctx->cap;
			c.rules = c.triggers = c.datarefs = c.results = c.dispose = (uint32_t)n; c.items = 4u*n; c.follow = n < 64 ? n : n/8; c.heap = n/2 + 1;
		}
		return ctx->cap.rules;
	}
	return -1;
}

} // extern "C"
