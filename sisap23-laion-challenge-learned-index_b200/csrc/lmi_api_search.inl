// lmi_api_search.inl -- navigation + search pipeline (included by lmi_api.cu)

namespace lmi {

// Host -> device copy of a caller buffer on `st`.  Pinned (or registered) memory goes straight to cudaMemcpyAsync;
// PAGEABLE memory -- what numpy hands the drop-in search() -- is first gathered into a pinned staging buffer by a few
// host threads (the driver's own pageable path is a single-threaded staged copy and blocks the caller for ~2 ms at
// the 15 MB of a 10k-query batch).  `pin` is a grow-only pinned buffer owned by the index.
static bool host_ptr_is_pinned(const void* p) {
  cudaPointerAttributes a{};
  if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return a.type == cudaMemoryTypeHost || a.type == cudaMemoryTypeManaged;
}

// A few persistent host threads that copy slices of a buffer (created on first use, shared by all indexes of the
// process; spawning threads per call costs more than the copy of a 15 MB query batch).
class CopyPool {
 public:
  static CopyPool& get() { static CopyPool* p = new CopyPool(); return *p; }     // never destroyed: its helpers outlive main()
  void copy(void* dst, const void* src, size_t bytes) {
    if (bytes < ((size_t)1 << 20) || workers_.empty()) { memcpy(dst, src, bytes); return; }
    std::lock_guard<std::mutex> one_job(call_mu_);          // one copy at a time (indexes on several host threads share the pool)
    std::unique_lock<std::mutex> lk(mu_);
    dst_ = (char*)dst; src_ = (const char*)src; bytes_ = bytes;
    parts_ = (int)workers_.size() + 1;
    next_ = 0; left_ = parts_;
    ++gen_;
    cv_.notify_all();
    lk.unlock();
    drain();
    lk.lock();
    done_.wait(lk, [&] { return left_ == 0; });
  }
 private:
  CopyPool() {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const int n = (int)std::min(7u, hw > 1 ? hw - 1 : 0u);
    for (int i = 0; i < n; ++i) workers_.emplace_back([this] { loop(); });
    for (auto& t : workers_) t.detach();        // process-lifetime helpers (no join at exit: they only wait on cv_)
  }
  void drain() {
    for (;;) {
      int p;
      { std::lock_guard<std::mutex> g(mu_); if (next_ >= parts_) return; p = next_++; }
      const size_t lo = bytes_ * p / parts_ / 64 * 64, hi = (p + 1 == parts_) ? bytes_ : bytes_ * (p + 1) / parts_ / 64 * 64;
      memcpy(dst_ + lo, src_ + lo, hi - lo);
      { std::lock_guard<std::mutex> g(mu_); if (--left_ == 0) done_.notify_all(); }
    }
  }
  void loop() {
    uint64_t seen = 0;
    for (;;) {
      { std::unique_lock<std::mutex> lk(mu_); cv_.wait(lk, [&] { return gen_ != seen; }); seen = gen_; }
      drain();
    }
  }
  std::mutex mu_, call_mu_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> workers_;
  char* dst_ = nullptr; const char* src_ = nullptr; size_t bytes_ = 0;
  int parts_ = 0, next_ = 0, left_ = 0;
  uint64_t gen_ = 0;
};

static void parallel_memcpy(void* dst, const void* src, size_t bytes) { CopyPool::get().copy(dst, src, bytes); }

// cudaFuncAttributeMaxDynamicSharedMemorySize is per function AND per device: it is set before every launch (a few
// microseconds) instead of being cached per process, so a second index on another GPU, or another index with a
// different shape, can never launch with a stale limit.
template <class K>
static cudaError_t opt_in_smem(K kern, size_t smem) {
  return cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
}

static int launch_nav(lmi_index* idx, const float* x_dev, int64_t n, int n_top, int32_t* cls_dev, float* probs_dev,
                      cudaStream_t st) {
  if (n <= 0) return LMI_OK;
  const int64_t grid = (n + kNavTQ - 1) / kNavTQ;
  LMI_CUDA(opt_in_smem(nav_kernel, nav_smem_bytes(idx->layers)));     // per device, and indexes with different MLPs share the function
  nav_kernel<<<(unsigned)grid, kNavThreads, nav_smem_bytes(idx->layers), st>>>(x_dev, n, idx->layers, n_top, cls_dev, probs_dev);
  LMI_CUDA(cudaGetLastError());
  return LMI_OK;
}

template <int KP>
static int launch_rescore(lmi_index* idx, const uint16_t* q16_dev, int nb, cudaStream_t st) {
  const int d = idx->d;
  const int stage_box = idx->tmap_x3_ok ? 1 : 0;      // one 3-D TMA box per pipeline stage (needs d % 256 == 0)
#ifdef LMI_EXPERIMENTS
  const int dbg = idx->dbg_flags;
#else
  const int dbg = 0;
#endif
  bool done = false;
  if constexpr (KP <= 128) {
    if (idx->kernel_now == LMI_RESCORE_TC) {
      done = true;
      // 32 KB stages beside the [k' + 16][128] x 8 B candidate lists: k' = 128 (k = 100) leaves room for two
      constexpr int STAGES = (KP <= 16) ? 5 : (KP <= 64) ? 4 : 2;
      auto kern = rescore_ts_kernel<KP, STAGES>;
      constexpr size_t smem = ts_smem_bytes<KP, STAGES>();
      LMI_CUDA(opt_in_smem(kern, smem));
      kern<<<idx->num_sms & ~1, kTsThreads, smem, st>>>(idx->tmap_x64, idx->tmap_x32, idx->tmap_xpf, idx->tmap_x3_64, q16_dev, idx->slot_pair.as<int32_t>(), nb, d, idx->inv_norm,
                                                   idx->items.as<WorkItem>(), idx->scalars.as<PlanScalars>(),
                                                   idx->qthr.as<uint32_t>(), idx->plists.as<Cand>(), stage_box, dbg);
      idx->rescore_name = "rescore_ts_kernel";
    }
  }
  if constexpr (KP <= 128) {
    if (idx->kernel_now == LMI_RESCORE_TC_PAIR) {
      done = true;
      auto launch = [&](auto kern, size_t smem, int threads) -> cudaError_t {
        cudaError_t e = opt_in_smem(kern, smem);
        if (e != cudaSuccess) return e;
        kern<<<idx->num_sms & ~1, threads, smem, st>>>(idx->tmap_x32, idx->tmap_x3_32, q16_dev, idx->slot_pair.as<int32_t>(), nb, d, idx->inv_norm,
                                                       idx->items.as<WorkItem>(), idx->scalars.as<PlanScalars>(),
                                                       idx->qthr.as<uint32_t>(), idx->plists.as<Cand>(),
                                                       stage_box | (idx->l2_hint ? 2 : 0) | (idx->lockstep_now ? 4 : 0), dbg, idx->progress.as<int32_t>());
        return cudaSuccess;
      };
      // 16 KB stages per CTA (half tiles) beside SETS x [k' + 16][128] x 8 B candidate lists.  Two epilogue warp
      // sets (one per accumulator) while their lists fit: k' <= 32; option "epi_sets" = 1 forces one set.
      bool two_sets = false;
      if constexpr (KP <= 32) {
        if (idx->epi_sets != 1) {
          two_sets = true;
          constexpr int STAGES = (KP <= 16) ? 9 : 7;
          static_assert(ts2_smem_bytes<KP, STAGES, 2>() <= 227 * 1024, "rescore_ts2 (2 sets): shared memory budget");
          LMI_CUDA(launch(rescore_ts2_kernel<KP, STAGES, 2>, ts2_smem_bytes<KP, STAGES, 2>(), ts2_threads(2)));
        }
      }
      if (!two_sets) {
        constexpr int STAGES = (KP <= 32) ? 10 : (KP <= 64) ? 8 : 4;
        static_assert(ts2_smem_bytes<KP, STAGES, 1>() <= 227 * 1024, "rescore_ts2: shared memory budget");
        LMI_CUDA(launch(rescore_ts2_kernel<KP, STAGES, 1>, ts2_smem_bytes<KP, STAGES, 1>(), ts2_threads(1)));
      }
      idx->rescore_name = "rescore_ts2_kernel";
    }
  }
  if constexpr (KP <= 64) {
    if (idx->kernel_now == LMI_RESCORE_TC_SS) {
      done = true;
      constexpr int STAGES = (KP <= 16) ? 4 : 3;
      auto kern = rescore_tc_kernel<KP, STAGES>;
      constexpr size_t smem = tc_smem_bytes<KP, STAGES>();
      LMI_CUDA(opt_in_smem(kern, smem));
      kern<<<idx->num_sms, kTcThreads, smem, st>>>(idx->tmap_q, idx->tmap_x, d, idx->inv_norm, idx->items.as<WorkItem>(),
                                                   idx->scalars.as<PlanScalars>(), idx->plists.as<Cand>());
      idx->rescore_name = "rescore_tc_kernel";
    }
  }
  if (!done) {
    auto kern = rescore_simt_kernel<KP>;
    constexpr size_t smem = simt_smem_bytes<KP>();
    LMI_CUDA(opt_in_smem(kern, smem));
    kern<<<idx->num_sms * 2, kTileQ, smem, st>>>(q16_dev, idx->slot_pair.as<int32_t>(), nb, d, idx->x, idx->inv_norm,
                                                 idx->items.as<WorkItem>(), idx->scalars.as<PlanScalars>(),
                                                 idx->plists.as<Cand>());
    idx->rescore_name = "rescore_simt_kernel";
  }
  LMI_CUDA(cudaGetLastError());
  return LMI_OK;
}

// Objects per work item.  Small chunks balance the persistent grid; large chunks let a row's candidate
// list live longer (fewer insertions: a list that has seen n objects admits ~k'/n of the next ones, and
// every chunk starts a new list) and leave fewer partial lists to merge.  Measured on 10M / nb = 4
// (profiles/r1_chunk_sweep.txt): k = 10 is best around 32k objects (1,200 work-list entries on 74
// clusters), k = 100 -- where insertions dominate -- at 64k.  So: aim at ~13 (k' <= 32) or ~8 entries per
// cluster, from the expected number of query tiles on this shard.
static int choose_chunk(const lmi_index* idx, int64_t nq, int nb, int KP) {
  if (idx->chunk_objects > 0) return idx->chunk_objects;
  const double clusters = std::max(1, idx->num_sms / 2);
  const double target = clusters * (KP >= 64 ? 8.0 : 13.0);
  const double qtiles = (double)nq * nb * idx->pair_fraction / kTileQ + 0.5 * (double)idx->n_nonempty;
  const double avg_bucket = (double)idx->n_local / (double)std::max<int64_t>(idx->n_nonempty, 1);
  double c = qtiles * avg_bucket / target;
  // Floor: an item must be long enough to amortise what it pays once (128 query rows into tensor memory, pipeline fill,
  // one partial list per row for K4).  Measured per rank of an 8-way bucket shard (profiles/r2c_shard_emulate_n8_buckets.json):
  // 4,096-object chunks 1.33 ms per partial search (K3 0.93 + finalize 0.22), 8,192: 1.16, 16,384: 1.15 (0.83 + 0.13),
  // 32,768: 1.27.  For k' >= 64 the K4 selection runs on registers up to 1,536 entries per query: no more than
  // 1536 / (nb * k') chunks per bucket where the buckets allow it.
  const double floor_c = (KP >= 64) ? std::max(12288.0, (double)idx->h_sizes_sorted[0] * nb * KP / 1536.0) : 12288.0;
  c = std::min(std::max(c, floor_c), (double)(1 << 22));
  return (int)(((int64_t)c + 255) / 256 * 256);
}

// The whole hot path on `st`.  Exactly one of q_nav_dev / cls64_dev is used to obtain the bucket
// list: navigation (search) or caller-supplied categories (search_rank, nb = 1).
static int run_search(lmi_index* idx, const float* q_nav_dev, const int64_t* cls64_dev, const uint16_t* q16_dev,
                      int64_t nq, int nb, int k, const double* thr_dev, int mode, double* out_d, uint32_t* out_i,
                      Partial* out_p, cudaStream_t st, cudaEvent_t q16_ready = nullptr, const int32_t* cls32_dev = nullptr) {
  if (!idx->has_store) return set_error(LMI_ERR_STATE, "bucket store not built (lmi_store_* / lmi_set_buckets)");
  if (q_nav_dev && !idx->has_mlp) return set_error(LMI_ERR_STATE, "Model is not trained, call `build` first.");   // LearnedIndex.py:60
  if (nq <= 0 || nq > (1 << 24)) return set_error(LMI_ERR_ARG, "nq out of range");
  if (nb < 1 || nb > idx->C) return set_error(LMI_ERR_ARG, "n_buckets must be in [1, n_categories]");
  if (k < 1) return set_error(LMI_ERR_ARG, "k must be >= 1");
  const int KP = next_kprime(k + idx->kprime_extra);
  if (KP < 0) return set_error(LMI_ERR_ARG, "k too large (k + kprime_extra <= 128)");
  if (nq * nb >= (1LL << 31)) return set_error(LMI_ERR_ARG, "nq * n_buckets too large");
  LMI_CUDA(cudaSetDevice(idx->device));
  const int C = idx->C, d = idx->d;
  const int64_t n_pairs = nq * nb;

  // ---- workspace ----
  LMI_CUDA(idx->top_cls.reserve((size_t)n_pairs * 4));
  LMI_CUDA(idx->m_b.reserve((size_t)C * 4));
  LMI_CUDA(idx->qoff.reserve((size_t)(C + 1) * 4));
  LMI_CUDA(idx->chunks_b.reserve((size_t)C * 4));
  LMI_CUDA(idx->plist_base.reserve((size_t)C * 8));
  LMI_CUDA(idx->slot_pair.reserve((size_t)n_pairs * 4));
  LMI_CUDA(idx->pair_slot.reserve((size_t)n_pairs * 4));
  LMI_CUDA(idx->scalars.reserve(sizeof(PlanScalars)));
  LMI_CUDA(idx->uncert.reserve((size_t)nq * 4));
  const int64_t qtiles_max = (nq + kTileQ - 1) / kTileQ;
  const int chunk = choose_chunk(idx, nq, nb, KP);
  idx->chunk_now = chunk;
  int64_t total_chunks = 0, max_chunks = 0, top_nb_chunks = 0;         // h_sizes_sorted is descending
  for (int c = 0; c < C; ++c) {
    const int64_t ch = (idx->h_sizes_sorted[c] + chunk - 1) / chunk;
    total_chunks += ch;
    if (c == 0) max_chunks = ch;
    if (c < nb) top_nb_chunks += ch;
  }
  // entries: at most qtiles_b * chunks_b + 1 (dummy) per bucket
  // (pair mode of the CTA-pair kernel: an odd last tile becomes two entries -> one more per bucket and chunk)
  // (item_order 2 lists a three-tile bucket as three half pairs: up to 2 entries per tile)
  const int64_t items_a = (2 * qtiles_max + 1) * total_chunks + C;
  const int64_t items_b = (2 * n_pairs / kTileQ + 2 * C) * std::max<int64_t>(max_chunks, 1) + C;
  const int64_t max_items = std::max<int64_t>(1, std::min(items_a, items_b));
  if (max_items >= (1LL << 31)) return set_error(LMI_ERR_NOMEM, "work list too large; raise chunk_objects");
  LMI_CUDA(idx->items.reserve((size_t)max_items * sizeof(WorkItem)));
  LMI_CUDA(idx->progress.reserve((size_t)(max_items / 2 + 2) * sizeof(int32_t)));
  const int64_t max_plists = std::max<int64_t>(1, nq * top_nb_chunks);
  LMI_CUDA(idx->plists.reserve((size_t)max_plists * KP * sizeof(Cand)));
  // an empty shard has no tensor map and no work items: the (idle) CUDA-core kernel is launched instead
  int kernel = LMI_RESCORE_SIMT;
  if (idx->rescore != LMI_RESCORE_SIMT && idx->tmap_x_ok) {
    if (idx->rescore == LMI_RESCORE_TC && d <= 2 * kTsACols) kernel = LMI_RESCORE_TC;        // k' <= 128
    else if (idx->rescore == LMI_RESCORE_TC_PAIR && d <= 2 * kTsACols) kernel = LMI_RESCORE_TC_PAIR;
    else if (KP <= 64) kernel = LMI_RESCORE_TC_SS;
  }
  idx->kernel_now = kernel;
  // lock-step followers (option "lockstep": 1 on, 0 off, -1 automatic = on a sharded index) need the bucket-major list
  idx->lockstep_now = kernel == LMI_RESCORE_TC_PAIR && (idx->lockstep == 1 || (idx->lockstep < 0 && idx->world > 1));
  if (idx->lockstep_now) LMI_CUDA(cudaMemsetAsync(idx->progress.p, 0, (size_t)(max_items / 2 + 2) * sizeof(int32_t), st));
  const bool use_tc = (kernel == LMI_RESCORE_TC_SS);       // SS kernel: needs the bucket-ordered query copy
  if (kernel == LMI_RESCORE_TC || kernel == LMI_RESCORE_TC_PAIR) {
    LMI_CUDA(idx->qthr.reserve((size_t)nq * 4));
    LMI_CUDA(cudaMemsetAsync(idx->qthr.p, 0, (size_t)nq * 4, st));
  }
  if (use_tc) {
    // +128 rows of slack: the last tile's TMA box may start up to 127 rows before the end
    LMI_CUDA(idx->qg.reserve((size_t)(n_pairs + kTileQ) * d * 2));
    const size_t rows_cap = idx->qg.cap / ((size_t)d * 2);
    if (idx->tmap_q_ptr != idx->qg.p || idx->tmap_q_rows != rows_cap) {
      int rc = make_tmap(&idx->tmap_q, idx->qg.p, rows_cap, d, kTileQ);
      if (rc != LMI_OK) return rc;
      idx->tmap_q_ptr = idx->qg.p; idx->tmap_q_rows = rows_cap;
    }
  }

  const bool timing = idx->timing != 0;
  if (timing) LMI_CUDA(cudaEventRecord(idx->ev[0], st));
  // ---- K1 navigation ----
  int32_t* top = idx->top_cls.as<int32_t>();
  if (cls32_dev) {
    top = const_cast<int32_t*>(cls32_dev);               // bucket lists navigated elsewhere (multi-GPU: every rank navigates a slice)
  } else if (q_nav_dev) {
    int rc = launch_nav(idx, q_nav_dev, nq, nb, top, nullptr, st);
    if (rc != LMI_OK) return rc;
  } else {
    cast_i64_to_i32_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, st>>>(cls64_dev, n_pairs, top);
  }
  if (timing) LMI_CUDA(cudaEventRecord(idx->ev[1], st));
  // ---- K2 plan ----
  LMI_CUDA(cudaMemsetAsync(idx->m_b.p, 0, (size_t)C * 4, st));
  plan_count_kernel<<<idx->num_sms * 2, 256, 0, st>>>(top, n_pairs, C, idx->m_b.as<int32_t>());
  plan_build_kernel<<<1, 1024, 0, st>>>(idx->m_b.as<int32_t>(), idx->bucket_ptr.as<int64_t>(), C, chunk,
                                        idx->qoff.as<int32_t>(), idx->chunks_b.as<int32_t>(), idx->plist_base.as<int64_t>(),
                                        idx->items.as<WorkItem>(), (int)max_items, idx->scalars.as<PlanScalars>(),
                                        kernel == LMI_RESCORE_TC_PAIR ? ((idx->item_order == 1 || idx->lockstep_now) ? 2 : idx->item_order == 2 ? 3 : 1) : 0);
  plan_scatter_kernel<<<C, kScatterThreads, 0, st>>>(top, n_pairs, C, idx->qoff.as<int32_t>(), idx->bucket_ptr.as<int64_t>(),
                                         idx->slot_pair.as<int32_t>(), idx->pair_slot.as<int32_t>());
  plan_mark_invalid_kernel<<<(unsigned)((n_pairs + 255) / 256), 256, 0, st>>>(top, n_pairs, C, idx->pair_slot.as<int32_t>());
  // the fp16 queries are first read here (host-buffer entry point: they were uploaded on the copy stream
  // while K1/K2 ran)
  if (q16_ready) LMI_CUDA(cudaStreamWaitEvent(st, q16_ready, 0));
  if (use_tc)   // A operand of the tensor-core GEMM: query rows in bucket order
    gather_queries_kernel<<<idx->num_sms * 4, 256, 0, st>>>(q16_dev, idx->slot_pair.as<int32_t>(),
                                                            idx->scalars.as<PlanScalars>(), nb, d, idx->qg.as<uint16_t>());
  LMI_CUDA(cudaGetLastError());
  cudaEvent_t* ring = idx->ev_k3[idx->n_since_stats % lmi_index::kEvRing];
  if (timing) { LMI_CUDA(cudaEventRecord(idx->ev[2], st)); LMI_CUDA(cudaEventRecord(ring[0], st)); }
  // ---- K3 rescoring ----
  int rc = LMI_OK;
  switch (KP) {
    case 16: rc = launch_rescore<16>(idx, q16_dev, nb, st); break;
    case 32: rc = launch_rescore<32>(idx, q16_dev, nb, st); break;
    case 64: rc = launch_rescore<64>(idx, q16_dev, nb, st); break;
    default: rc = launch_rescore<128>(idx, q16_dev, nb, st); break;
  }
  if (rc != LMI_OK) return rc;
  if (timing) { LMI_CUDA(cudaEventRecord(ring[1], st)); LMI_CUDA(cudaEventRecord(idx->ev[3], st)); ++idx->n_since_stats; }
  // ---- K4 + K5 finalize ----
  auto fin = (KP >= 64) ? finalize_kernel<true> : finalize_kernel<false>;
  fin<<<(unsigned)((nq + kFinWarps - 1) / kFinWarps), kFinWarps * 32, 0, st>>>(
      nq, nb, k, KP, d, top, idx->pair_slot.as<int32_t>(), idx->qoff.as<int32_t>(), idx->chunks_b.as<int32_t>(),
      idx->plist_base.as<int64_t>(), idx->plists.as<Cand>(), q16_dev, idx->x, idx->ids, thr_dev,
      idx->scalars.as<PlanScalars>(), idx->uncert.as<int32_t>(), mode, out_d, out_i, out_p);
  // exactness fallback: queries whose certificate failed are re-derived in float64 from every probed object
  // (exits at once when the list is empty, which is the usual case)
  exact_fixup_kernel<<<idx->num_sms, kFixWarps * 32, 0, st>>>(
      nq, nb, k, d, top, idx->pair_slot.as<int32_t>(), idx->bucket_ptr.as<int64_t>(), q16_dev, idx->x, idx->ids, thr_dev,
      idx->scalars.as<PlanScalars>(), idx->uncert.as<int32_t>(), mode, out_d, out_i, out_p);
  LMI_CUDA(cudaGetLastError());
  if (timing) LMI_CUDA(cudaEventRecord(idx->ev[4], st));
  LMI_CUDA(cudaMemcpyAsync(idx->h_scalars, idx->scalars.p, sizeof(PlanScalars), cudaMemcpyDeviceToHost, st));
  idx->stats.nq = (int32_t)nq; idx->stats.n_buckets = nb; idx->stats.k = k; idx->stats.kprime = KP;
  idx->stats.n_objects_local = idx->n_local;
  idx->stats.chunk_objects = chunk;
  idx->last_stream = st;
  idx->stats_pending = true;
  return LMI_OK;
}

// results -> caller buffers (through pinned staging when the caller's are pageable), synchronous on return
static int copy_results_out(lmi_index* idx, int64_t nq, int k, double* out_dists, uint32_t* out_ids, cudaStream_t st) {
  const size_t b_d = (size_t)nq * k * 8, b_i = (size_t)nq * k * 4;
  if (host_ptr_is_pinned(out_dists) && host_ptr_is_pinned(out_ids)) {
    LMI_CUDA(cudaMemcpyAsync(out_dists, idx->ws_out_d.p, b_d, cudaMemcpyDeviceToHost, st));
    LMI_CUDA(cudaMemcpyAsync(out_ids, idx->ws_out_i.p, b_i, cudaMemcpyDeviceToHost, st));
    LMI_CUDA(cudaStreamSynchronize(st));
    return LMI_OK;
  }
  LMI_CUDA(idx->pin_out.reserve(b_d + b_i));
  LMI_CUDA(cudaMemcpyAsync(idx->pin_out.p, idx->ws_out_d.p, b_d, cudaMemcpyDeviceToHost, st));
  LMI_CUDA(cudaMemcpyAsync((char*)idx->pin_out.p + b_d, idx->ws_out_i.p, b_i, cudaMemcpyDeviceToHost, st));
  LMI_CUDA(cudaStreamSynchronize(st));
  memcpy(out_dists, idx->pin_out.p, b_d);
  memcpy(out_ids, (char*)idx->pin_out.p + b_d, b_i);
  return LMI_OK;
}

}  // namespace lmi

extern "C" {

int lmi_navigate(lmi_index* idx, const float* q_nav, int64_t nq, int n_top, int64_t* out_classes, float* out_probs) {
  if (!idx || !q_nav || !out_classes) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_mlp) return set_error(LMI_ERR_STATE, "Model is not trained, call `build` first.");
  if (n_top < 1 || n_top > idx->C) return set_error(LMI_ERR_ARG, "n_top must be in [1, n_categories]");
  if (nq < 0) return set_error(LMI_ERR_ARG, "nq < 0");
  LMI_CUDA(cudaSetDevice(idx->device));
  cudaStream_t st = idx->stream;
  const int64_t step = 1 << 20;
  std::vector<int32_t> h_cls;
  for (int64_t lo = 0; lo < nq; lo += step) {
    const int64_t cnt = std::min(step, nq - lo);
    LMI_CUDA(idx->nav_x.reserve((size_t)cnt * idx->d_nav * 4));
    LMI_CUDA(idx->nav_cls.reserve((size_t)cnt * n_top * 4));
    if (out_probs) LMI_CUDA(idx->nav_probs.reserve((size_t)cnt * n_top * 4));
    LMI_CUDA(cudaMemcpyAsync(idx->nav_x.p, q_nav + (size_t)lo * idx->d_nav, (size_t)cnt * idx->d_nav * 4, cudaMemcpyHostToDevice, st));
    int rc = launch_nav(idx, idx->nav_x.as<float>(), cnt, n_top, idx->nav_cls.as<int32_t>(),
                        out_probs ? idx->nav_probs.as<float>() : nullptr, st);
    if (rc != LMI_OK) return rc;
    h_cls.resize((size_t)cnt * n_top);
    LMI_CUDA(cudaMemcpyAsync(h_cls.data(), idx->nav_cls.p, h_cls.size() * 4, cudaMemcpyDeviceToHost, st));
    if (out_probs)
      LMI_CUDA(cudaMemcpyAsync(out_probs + (size_t)lo * n_top, idx->nav_probs.p, (size_t)cnt * n_top * 4, cudaMemcpyDeviceToHost, st));
    LMI_CUDA(cudaStreamSynchronize(st));
    for (size_t i = 0; i < h_cls.size(); ++i) out_classes[(size_t)lo * n_top + i] = h_cls[i];
  }
  return LMI_OK;
}

int lmi_predict(lmi_index* idx, const float* x_nav, int64_t n, int64_t* out_category) {
  return lmi_navigate(idx, x_nav, n, 1, out_category, nullptr);
}

int lmi_search_device(lmi_index* idx, const float* q_nav_dev, const uint16_t* q_f16_dev, int64_t nq, int n_buckets, int k,
                      double* out_dists_dev, uint32_t* out_ids_dev, void* stream) {
  if (!idx || !q_nav_dev || !q_f16_dev || !out_dists_dev || !out_ids_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  return run_search(idx, q_nav_dev, nullptr, q_f16_dev, nq, n_buckets, k, nullptr, 0, out_dists_dev, out_ids_dev, nullptr,
                    reinterpret_cast<cudaStream_t>(stream));
}

int lmi_search(lmi_index* idx, const float* q_nav, const uint16_t* q_f16, int64_t nq, int n_buckets, int k,
               double* out_dists, uint32_t* out_ids) {
  if (!idx || !q_nav || !q_f16 || !out_dists || !out_ids) return set_error(LMI_ERR_ARG, "NULL argument");
  if (nq <= 0) return set_error(LMI_ERR_ARG, "nq must be positive");
  if (k < 1) return set_error(LMI_ERR_ARG, "k must be >= 1");
  LMI_CUDA(cudaSetDevice(idx->device));
  cudaStream_t st = idx->stream;
  LMI_CUDA(idx->ws_qnav.reserve((size_t)nq * idx->d_nav * 4));
  LMI_CUDA(idx->ws_q16.reserve((size_t)nq * idx->d * 2));
  LMI_CUDA(idx->ws_out_d.reserve((size_t)nq * k * 8));
  LMI_CUDA(idx->ws_out_i.reserve((size_t)nq * k * 4));
  // navigation needs only the pca96 queries: the (4x larger) fp16 upload overlaps K1/K2 on a second stream.
  // (ws_q16 is free: the previous call on this index synchronised its stream before returning)
  const size_t b_nav = (size_t)nq * idx->d_nav * 4, b_q16 = (size_t)nq * idx->d * 2;
  const void *src_nav = q_nav, *src_q16 = q_f16;
  if (!host_ptr_is_pinned(q_nav) || !host_ptr_is_pinned(q_f16)) {       // pageable caller buffers (numpy): see parallel_memcpy
    LMI_CUDA(idx->pin_in.reserve(b_nav + b_q16));
    parallel_memcpy(idx->pin_in.p, q_nav, b_nav);
    src_nav = idx->pin_in.p;
  }
  LMI_CUDA(cudaMemcpyAsync(idx->ws_qnav.p, src_nav, b_nav, cudaMemcpyHostToDevice, st));
  if (src_nav != q_nav) {
    // navigation is already running while the 4x larger block is gathered; 4 pieces so the H2D copy of one piece
    // overlaps the host-side gather of the next
    char* stage = (char*)idx->pin_in.p + b_nav;
    const int n_piece = b_q16 >= ((size_t)4 << 20) ? 4 : 1;
    for (int c = 0; c < n_piece; ++c) {
      const size_t lo = b_q16 * c / n_piece / 256 * 256, hi = (c + 1 == n_piece) ? b_q16 : b_q16 * (c + 1) / n_piece / 256 * 256;
      parallel_memcpy(stage + lo, (const char*)q_f16 + lo, hi - lo);
      LMI_CUDA(cudaMemcpyAsync((char*)idx->ws_q16.p + lo, stage + lo, hi - lo, cudaMemcpyHostToDevice, idx->copy_stream));
    }
  } else {
    LMI_CUDA(cudaMemcpyAsync(idx->ws_q16.p, src_q16, b_q16, cudaMemcpyHostToDevice, idx->copy_stream));
  }
  LMI_CUDA(cudaEventRecord(idx->ev_q16, idx->copy_stream));
  int rc = run_search(idx, idx->ws_qnav.as<float>(), nullptr, idx->ws_q16.as<uint16_t>(), nq, n_buckets, k, nullptr, 0,
                      idx->ws_out_d.as<double>(), idx->ws_out_i.as<uint32_t>(), nullptr, st, idx->ev_q16);
  if (rc != LMI_OK) { cudaStreamSynchronize(idx->copy_stream); return rc; }
  return copy_results_out(idx, nq, k, out_dists, out_ids, st);
}

int lmi_search_rank(lmi_index* idx, const uint16_t* q_f16, int64_t nq, const int64_t* category_per_query, int k,
                    const double* threshold, double* out_dists, uint32_t* out_ids) {
  if (!idx || !q_f16 || !category_per_query || !out_dists || !out_ids) return set_error(LMI_ERR_ARG, "NULL argument");
  if (nq <= 0) return set_error(LMI_ERR_ARG, "nq must be positive");
  if (k < 1) return set_error(LMI_ERR_ARG, "k must be >= 1");
  LMI_CUDA(cudaSetDevice(idx->device));
  cudaStream_t st = idx->stream;
  LMI_CUDA(idx->ws_cls64.reserve((size_t)nq * 8));
  LMI_CUDA(idx->ws_q16.reserve((size_t)nq * idx->d * 2));
  LMI_CUDA(idx->ws_out_d.reserve((size_t)nq * k * 8));
  LMI_CUDA(idx->ws_out_i.reserve((size_t)nq * k * 4));
  LMI_CUDA(cudaMemcpyAsync(idx->ws_cls64.p, category_per_query, (size_t)nq * 8, cudaMemcpyHostToDevice, st));
  LMI_CUDA(cudaMemcpyAsync(idx->ws_q16.p, q_f16, (size_t)nq * idx->d * 2, cudaMemcpyHostToDevice, st));
  const double* thr_dev = nullptr;
  if (threshold) {
    LMI_CUDA(idx->ws_thr.reserve((size_t)nq * 8));
    LMI_CUDA(cudaMemcpyAsync(idx->ws_thr.p, threshold, (size_t)nq * 8, cudaMemcpyHostToDevice, st));
    thr_dev = idx->ws_thr.as<double>();
  }
  int rc = run_search(idx, nullptr, idx->ws_cls64.as<int64_t>(), idx->ws_q16.as<uint16_t>(), nq, 1, k, thr_dev, 0,
                      idx->ws_out_d.as<double>(), idx->ws_out_i.as<uint32_t>(), nullptr, st);
  if (rc != LMI_OK) return rc;
  return copy_results_out(idx, nq, k, out_dists, out_ids, st);
}

int64_t lmi_partial_bytes(int64_t nq, int k) { return nq * (int64_t)k * (int64_t)sizeof(Partial); }

int lmi_search_partial_device(lmi_index* idx, const float* q_nav_dev, const uint16_t* q_f16_dev, int64_t nq, int n_buckets,
                              int k, void* partial_dev, void* stream) {
  if (!idx || !q_nav_dev || !q_f16_dev || !partial_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  return run_search(idx, q_nav_dev, nullptr, q_f16_dev, nq, n_buckets, k, nullptr, 1, nullptr, nullptr,
                    reinterpret_cast<Partial*>(partial_dev), reinterpret_cast<cudaStream_t>(stream));
}

int lmi_navigate_device(lmi_index* idx, const float* q_nav_dev, int64_t nq, int n_top, int32_t* out_classes_dev, void* stream) {
  if (!idx || !q_nav_dev || !out_classes_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_mlp) return set_error(LMI_ERR_STATE, "Model is not trained, call `build` first.");
  if (n_top < 1 || n_top > idx->C) return set_error(LMI_ERR_ARG, "n_top must be in [1, n_categories]");
  if (nq < 0) return set_error(LMI_ERR_ARG, "nq < 0");
  LMI_CUDA(cudaSetDevice(idx->device));
  return launch_nav(idx, q_nav_dev, nq, n_top, out_classes_dev, nullptr, reinterpret_cast<cudaStream_t>(stream));
}

int lmi_search_partial_classes_device(lmi_index* idx, const int32_t* classes_dev, const uint16_t* q_f16_dev, int64_t nq,
                                      int n_buckets, int k, void* partial_dev, void* stream) {
  if (!idx || !classes_dev || !q_f16_dev || !partial_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  return run_search(idx, nullptr, nullptr, q_f16_dev, nq, n_buckets, k, nullptr, 1, nullptr, nullptr,
                    reinterpret_cast<Partial*>(partial_dev), reinterpret_cast<cudaStream_t>(stream), nullptr, classes_dev);
}

int lmi_merge_partials_device(lmi_index* idx, const void* gathered_dev, int world, int64_t nq, int k, double* out_dists_dev,
                              uint32_t* out_ids_dev, void* stream) {
  if (!idx || !gathered_dev || !out_dists_dev || !out_ids_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  if (world < 1 || nq <= 0 || k < 1) return set_error(LMI_ERR_ARG, "bad world/nq/k");
  LMI_CUDA(cudaSetDevice(idx->device));
  auto merge = ((int64_t)world * k <= kMergeSmemE) ? merge_partials_kernel<true> : merge_partials_kernel<false>;
  merge<<<(unsigned)((nq + kFinWarps - 1) / kFinWarps), kFinWarps * 32, 0, reinterpret_cast<cudaStream_t>(stream)>>>(
      reinterpret_cast<const Partial*>(gathered_dev), world, nq, k, out_dists_dev, out_ids_dev);
  LMI_CUDA(cudaGetLastError());
  return LMI_OK;
}

int lmi_get_stats(lmi_index* idx, lmi_stats* out) {
  if (!idx || !out) return set_error(LMI_ERR_ARG, "NULL argument");
  LMI_CUDA(cudaSetDevice(idx->device));
  if (idx->stats_pending) {
    LMI_CUDA(cudaStreamSynchronize(idx->last_stream));
    const PlanScalars& s = *idx->h_scalars;
    idx->stats.n_work_items = s.n_items;
    idx->stats.pairs = s.pairs;
    idx->stats.bucket_bytes = s.bucket_objs * idx->d * 2;
    idx->stats.n_partial_lists = s.n_plists;
    idx->stats.padded_pairs = s.padded_pairs;
    idx->stats.k3_sm_mhz = s.k3_ns ? (float)((double)s.k3_clk * 1e3 / (double)s.k3_ns) : 0.f;
    idx->stats.n_uncertified = (int64_t)s.n_uncertified;
    double e; unsigned long long bits = s.max_err_bits; memcpy(&e, &bits, 8);
    idx->stats.max_abs_score_err = e;
    if (idx->timing) {
      float ms[4] = {0, 0, 0, 0};
      for (int i = 0; i < 4; ++i) cudaEventElapsedTime(&ms[i], idx->ev[i], idx->ev[i + 1]);
      idx->stats.ms_navigate = ms[0]; idx->stats.ms_plan = ms[1]; idx->stats.ms_rescore = ms[2]; idx->stats.ms_finalize = ms[3];
      cudaEventElapsedTime(&idx->stats.ms_total, idx->ev[0], idx->ev[4]);
      // average rescoring-kernel duration over every search since the previous lmi_get_stats (ring-limited)
      const int64_t n = std::min<int64_t>(idx->n_since_stats, lmi_index::kEvRing);
      double sum = 0;
      for (int64_t i = 0; i < n; ++i) {
        float t = 0;
        cudaEventElapsedTime(&t, idx->ev_k3[(idx->n_since_stats - 1 - i) % lmi_index::kEvRing][0],
                             idx->ev_k3[(idx->n_since_stats - 1 - i) % lmi_index::kEvRing][1]);
        sum += t;
      }
      idx->stats.ms_rescore_avg = n ? (float)(sum / n) : 0.f;
      idx->stats.n_rescore_timed = (int32_t)n;
    }
    idx->n_since_stats = 0;
    idx->stats_pending = false;
  }
  *out = idx->stats;
  return LMI_OK;
}

const char* lmi_rescore_kernel_name(lmi_index* idx) { return idx ? idx->rescore_name.c_str() : ""; }

}  // extern "C"
