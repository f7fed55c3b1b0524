// lmi_api_store.inl -- bucket store build (included by lmi_api.cu)

namespace lmi {

// longest-processing-time assignment of buckets to ranks; expected work of a bucket ~ n_b * m_b
// with m_b ~ n_b (queries follow the data distribution), so the weight is n_b^2.
static std::vector<int32_t> lpt_owners(const std::vector<int64_t>& sizes, int world) {
  const int C = (int)sizes.size();
  std::vector<int> order(C);
  for (int i = 0; i < C; ++i) order[i] = i;
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return sizes[a] > sizes[b]; });
  std::vector<double> load(world, 0.0);
  std::vector<int32_t> owner(C, 0);
  for (int c : order) {
    int best = 0;
    for (int r = 1; r < world; ++r) if (load[r] < load[best]) best = r;
    owner[c] = best;
    load[best] += (double)sizes[c] * (double)sizes[c];
  }
  return owner;
}

}  // namespace lmi

extern "C" {

int lmi_store_begin(lmi_index* idx, const int64_t* pred_categories, int64_t n_total, int64_t id_base, int on_device) {
  if (!idx || !pred_categories) return set_error(LMI_ERR_ARG, "NULL argument");
  if (n_total <= 0 || n_total >= (1LL << 31)) return set_error(LMI_ERR_ARG, "n_total must be in [1, 2^31)");
  if (id_base < 0 || id_base + n_total > 0xffffffffLL) return set_error(LMI_ERR_ARG, "ids do not fit uint32 (reference: LearnedIndex.py:137)");
  LMI_CUDA(cudaSetDevice(idx->device));
  const int C = idx->C;
  if (idx->x) { cudaFree(idx->x); idx->x = nullptr; }
  if (idx->inv_norm) { cudaFree(idx->inv_norm); idx->inv_norm = nullptr; }
  if (idx->ids) { cudaFree(idx->ids); idx->ids = nullptr; }
  idx->has_store = false; idx->tmap_x_ok = false;
  idx->n_total = n_total; idx->id_base = id_base; idx->rows_seen = 0;

  const int n_blocks = (int)((n_total + kStoreBlockObjs - 1) / kStoreBlockObjs);
  LMI_CUDA(idx->cats.reserve((size_t)n_total * 8));
  LMI_CUDA(idx->pos.reserve((size_t)n_total * 8));
  LMI_CUDA(idx->block_hist.reserve((size_t)n_blocks * C * 4));
  LMI_CUDA(idx->bucket_size.reserve((size_t)C * 8));
  LMI_CUDA(idx->keep_lo.reserve((size_t)C * 8));
  LMI_CUDA(idx->keep_hi.reserve((size_t)C * 8));
  LMI_CUDA(idx->owner.reserve((size_t)C * 4));
  LMI_CUDA(idx->bucket_ptr.reserve((size_t)(C + 1) * 8));
  cudaStream_t st = idx->stream;
  LMI_CUDA(cudaMemcpyAsync(idx->cats.p, pred_categories, (size_t)n_total * 8,
                           on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice, st));
  const size_t hist_smem = (size_t)C * sizeof(int);
  if (hist_smem > 48 * 1024) {
    LMI_CUDA(cudaFuncSetAttribute(store_hist_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_smem));
    LMI_CUDA(cudaFuncSetAttribute(store_assign_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)hist_smem));
  }
  store_hist_kernel<<<n_blocks, kStoreThreads, hist_smem, st>>>(idx->cats.as<int64_t>(), n_total, C, idx->block_hist.as<int32_t>());
  store_scan_blocks_kernel<<<(C + 127) / 128, 128, 0, st>>>(idx->block_hist.as<int32_t>(), n_blocks, C, idx->bucket_size.as<int64_t>());
  LMI_CUDA(cudaGetLastError());
  std::vector<int64_t> sizes(C);
  LMI_CUDA(cudaMemcpyAsync(sizes.data(), idx->bucket_size.p, (size_t)C * 8, cudaMemcpyDeviceToHost, st));
  LMI_CUDA(cudaStreamSynchronize(st));
  int64_t seen = 0;
  for (int64_t s : sizes) seen += s;
  if (seen != n_total) return set_error(LMI_ERR_ARG, "pred_categories holds values outside [0, n_categories)");
  std::vector<int32_t> owner = lpt_owners(sizes, idx->world);
  LMI_CUDA(cudaMemcpyAsync(idx->owner.p, owner.data(), (size_t)C * 4, cudaMemcpyHostToDevice, st));
  store_layout_kernel<<<1, 1, 0, st>>>(idx->bucket_size.as<int64_t>(), C, idx->rank, idx->world, idx->policy,
                                       idx->owner.as<int32_t>(), idx->keep_lo.as<int64_t>(), idx->keep_hi.as<int64_t>(),
                                       idx->bucket_ptr.as<int64_t>());
  store_assign_kernel<<<n_blocks, kStoreThreads, hist_smem, st>>>(idx->cats.as<int64_t>(), n_total, C, idx->block_hist.as<int32_t>(),
                                                                 idx->keep_lo.as<int64_t>(), idx->keep_hi.as<int64_t>(),
                                                                 idx->bucket_ptr.as<int64_t>(), idx->pos.as<int64_t>());
  LMI_CUDA(cudaGetLastError());
  idx->h_bucket_ptr.assign(C + 1, 0);
  LMI_CUDA(cudaMemcpyAsync(idx->h_bucket_ptr.data(), idx->bucket_ptr.p, (size_t)(C + 1) * 8, cudaMemcpyDeviceToHost, st));
  LMI_CUDA(cudaStreamSynchronize(st));
  idx->n_local = idx->h_bucket_ptr[C];
  const size_t rows_alloc = (size_t)std::max<int64_t>(idx->n_local, 1);
  LMI_CUDA(cudaMalloc(&idx->x, rows_alloc * idx->d * 2));
  LMI_CUDA(cudaMalloc(&idx->inv_norm, rows_alloc * 4));
  LMI_CUDA(cudaMalloc(&idx->ids, rows_alloc * 4));
  // bucket-size bookkeeping for per-search workspace bounds and the chunk-size choice
  idx->h_sizes_sorted.assign(C, 0);
  idx->n_nonempty = 0;
  int64_t global_of_local = 0;                     // global size of the buckets this shard holds part of
  for (int c = 0; c < C; ++c) {
    const int64_t n = idx->h_bucket_ptr[c + 1] - idx->h_bucket_ptr[c];
    idx->h_sizes_sorted[c] = n;
    if (n > 0) { ++idx->n_nonempty; global_of_local += sizes[c]; }
  }
  idx->pair_fraction = n_total > 0 ? (double)global_of_local / (double)n_total : 1.0;
  std::sort(idx->h_sizes_sorted.begin(), idx->h_sizes_sorted.end(), [](int64_t a, int64_t b) { return a > b; });
  idx->store_open = true;
  return LMI_OK;
}

int lmi_store_append(lmi_index* idx, const uint16_t* rows_f16, int64_t first_row, int64_t rows, int on_device) {
  if (!idx || !rows_f16) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->store_open) return set_error(LMI_ERR_STATE, "lmi_store_begin first");
  if (first_row < 0 || rows < 0 || first_row + rows > idx->n_total) return set_error(LMI_ERR_ARG, "row range outside [0, n_total)");
  LMI_CUDA(cudaSetDevice(idx->device));
  cudaStream_t st = idx->stream;
  const int d = idx->d;
  const int64_t step = on_device ? rows : std::min<int64_t>(rows, (int64_t)(256u << 20) / (d * 2));   // 256 MB staging
  for (int64_t lo = 0; lo < rows; lo += std::max<int64_t>(step, 1)) {
    const int64_t cnt = std::min(step, rows - lo);
    const uint16_t* src = rows_f16 + (size_t)lo * d;
    if (!on_device) {
      LMI_CUDA(idx->stage_rows.reserve((size_t)cnt * d * 2));
      LMI_CUDA(cudaMemcpyAsync(idx->stage_rows.p, src, (size_t)cnt * d * 2, cudaMemcpyHostToDevice, st));
      src = idx->stage_rows.as<uint16_t>();
    }
    const int grid = (int)std::min<int64_t>((cnt + 7) / 8, 148 * 16);
    if (grid > 0)
      store_scatter_rows_kernel<<<grid, 256, 0, st>>>(src, first_row + lo, cnt, d, idx->pos.as<int64_t>(), idx->id_base,
                                                      idx->x, idx->inv_norm, idx->ids);
    LMI_CUDA(cudaGetLastError());
    if (!on_device) LMI_CUDA(cudaStreamSynchronize(st));       // staging buffer is reused
  }
  idx->rows_seen += rows;
  return LMI_OK;
}

int lmi_store_end(lmi_index* idx) {
  if (!idx) return set_error(LMI_ERR_ARG, "NULL index");
  if (!idx->store_open) return set_error(LMI_ERR_STATE, "lmi_store_begin first");
  LMI_CUDA(cudaSetDevice(idx->device));
  LMI_CUDA(cudaStreamSynchronize(idx->stream));
  if (idx->rows_seen != idx->n_total) return set_error(LMI_ERR_STATE, "not every row was appended exactly once");
  // build-time scratch is no longer needed (bucket_ptr stays: the planner reads it)
  idx->cats.release(); idx->pos.release(); idx->block_hist.release(); idx->stage_rows.release();
  if (idx->n_local > 0) {
    int rc = make_tmap(&idx->tmap_x, idx->x, (uint64_t)idx->n_local, idx->d, kTcBlockN);
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_x64, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN);
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_x32, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN / 2);
    // L2 prefetch boxes: 64 rows x up to 256 halves, no swizzle (nothing lands in shared memory)
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_xpf, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN,
                                     std::min(idx->d, kTsAtomsPerStage * kTsAtomK), CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc != LMI_OK) return rc;
    idx->tmap_x_ok = true;
    idx->tmap_x3_ok = false;
    if (idx->d % (kTsAtomsPerStage * kTsAtomK) == 0 &&
        make_tmap3(&idx->tmap_x3_64, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN, kTsAtomsPerStage) == LMI_OK &&
        make_tmap3(&idx->tmap_x3_32, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN / 2, kTsAtomsPerStage) == LMI_OK)
      idx->tmap_x3_ok = true;
  }
  idx->store_open = false;
  idx->has_store = true;
  return LMI_OK;
}

int lmi_set_buckets(lmi_index* idx, const uint16_t* data_f16, const int64_t* pred_categories, int64_t n_total, int64_t id_base) {
  int rc = lmi_store_begin(idx, pred_categories, n_total, id_base, 0);
  if (rc != LMI_OK) return rc;
  rc = lmi_store_append(idx, data_f16, 0, n_total, 0);
  if (rc != LMI_OK) return rc;
  return lmi_store_end(idx);
}

int lmi_get_bucket_ptr(lmi_index* idx, int64_t* out_bucket_ptr) {
  if (!idx || !out_bucket_ptr) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_store) return set_error(LMI_ERR_STATE, "no store");
  memcpy(out_bucket_ptr, idx->h_bucket_ptr.data(), (size_t)(idx->C + 1) * 8);
  return LMI_OK;
}

}  // extern "C"
