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
  idx->h_global_sizes = sizes;          // (finish_store derives pair_fraction / chunking bounds from these)
  idx->store_open = true;
  return LMI_OK;
}

// Host -> device upload of one row range.  The source may be row-major (col_stride = 1; row_stride >= d) or
// feature-major (row_stride = 1: the block of a pandas DataFrame built as search.py:83-84 builds it, whose
// .to_numpy() is an F-ordered view -- uploading it as it lies avoids a 15 GB host-side transpose at 10M).  Chunks
// of <= 64 MB are copied into one of two PINNED staging buffers by a few host threads while the previous
// chunk's cudaMemcpyAsync + scatter kernel run on the stream.
static int append_host(lmi_index* idx, const uint16_t* base, int64_t first_row, int64_t rows, int64_t row_stride,
                       int64_t col_stride) {
  const int d = idx->d;
  cudaStream_t st = idx->stream;
  const bool feature_major = (col_stride != 1);
  const size_t kStage = (size_t)64 << 20;
  const int64_t step = std::max<int64_t>(64, (int64_t)(kStage / ((size_t)d * 2)) / 64 * 64);
  for (int b = 0; b < 2; ++b) {
    if (!idx->pin_stage[b]) {
      LMI_CUDA(cudaMallocHost(&idx->pin_stage[b], kStage));
      LMI_CUDA(cudaEventCreateWithFlags(&idx->pin_ev[b], cudaEventDisableTiming));
    }
  }
  LMI_CUDA(idx->stage_rows.reserve(kStage));
  if (feature_major) LMI_CUDA(idx->stage_rows_t.reserve(kStage));
  const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
  const int n_thr = (int)std::min<unsigned>(8, hw);
  int buf = 0;
  for (int64_t lo = 0; lo < rows; lo += step, buf ^= 1) {
    const int64_t cnt = std::min(step, rows - lo);
    uint16_t* pin = reinterpret_cast<uint16_t*>(idx->pin_stage[buf]);
    LMI_CUDA(cudaEventSynchronize(idx->pin_ev[buf]));               // the copy that last used this buffer is done
    // ---- host: gather the chunk into pinned memory ([cnt][d] row-major, or [d][cnt] feature-major) ----
    auto work = [&](int t) {
      if (!feature_major) {
        const int64_t r0 = cnt * t / n_thr, r1 = cnt * (t + 1) / n_thr;
        if (row_stride == d) memcpy(pin + (size_t)r0 * d, base + (size_t)(lo + r0) * d, (size_t)(r1 - r0) * d * 2);
        else for (int64_t r = r0; r < r1; ++r) memcpy(pin + (size_t)r * d, base + (size_t)(lo + r) * row_stride, (size_t)d * 2);
      } else {
        const int f0 = (int)((int64_t)d * t / n_thr), f1 = (int)((int64_t)d * (t + 1) / n_thr);
        for (int f = f0; f < f1; ++f) memcpy(pin + (size_t)f * cnt, base + (size_t)f * col_stride + lo, (size_t)cnt * 2);
      }
    };
    if (n_thr == 1 || cnt * d < (1 << 20)) { for (int t = 0; t < n_thr; ++t) work(t); }
    else {
      std::vector<std::thread> pool;
      for (int t = 1; t < n_thr; ++t) pool.emplace_back(work, t);
      work(0);
      for (auto& th : pool) th.join();
    }
    // ---- device: H2D, (transpose,) scatter into the CSR store ----
    const uint16_t* src = idx->stage_rows.as<uint16_t>();
    if (!feature_major) {
      LMI_CUDA(cudaMemcpyAsync(idx->stage_rows.p, pin, (size_t)cnt * d * 2, cudaMemcpyHostToDevice, st));
    } else {
      LMI_CUDA(cudaMemcpyAsync(idx->stage_rows_t.p, pin, (size_t)cnt * d * 2, cudaMemcpyHostToDevice, st));
      const dim3 grid((unsigned)((cnt + 63) / 64), (unsigned)((d + 63) / 64));
      store_transpose_kernel<<<grid, 256, 0, st>>>(idx->stage_rows_t.as<uint16_t>(), cnt, d, cnt, idx->stage_rows.as<uint16_t>());
    }
    LMI_CUDA(cudaEventRecord(idx->pin_ev[buf], st));
    const int grid = (int)std::min<int64_t>((cnt + 7) / 8, 148 * 16);
    store_scatter_rows_kernel<<<grid, 256, 0, st>>>(src, first_row + lo, cnt, d, idx->pos.as<int64_t>(), idx->id_base,
                                                    idx->x, idx->inv_norm, idx->ids);
    LMI_CUDA(cudaGetLastError());
  }
  LMI_CUDA(cudaStreamSynchronize(st));
  return LMI_OK;
}

int lmi_store_append_strided(lmi_index* idx, const uint16_t* base, int64_t first_row, int64_t rows, int64_t row_stride,
                             int64_t col_stride, int on_device) {
  if (!idx || !base) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->store_open) return set_error(LMI_ERR_STATE, "lmi_store_begin first");
  if (first_row < 0 || rows < 0 || first_row + rows > idx->n_total) return set_error(LMI_ERR_ARG, "row range outside [0, n_total)");
  const int d = idx->d;
  const bool row_major = (col_stride == 1 && row_stride >= d);
  const bool feature_major = (row_stride == 1 && col_stride >= rows && d > 1);
  if (!row_major && !feature_major)
    return set_error(LMI_ERR_ARG, "source must be row-major (col_stride 1, row_stride >= d) or feature-major (row_stride 1, col_stride >= rows)");
  LMI_CUDA(cudaSetDevice(idx->device));
  if (rows == 0) return LMI_OK;
  if (!on_device) {
    int rc = append_host(idx, base, first_row, rows, row_stride, row_major ? 1 : col_stride);
    if (rc != LMI_OK) return rc;
  } else {
    cudaStream_t st = idx->stream;
    const uint16_t* src = base;
    if (feature_major || row_stride != d) {              // bring the chunk to dense row-major first
      const int64_t step = std::max<int64_t>(64, (int64_t)(((size_t)256 << 20) / ((size_t)d * 2)));
      LMI_CUDA(idx->stage_rows.reserve((size_t)std::min(step, rows) * d * 2));
      for (int64_t lo = 0; lo < rows; lo += step) {
        const int64_t cnt = std::min(step, rows - lo);
        if (feature_major) {
          const dim3 grid((unsigned)((cnt + 63) / 64), (unsigned)((d + 63) / 64));
          store_transpose_kernel<<<grid, 256, 0, st>>>(base + lo, cnt, d, col_stride, idx->stage_rows.as<uint16_t>());
        } else {
          LMI_CUDA(cudaMemcpy2DAsync(idx->stage_rows.p, (size_t)d * 2, base + (size_t)lo * row_stride, (size_t)row_stride * 2,
                                     (size_t)d * 2, (size_t)cnt, cudaMemcpyDeviceToDevice, st));
        }
        const int grid = (int)std::min<int64_t>((cnt + 7) / 8, 148 * 16);
        store_scatter_rows_kernel<<<grid, 256, 0, st>>>(idx->stage_rows.as<uint16_t>(), first_row + lo, cnt, d, idx->pos.as<int64_t>(),
                                                        idx->id_base, idx->x, idx->inv_norm, idx->ids);
        LMI_CUDA(cudaGetLastError());
      }
    } else {
      const int grid = (int)std::min<int64_t>((rows + 7) / 8, 148 * 16);
      store_scatter_rows_kernel<<<grid, 256, 0, st>>>(src, first_row, rows, d, idx->pos.as<int64_t>(), idx->id_base,
                                                      idx->x, idx->inv_norm, idx->ids);
      LMI_CUDA(cudaGetLastError());
    }
  }
  idx->rows_seen += rows;
  return LMI_OK;
}

int lmi_store_append(lmi_index* idx, const uint16_t* rows_f16, int64_t first_row, int64_t rows, int on_device) {
  return lmi_store_append_strided(idx, rows_f16, first_row, rows, idx ? idx->d : 0, 1, on_device);
}

namespace lmi {
// host bookkeeping + tensor maps over a complete store (end of a build, or after a load from the cache file)
static int finish_store(lmi_index* idx, const std::vector<int64_t>* global_sizes) {
  const int C = idx->C;
  idx->n_local = idx->h_bucket_ptr[C];
  idx->h_sizes_sorted.assign(C, 0);
  idx->n_nonempty = 0;
  int64_t global_of_local = 0;                     // global size of the buckets this shard holds part of
  for (int c = 0; c < C; ++c) {
    const int64_t n = idx->h_bucket_ptr[c + 1] - idx->h_bucket_ptr[c];
    idx->h_sizes_sorted[c] = n;
    if (n > 0) { ++idx->n_nonempty; if (global_sizes) global_of_local += (*global_sizes)[c]; }
  }
  if (global_sizes) idx->pair_fraction = idx->n_total > 0 ? (double)global_of_local / (double)idx->n_total : 1.0;
  std::sort(idx->h_sizes_sorted.begin(), idx->h_sizes_sorted.end(), [](int64_t a, int64_t b) { return a > b; });
  idx->tmap_x_ok = false;
  idx->tmap_x3_ok = false;
  if (idx->n_local > 0) {
    int rc = make_tmap(&idx->tmap_x, idx->x, (uint64_t)idx->n_local, idx->d, kTcBlockN);
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_x64, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN);
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_x32, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN / 2);
    // L2 prefetch boxes: 64 rows x up to 256 halves, no swizzle (nothing lands in shared memory)
    if (rc == LMI_OK) rc = make_tmap(&idx->tmap_xpf, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN,
                                     std::min(idx->d, kTsAtomsPerStage * kTsAtomK), CU_TENSOR_MAP_SWIZZLE_NONE);
    if (rc != LMI_OK) return rc;
    idx->tmap_x_ok = true;
    if (idx->d % (kTsAtomsPerStage * kTsAtomK) == 0 &&
        make_tmap3(&idx->tmap_x3_64, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN, kTsAtomsPerStage) == LMI_OK &&
        make_tmap3(&idx->tmap_x3_32, idx->x, (uint64_t)idx->n_local, idx->d, kTsBlockN / 2, kTsAtomsPerStage) == LMI_OK)
      idx->tmap_x3_ok = true;
  }
  idx->store_open = false;
  idx->has_store = true;
  return LMI_OK;
}

static void release_build_scratch(lmi_index* idx) {
  idx->cats.release(); idx->pos.release(); idx->block_hist.release(); idx->stage_rows.release(); idx->stage_rows_t.release();
  for (int b = 0; b < 2; ++b) {
    if (idx->pin_stage[b]) { cudaFreeHost(idx->pin_stage[b]); idx->pin_stage[b] = nullptr; }
    if (idx->pin_ev[b]) { cudaEventDestroy(idx->pin_ev[b]); idx->pin_ev[b] = nullptr; }
  }
}
}  // namespace lmi

int lmi_store_end(lmi_index* idx) {
  if (!idx) return set_error(LMI_ERR_ARG, "NULL index");
  if (!idx->store_open) return set_error(LMI_ERR_STATE, "lmi_store_begin first");
  LMI_CUDA(cudaSetDevice(idx->device));
  LMI_CUDA(cudaStreamSynchronize(idx->stream));
  if (idx->rows_seen != idx->n_total) return set_error(LMI_ERR_STATE, "not every row was appended exactly once");
  release_build_scratch(idx);            // (bucket_ptr stays: the planner reads it)
  return finish_store(idx, &idx->h_global_sizes);
}

int lmi_set_buckets(lmi_index* idx, const uint16_t* data_f16, const int64_t* pred_categories, int64_t n_total, int64_t id_base) {
  int rc = lmi_store_begin(idx, pred_categories, n_total, id_base, 0);
  if (rc != LMI_OK) return rc;
  rc = lmi_store_append(idx, data_f16, 0, n_total, 0);
  if (rc != LMI_OK) return rc;
  return lmi_store_end(idx);
}

// ---- on-disk cache of the device layout (the reference has no loader at all: utils.py:46-60
//      pickles the model only, so "load" there = re-running groupby + .loc on every search) ----
namespace lmi {
struct StoreFileHeader {
  char magic[8];               // "LMIB200S"
  uint32_t version, d, C, policy;
  int64_t n_total, n_local, id_base;
  int32_t rank, world;
  double pair_fraction;
  uint8_t key[32];             // caller's fingerprint of (pred_categories, data): checked on load
};
static const char kStoreMagic[8] = {'L', 'M', 'I', 'B', '2', '0', '0', 'S'};
}  // namespace lmi

int lmi_store_save(lmi_index* idx, const char* path, const uint8_t* key32) {
  if (!idx || !path || !key32) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_store) return set_error(LMI_ERR_STATE, "no store to save");
  LMI_CUDA(cudaSetDevice(idx->device));
  LMI_CUDA(cudaStreamSynchronize(idx->stream));
  const std::string tmp = std::string(path) + ".tmp";
  FILE* f = fopen(tmp.c_str(), "wb");
  if (!f) return set_error(LMI_ERR_ARG, std::string("cannot open ") + tmp);
  StoreFileHeader h{};
  memcpy(h.magic, kStoreMagic, 8);
  h.version = 1; h.d = (uint32_t)idx->d; h.C = (uint32_t)idx->C; h.policy = (uint32_t)idx->policy;
  h.n_total = idx->n_total; h.n_local = idx->n_local; h.id_base = idx->id_base; h.rank = idx->rank; h.world = idx->world;
  h.pair_fraction = idx->pair_fraction;
  memcpy(h.key, key32, 32);
  bool ok = fwrite(&h, sizeof(h), 1, f) == 1 &&
            fwrite(idx->h_bucket_ptr.data(), 8, (size_t)idx->C + 1, f) == (size_t)idx->C + 1;
  const size_t kBuf = (size_t)64 << 20;
  void* pin = nullptr;
  if (ok && cudaMallocHost(&pin, kBuf) != cudaSuccess) ok = false;
  auto dump = [&](const void* dev, size_t bytes) {
    for (size_t off = 0; ok && off < bytes; off += kBuf) {
      const size_t n = std::min(kBuf, bytes - off);
      if (cudaMemcpy(pin, (const char*)dev + off, n, cudaMemcpyDeviceToHost) != cudaSuccess || fwrite(pin, 1, n, f) != n) ok = false;
    }
  };
  if (ok && idx->n_local > 0) {
    dump(idx->ids, (size_t)idx->n_local * 4);
    dump(idx->inv_norm, (size_t)idx->n_local * 4);
    dump(idx->x, (size_t)idx->n_local * idx->d * 2);
  }
  if (pin) cudaFreeHost(pin);
  if (fclose(f) != 0) ok = false;
  if (!ok) { remove(tmp.c_str()); return set_error(LMI_ERR_CUDA, std::string("writing ") + path + " failed"); }
  if (rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return set_error(LMI_ERR_ARG, std::string("cannot rename to ") + path); }
  return LMI_OK;
}

int lmi_store_load(lmi_index* idx, const char* path, const uint8_t* key32) {
  if (!idx || !path || !key32) return set_error(LMI_ERR_ARG, "NULL argument");
  if (idx->store_open) return set_error(LMI_ERR_STATE, "a store build is in progress");
  FILE* f = fopen(path, "rb");
  if (!f) return set_error(LMI_ERR_ARG, std::string("cannot open ") + path);
  StoreFileHeader h{};
  std::vector<int64_t> bptr;
  auto fail = [&](int code, const std::string& msg) { fclose(f); return set_error(code, msg); };
  if (fread(&h, sizeof(h), 1, f) != 1 || memcmp(h.magic, kStoreMagic, 8) != 0 || h.version != 1)
    return fail(LMI_ERR_ARG, std::string(path) + " is not an lmi_b200 store file");
  if ((int)h.d != idx->d || (int)h.C != idx->C || h.rank != idx->rank || h.world != idx->world || (int)h.policy != idx->policy)
    return fail(LMI_ERR_STATE, "store file was written for another index shape / shard");
  if (memcmp(h.key, key32, 32) != 0) return fail(LMI_ERR_STATE, "store file key mismatch (other pred_categories / data)");
  if (h.n_local < 0 || h.n_total <= 0 || h.n_local > h.n_total) return fail(LMI_ERR_ARG, "corrupt store file header");
  bptr.resize((size_t)idx->C + 1);
  if (fread(bptr.data(), 8, bptr.size(), f) != bptr.size() || bptr[0] != 0 || bptr[idx->C] != h.n_local)
    return fail(LMI_ERR_ARG, "corrupt store file (bucket_ptr)");
  if (cudaSetDevice(idx->device) != cudaSuccess) return fail(LMI_ERR_CUDA, "cudaSetDevice failed");
  if (idx->x) { cudaFree(idx->x); idx->x = nullptr; }
  if (idx->inv_norm) { cudaFree(idx->inv_norm); idx->inv_norm = nullptr; }
  if (idx->ids) { cudaFree(idx->ids); idx->ids = nullptr; }
  idx->has_store = false; idx->tmap_x_ok = false;
  const size_t rows_alloc = (size_t)std::max<int64_t>(h.n_local, 1);
  if (cudaMalloc(&idx->x, rows_alloc * idx->d * 2) != cudaSuccess || cudaMalloc(&idx->inv_norm, rows_alloc * 4) != cudaSuccess ||
      cudaMalloc(&idx->ids, rows_alloc * 4) != cudaSuccess)
    return fail(LMI_ERR_NOMEM, "out of device memory loading the store");
  const size_t kBuf = (size_t)64 << 20;
  void* pin[2] = {nullptr, nullptr};
  cudaEvent_t ev[2] = {nullptr, nullptr};
  bool ok = true;
  for (int b = 0; b < 2 && ok; ++b)
    ok = cudaMallocHost(&pin[b], kBuf) == cudaSuccess && cudaEventCreateWithFlags(&ev[b], cudaEventDisableTiming) == cudaSuccess;
  int buf = 0;
  auto slurp = [&](void* dev, size_t bytes) {          // fread into one pinned buffer while the other one is in flight
    for (size_t off = 0; ok && off < bytes; off += kBuf, buf ^= 1) {
      const size_t n = std::min(kBuf, bytes - off);
      if (cudaEventSynchronize(ev[buf]) != cudaSuccess || fread(pin[buf], 1, n, f) != n ||
          cudaMemcpyAsync((char*)dev + off, pin[buf], n, cudaMemcpyHostToDevice, idx->stream) != cudaSuccess ||
          cudaEventRecord(ev[buf], idx->stream) != cudaSuccess)
        ok = false;
    }
  };
  if (ok && h.n_local > 0) {
    slurp(idx->ids, (size_t)h.n_local * 4);
    slurp(idx->inv_norm, (size_t)h.n_local * 4);
    slurp(idx->x, (size_t)h.n_local * idx->d * 2);
  }
  if (cudaStreamSynchronize(idx->stream) != cudaSuccess) ok = false;
  for (int b = 0; b < 2; ++b) { if (pin[b]) cudaFreeHost(pin[b]); if (ev[b]) cudaEventDestroy(ev[b]); }
  fclose(f);
  if (!ok) return set_error(LMI_ERR_CUDA, std::string("reading ") + path + " failed");
  idx->n_total = h.n_total; idx->id_base = h.id_base; idx->rows_seen = h.n_total;
  idx->h_bucket_ptr = bptr;
  idx->pair_fraction = h.pair_fraction;
  LMI_CUDA(idx->bucket_ptr.reserve((size_t)(idx->C + 1) * 8));
  LMI_CUDA(cudaMemcpy(idx->bucket_ptr.p, bptr.data(), bptr.size() * 8, cudaMemcpyHostToDevice));
  return finish_store(idx, nullptr);
}

int lmi_get_bucket_ptr(lmi_index* idx, int64_t* out_bucket_ptr) {
  if (!idx || !out_bucket_ptr) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_store) return set_error(LMI_ERR_STATE, "no store");
  memcpy(out_bucket_ptr, idx->h_bucket_ptr.data(), (size_t)(idx->C + 1) * 8);
  return LMI_OK;
}

}  // extern "C"
