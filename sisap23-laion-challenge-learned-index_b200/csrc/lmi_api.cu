// lmi_api.cu -- C-ABI of liblmi_b200.so (see include/lmi_b200.h).  Host orchestration only; the
// kernels live in nav.cuh (K1), plan.cuh (K2), rescore_tc.cuh / rescore_simt.cuh (K3),
// finalize.cuh (K4+K5) and store.cuh (bucket store build).
#include "index_state.cuh"
#include "store.cuh"
#include "rescore_simt.cuh"
#include "rescore_tc.cuh"
#include "rescore_ts.cuh"
#include "rescore_ts2.cuh"
#include "finalize.cuh"

namespace lmi {

thread_local std::string g_last_error;
int set_error(int code, const std::string& msg) { g_last_error = msg; return code; }

static int next_kprime(int need) {
  for (int kp : {16, 32, 64, 128}) if (need <= kp) return kp;
  return -1;
}

// cuTensorMapEncodeTiled through the runtime (no link-time dependency on libcuda)
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D fp16 row-major [rows, d] tensor, box = [64 halves (128 B), box_rows], 128-byte swizzle
static int make_tmap(CUtensorMap* map, const void* base, uint64_t rows, int d, int box_rows,
                     int box_cols = kTcBlockK, CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B) {
  const CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return set_error(LMI_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t gdim[2] = {(cuuint64_t)d, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)d * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, swz, promo,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(LMI_ERR_CUDA, "cuTensorMapEncodeTiled failed, code " + std::to_string((int)r));
  return LMI_OK;
}

// The same [rows, d] fp16 matrix seen as a 3-D tensor (64 halves, rows, d/64 K-atoms): a box of
// (64, box_rows, atoms) lands in shared memory as [atom][row][128 B] with the 128-byte swizzle -- exactly the
// K-major UMMA operand layout of one pipeline stage -- so ONE cp.async.bulk.tensor fills a stage instead of
// one per 64-half atom.
static int make_tmap3(CUtensorMap* map, const void* base, uint64_t rows, int d, int box_rows, int box_atoms) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return set_error(LMI_ERR_CUDA, "cuTensorMapEncodeTiled entry point not found");
  cuuint64_t gdim[3] = {64, (cuuint64_t)rows, (cuuint64_t)(d / 64)};
  cuuint64_t gstride[2] = {(cuuint64_t)d * 2, 128};
  cuuint32_t box[3] = {64, (cuuint32_t)box_rows, (cuuint32_t)box_atoms};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), gdim, gstride, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return set_error(LMI_ERR_CUDA, "cuTensorMapEncodeTiled (3-D) failed, code " + std::to_string((int)r));
  return LMI_OK;
}

}  // namespace lmi

using namespace lmi;

extern "C" {

const char* lmi_last_error(void) { return g_last_error.c_str(); }
const char* lmi_version(void) { return "lmi_b200 0.1.0 (sm_100a)"; }

int lmi_create(lmi_index** out, int device, int d_nav, int d, int n_categories) {
  if (!out) return set_error(LMI_ERR_ARG, "out is NULL");
  if (d_nav <= 0 || d_nav > 4096) return set_error(LMI_ERR_ARG, "d_nav out of range");
  if (d <= 0 || d % 64 != 0 || d > 1024) return set_error(LMI_ERR_ARG, "d must be a multiple of 64, <= 1024");
  if (n_categories <= 0 || n_categories > 65536) return set_error(LMI_ERR_ARG, "n_categories out of range");
  int ndev = 0;
  LMI_CUDA(cudaGetDeviceCount(&ndev));
  if (device < 0 || device >= ndev) return set_error(LMI_ERR_ARG, "no such CUDA device");
  LMI_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  LMI_CUDA(cudaGetDeviceProperties(&prop, device));
  if (prop.major != 10)
    return set_error(LMI_ERR_CUDA, std::string("lmi_b200 is built for sm_100a only; device is sm_") +
                                       std::to_string(prop.major) + std::to_string(prop.minor));
  lmi_index* idx = new lmi_index();
  idx->device = device;
  idx->num_sms = prop.multiProcessorCount;
  idx->d_nav = d_nav; idx->d = d; idx->C = n_categories;
  LMI_CUDA(cudaStreamCreateWithFlags(&idx->stream, cudaStreamNonBlocking));
  LMI_CUDA(cudaStreamCreateWithFlags(&idx->copy_stream, cudaStreamNonBlocking));
  LMI_CUDA(cudaEventCreateWithFlags(&idx->ev_q16, cudaEventDisableTiming));
  for (auto& e : idx->ev) LMI_CUDA(cudaEventCreate(&e));
  for (auto& pr : idx->ev_k3) for (auto& e : pr) LMI_CUDA(cudaEventCreate(&e));
  LMI_CUDA(cudaMallocHost(&idx->h_scalars, sizeof(PlanScalars)));
  memset(idx->h_scalars, 0, sizeof(PlanScalars));
  *out = idx;
  return LMI_OK;
}

void lmi_destroy(lmi_index* idx) {
  if (!idx) return;
  cudaSetDevice(idx->device);
  cudaDeviceSynchronize();
  for (void* p : idx->mlp_allocs) cudaFree(p);
  DevBuf* bufs[] = {&idx->cats, &idx->pos, &idx->block_hist, &idx->bucket_size, &idx->keep_lo, &idx->keep_hi,
                    &idx->owner, &idx->bucket_ptr, &idx->stage_rows, &idx->ws_qnav, &idx->ws_q16, &idx->ws_out_d,
                    &idx->ws_out_i, &idx->ws_thr, &idx->ws_cls64, &idx->top_cls, &idx->m_b, &idx->qoff,
                    &idx->chunks_b, &idx->plist_base, &idx->slot_pair, &idx->pair_slot, &idx->items, &idx->plists,
                    &idx->qg, &idx->scalars, &idx->qthr, &idx->nav_cls, &idx->nav_probs, &idx->nav_x, &idx->stage_rows_t, &idx->uncert};
  for (DevBuf* b : bufs) b->release();
  if (idx->x) cudaFree(idx->x);
  if (idx->inv_norm) cudaFree(idx->inv_norm);
  if (idx->ids) cudaFree(idx->ids);
  idx->pin_in.release(); idx->pin_out.release();
  if (idx->h_scalars) cudaFreeHost(idx->h_scalars);
  for (int b = 0; b < 2; ++b) { if (idx->pin_stage[b]) cudaFreeHost(idx->pin_stage[b]); if (idx->pin_ev[b]) cudaEventDestroy(idx->pin_ev[b]); }
  for (auto& e : idx->ev) if (e) cudaEventDestroy(e);
  for (auto& pr : idx->ev_k3) for (auto& e : pr) if (e) cudaEventDestroy(e);
  if (idx->stream) cudaStreamDestroy(idx->stream);
  if (idx->copy_stream) cudaStreamDestroy(idx->copy_stream);
  if (idx->ev_q16) cudaEventDestroy(idx->ev_q16);
  delete idx;
}

int lmi_set_option(lmi_index* idx, const char* name, int64_t value) {
  if (!idx || !name) return set_error(LMI_ERR_ARG, "NULL argument");
  std::string n(name);
  if (n == "rescore") {
    if (value != LMI_RESCORE_SIMT && value != LMI_RESCORE_TC && value != LMI_RESCORE_TC_SS && value != LMI_RESCORE_TC_PAIR)
      return set_error(LMI_ERR_ARG, "rescore: 0 (simt), 1 (tcgen05, queries in TMEM), 2 (tcgen05, smem operands) or 3 (tcgen05 CTA pairs)");
    idx->rescore = (int)value;
  } else if (n == "chunk_objects") {
    if (value != 0 && (value < 256 || value > (1 << 24))) return set_error(LMI_ERR_ARG, "chunk_objects out of range (0 = automatic)");
    idx->chunk_objects = (int)((value + 255) / 256 * 256);
  } else if (n == "kprime_extra") {
    if (value < 0 || value > 64) return set_error(LMI_ERR_ARG, "kprime_extra out of range");
    idx->kprime_extra = (int)value;
  } else if (n == "lockstep") {
    if (value < -1 || value > 1) return set_error(LMI_ERR_ARG, "lockstep: 0 (off), 1 (on) or -1 (on when the index is a shard)");
    idx->lockstep = (int)value;
  } else if (n == "l2_hint") {
    idx->l2_hint = value ? 1 : 0;
  } else if (n == "item_order") {
    if (value < 0 || value > 2) return set_error(LMI_ERR_ARG, "item_order: 0 (full pairs first), 1 (bucket-major) or 2 (0 + three-tile buckets as three half pairs)");
    idx->item_order = (int)value;
  } else if (n == "epi_sets") {
    if (value < 0 || value > 2) return set_error(LMI_ERR_ARG, "epi_sets: 0 (automatic), 1 or 2 epilogue warp sets in the CTA-pair kernel");
    idx->epi_sets = (int)value;
#ifdef LMI_EXPERIMENTS
  } else if (n == "dbg_flags") {          // kernel experiments: scripts/dev builds only (results invalid when non-zero)
    idx->dbg_flags = (int)value;
#endif
  } else if (n == "timing") {
    idx->timing = value ? 1 : 0;
  } else {
    return set_error(LMI_ERR_ARG, "unknown option " + n);
  }
  return LMI_OK;
}

int lmi_set_shard(lmi_index* idx, int rank, int world, int policy) {
  if (!idx) return set_error(LMI_ERR_ARG, "NULL index");
  if (world < 1 || rank < 0 || rank >= world) return set_error(LMI_ERR_ARG, "bad rank/world");
  if (policy != LMI_SHARD_BUCKETS && policy != LMI_SHARD_STRIPES) return set_error(LMI_ERR_ARG, "bad shard policy");
  if (idx->has_store || idx->store_open) return set_error(LMI_ERR_STATE, "set the shard before building the store");
  idx->rank = rank; idx->world = world; idx->policy = policy;
  return LMI_OK;
}

int lmi_set_mlp(lmi_index* idx, int n_layers, const int32_t* out_dims, const float* const* W, const float* const* b) {
  if (!idx || !out_dims || !W || !b) return set_error(LMI_ERR_ARG, "NULL argument");
  if (n_layers < 1 || n_layers > kNavMaxLayers) return set_error(LMI_ERR_ARG, "1..8 linear layers supported");
  if (out_dims[n_layers - 1] != idx->C) return set_error(LMI_ERR_ARG, "last layer width must equal n_categories");
  LMI_CUDA(cudaSetDevice(idx->device));
  for (void* p : idx->mlp_allocs) cudaFree(p);
  idx->mlp_allocs.clear();
  idx->has_mlp = false;           // until the new stack is complete: an error below must not leave pointers to freed weights
  idx->layers = NavLayers{};
  NavLayers L{};
  L.n_layers = n_layers;
  L.max_width = idx->d_nav;
  int in = idx->d_nav;
  for (int l = 0; l < n_layers; ++l) {
    const int out = out_dims[l];
    if (out <= 0 || out > 65536) return set_error(LMI_ERR_ARG, "layer width out of range");
    L.in_dim[l] = in; L.out_dim[l] = out;
    L.max_width = std::max(L.max_width, out);
    const int pitch = (out + 3) & ~3;                   // rows padded with zeros to a multiple of 4 floats (16-byte loads)
    L.pitch[l] = pitch;
    std::vector<float> wt((size_t)in * pitch, 0.f);     // transpose nn.Linear.weight [out,in] -> [in,pitch]
    for (int o = 0; o < out; ++o)
      for (int i = 0; i < in; ++i) wt[(size_t)i * pitch + o] = W[l][(size_t)o * in + i];
    float *dw = nullptr, *db = nullptr;
    LMI_CUDA(cudaMalloc(&dw, wt.size() * sizeof(float)));
    idx->mlp_allocs.push_back(dw);
    LMI_CUDA(cudaMalloc(&db, (size_t)out * sizeof(float)));
    idx->mlp_allocs.push_back(db);
    LMI_CUDA(cudaMemcpy(dw, wt.data(), wt.size() * sizeof(float), cudaMemcpyHostToDevice));
    LMI_CUDA(cudaMemcpy(db, b[l], (size_t)out * sizeof(float), cudaMemcpyHostToDevice));
    L.wt[l] = dw; L.bias[l] = db;
    in = out;
  }
  // activation ping-pong: buffer A holds the input and the outputs of layers 1, 3, ...; buffer B those of layers 0, 2, ...
  L.width_a = idx->d_nav; L.width_b = 0;
  for (int l = 0; l < n_layers; ++l) {
    int& w = (l & 1) ? L.width_a : L.width_b;
    w = std::max(w, L.out_dim[l]);
  }
  L.width_a = (L.width_a + 3) & ~3; L.width_b = (L.width_b + 3) & ~3;
  if (nav_smem_bytes(L) > 200 * 1024) return set_error(LMI_ERR_ARG, "MLP too wide for the navigation kernel (max width 800)");
  idx->layers = L;
  idx->has_mlp = true;            // (the shared-memory opt-in of nav_kernel is set at every launch: launch_nav)
  return LMI_OK;
}

}  // extern "C"

#include "lmi_api_store.inl"
#include "lmi_api_search.inl"

// ---- allocation-free navigation ----------------------------------------------------------------------------------
// lmi_navigate_device_noalloc (include/lmi_b200.h): predict_proba on device buffers for callers that cannot allocate
// (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  K1 writes its int32 bucket lists into a
// file-scope device array instead of a grow-only workspace, and a second kernel widens them to the int64 the
// reference's torch.topk returns.  No counterpart in the reference.
//
// The array is part of the library image: every GPU that loads liblmi_b200.so holds its 4 MiB until the process exits.
// All calls on a device share it, so two calls that overlap on the GPU (different streams, or different indexes on
// one device) race; calls on one stream are ordered and safe.  Nothing in it has to start at zero: K1 writes every
// entry [0, nq * n_top) before the widening kernel reads it, so the previous call's contents are never observed.
namespace lmi {

constexpr int64_t kNoAllocMaxEntries = (int64_t)1 << 20;          // nq * n_top <= 1,048,576 (e.g. 10,000 queries x 100 buckets)
static __device__ __align__(256) int32_t g_noalloc_classes[kNoAllocMaxEntries];

__global__ void __launch_bounds__(256)
noalloc_widen_kernel(const int32_t* __restrict__ cls32, int64_t n, int64_t* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int64_t)cls32[i];
}

// In host code `g_noalloc_classes` is only a handle for the runtime, not a device address: the address is asked for
// once per device, with that device current, and kept.
constexpr int kNoAllocMaxDevices = 64;
static std::mutex g_noalloc_mu;
static int32_t* g_noalloc_ptr[kNoAllocMaxDevices] = {};

static int noalloc_scratch(int device, int32_t** out) {
  if (device < 0 || device >= kNoAllocMaxDevices) return set_error(LMI_ERR_ARG, "device ordinal out of range for the static scratch table");
  std::lock_guard<std::mutex> g(g_noalloc_mu);
  if (!g_noalloc_ptr[device]) {
    void* p = nullptr;
    LMI_CUDA(cudaGetSymbolAddress(&p, g_noalloc_classes));          // current device = `device` (set by the caller)
    if (!p) return set_error(LMI_ERR_CUDA, "cudaGetSymbolAddress returned a null address");
    g_noalloc_ptr[device] = static_cast<int32_t*>(p);
  }
  *out = g_noalloc_ptr[device];
  return LMI_OK;
}

}  // namespace lmi

extern "C" int lmi_navigate_device_noalloc(lmi_index* idx, const float* q_nav_dev, int64_t nq, int n_top,
                                           int64_t* out_classes_dev, float* out_probs_dev, void* stream) {
  using namespace lmi;
  if (!idx || !q_nav_dev || !out_classes_dev) return set_error(LMI_ERR_ARG, "NULL argument");
  if (!idx->has_mlp) return set_error(LMI_ERR_STATE, "Model is not trained, call `build` first.");
  if (n_top < 1 || n_top > idx->C) return set_error(LMI_ERR_ARG, "n_top must be in [1, n_categories]");
  if (nq < 0) return set_error(LMI_ERR_ARG, "nq < 0");
  if (nq > kNoAllocMaxEntries / n_top)
    return set_error(LMI_ERR_ARG, "lmi_navigate_device_noalloc: nq * n_top = " + std::to_string(nq) + " * " + std::to_string(n_top) +
                                      " exceeds the static scratch of " + std::to_string(kNoAllocMaxEntries) +
                                      " entries; split the batch or use lmi_navigate_device");
  if (nq == 0) return LMI_OK;
  LMI_CUDA(cudaSetDevice(idx->device));
  int32_t* scratch = nullptr;
  int rc = noalloc_scratch(idx->device, &scratch);
  if (rc != LMI_OK) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  rc = launch_nav(idx, q_nav_dev, nq, n_top, scratch, out_probs_dev, st);
  if (rc != LMI_OK) return rc;
  const int64_t n = nq * n_top;
  noalloc_widen_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(scratch, n, out_classes_dev);
  LMI_CUDA(cudaGetLastError());
  return LMI_OK;
}
