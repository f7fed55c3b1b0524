// index_state.cuh -- host-side state of one lmi_index (one CUDA device / one shard).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>
#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <string>
#include <vector>

#include "../../include/lmi_b200.h"
#include "common.cuh"
#include "nav.cuh"
#include "plan.cuh"

namespace lmi {

extern thread_local std::string g_last_error;
int set_error(int code, const std::string& msg);

#define LMI_CUDA(expr)                                                                         \
  do {                                                                                         \
    cudaError_t _e = (expr);                                                                   \
    if (_e != cudaSuccess)                                                                     \
      return lmi::set_error(LMI_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e)); \
  } while (0)

// grow-only device buffer
struct DevBuf {
  void* p = nullptr;
  size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) { cudaFree(p); p = nullptr; cap = 0; }
    size_t want = bytes + bytes / 8 + 256;
    cudaError_t e = cudaMalloc(&p, want);
    if (e == cudaSuccess) cap = want;
    return e;
  }
  void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
  template <class T> T* as() const { return reinterpret_cast<T*>(p); }
};

// grow-only pinned host buffer
struct PinBuf {
  void* p = nullptr; size_t cap = 0;
  cudaError_t reserve(size_t bytes) {
    if (bytes <= cap) return cudaSuccess;
    if (p) { cudaFreeHost(p); p = nullptr; cap = 0; }
    cudaError_t e = cudaMallocHost(&p, bytes + bytes / 8 + 256);
    if (e == cudaSuccess) cap = bytes + bytes / 8 + 256;
    return e;
  }
  void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};


}  // namespace lmi

struct lmi_index {
  int device = 0;
  int num_sms = 148;
  int d_nav = 96, d = 768, C = 0;
  cudaStream_t stream = nullptr;            // stream of the host-buffer entry points
  cudaStream_t copy_stream = nullptr;       // uploads the fp16 queries while K1/K2 run (lmi_search)
  cudaEvent_t ev_q16 = nullptr;

  // options
  int rescore = LMI_RESCORE_TC_PAIR;
  int chunk_objects = 0;                    // objects per work item; 0 = chosen per search (choose_chunk)
  int kprime_extra = 6;
  int timing = 1;
  int dbg_flags = 0;                        // kernel experiments; honoured by -DLMI_EXPERIMENTS builds only
  int item_order = 0;                       // CTA-pair kernel work list: 0 = full pairs first, then half pairs; 1 = bucket-major
  int l2_hint = 0;                          // CTA-pair kernel: per-item L2 eviction priority on the chunk loads (option "l2_hint")
  int lockstep = 0;                         // CTA-pair kernel: half pairs follow their chunk's full pair in lock-step (option "lockstep": 0 off, 1 on, -1 on shards)
  bool lockstep_now = false;
  int epi_sets = 0;                         // CTA-pair kernel: epilogue warp sets (0 = automatic: 2 while k' <= 32)
  int rank = 0, world = 1, policy = LMI_SHARD_BUCKETS;

  // model
  lmi::NavLayers layers{};                  // device pointers inside
  std::vector<void*> mlp_allocs;
  bool has_mlp = false;

  // store
  bool store_open = false, has_store = false;
  int64_t n_total = 0, n_local = 0, id_base = 1;
  int64_t rows_seen = 0;
  lmi::DevBuf cats, pos, block_hist, bucket_size, keep_lo, keep_hi, owner, bucket_ptr;   // build-time
  lmi::DevBuf stage_rows, stage_rows_t;                                                  // host->device staging (row-major / feature-major)
  void* pin_stage[2] = {nullptr, nullptr};  // pinned host staging (double-buffered upload), freed by lmi_store_end
  cudaEvent_t pin_ev[2] = {nullptr, nullptr};
  std::vector<int64_t> h_global_sizes;      // bucket sizes over ALL ranks (set by lmi_store_begin)
  uint16_t* x = nullptr;                    // [n_local, d] fp16, CSR by bucket
  float* inv_norm = nullptr;                // [n_local]
  uint32_t* ids = nullptr;                  // [n_local]
  std::vector<int64_t> h_bucket_ptr;        // [C+1] local CSR (host copy)
  std::vector<int64_t> h_sizes_sorted;      // local bucket sizes, descending (workspace bounds for any chunk size)
  int64_t n_nonempty = 0;                   // local buckets holding objects
  double pair_fraction = 1.0;               // expected share of (query, bucket) pairs that hit a locally non-empty bucket
  int chunk_now = 0;                        // chunk size of the last search (stats)
  CUtensorMap tmap_x{};                     // store tensor map, box 256 x 64 (SS kernel)
  CUtensorMap tmap_x64{};                   // store tensor map, box 64 x 64 (TS kernel)
  CUtensorMap tmap_xpf{};                   // store tensor map, box 64 x 256 unswizzled (TS kernel, L2 prefetch)
  CUtensorMap tmap_x32{};                   // store tensor map, box 32 x 64 (TS kernel, shared pairs)
  CUtensorMap tmap_x3_64{}, tmap_x3_32{};     // 3-D views (64 halves, rows, K-atoms): one box = one pipeline stage
  bool tmap_x3_ok = false;
  bool tmap_x_ok = false;
  int kernel_now = LMI_RESCORE_SIMT;         // kernel chosen for the search in flight

  // search workspace (grow-only)
  lmi::DevBuf ws_qnav, ws_q16, ws_out_d, ws_out_i, ws_thr, ws_cls64;
  lmi::PinBuf pin_in, pin_out;               // staging for pageable caller buffers (lmi_search / lmi_search_rank)
  lmi::DevBuf top_cls, m_b, qoff, chunks_b, plist_base, slot_pair, pair_slot, items, plists, qg, scalars, qthr, uncert, progress;
  lmi::DevBuf nav_cls, nav_probs, nav_x;
  CUtensorMap tmap_q{};
  void* tmap_q_ptr = nullptr; size_t tmap_q_rows = 0;
  int64_t max_items = 0;

  // stats
  lmi::PlanScalars* h_scalars = nullptr;    // pinned
  cudaEvent_t ev[5]{};
  // ring of (start, stop) pairs around the rescoring kernel of the searches since the last lmi_get_stats
  static constexpr int kEvRing = 256;
  cudaEvent_t ev_k3[kEvRing][2]{};
  int64_t n_since_stats = 0;
  cudaStream_t last_stream = nullptr;
  lmi_stats stats{};
  bool stats_pending = false;
  std::string rescore_name = "none";
};
