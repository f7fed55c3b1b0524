// common.cuh -- shared helpers for the lmi_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

// Kernel experiments (switches that disable parts of a kernel; results invalid by construction) exist only in
// builds made with -DLMI_EXPERIMENTS (scripts/dev); in the shipped library the tests below are constant false
// and the branches disappear.
#ifdef LMI_EXPERIMENTS
#define LMI_DBG(flags, bit) (((flags) & (bit)) != 0)
#else
#define LMI_DBG(flags, bit) false
#endif

namespace lmi {

constexpr float kNegInf = -3.0e38f;          // "empty" score in top-k' lists
constexpr double kEmptyDist = 10000.0;       // reference: LearnedIndex.py:138
constexpr int kTileQ = 128;                  // query rows per work item (= TMEM lanes / UMMA M)
constexpr int kMaxKPrime = 128;

// One unit of rescoring work: 128 query slots of one bucket x one chunk of its objects.
struct __align__(16) WorkItem {
  int32_t slot_base;      // first slot (row of the bucket-ordered query list)
  int32_t rows;           // valid rows (1..128)
  int32_t obj_begin;      // first object (CSR position, local to this shard)
  int32_t obj_count;      // objects in this chunk
  int64_t plist_base;     // partial list id of row 0; row r writes list plist_base + r*plist_stride
  int32_t plist_stride;   // = chunks of this bucket
  int32_t bucket;
  int32_t flags;          // kItemShared: entries 2i and 2i+1 score the same object chunk
  int32_t pad;
};
constexpr int kItemShared = 1;
constexpr int kItemHalf = 2;      // CTA-pair kernel: half tile (<= 64 rows per CTA), scored with M = 128 pair MMAs
constexpr int kItemLastReader = 4; // CTA-pair kernel: no later work item streams this object chunk (its loads may leave L2 first)
constexpr int kItemLeads = 8;       // CTA-pair kernel, lock-step option: the NEXT pair of the list streams the same chunk twice as fast; publish progress
constexpr int kItemFollows = 16;    //   ... this pair follows the previous one (same chunk): it loads a tile only once the leader has asked for it

struct __align__(8) Cand {  // entry of a partial top-k' list
  float score;              // dot(q, x) * inv_norm(x)  (cosine times |q|)
  int32_t pos;              // CSR position of the object, -1 = empty
};

// final per-rank result entry exchanged between GPUs (16 bytes)
struct __align__(16) Partial {
  double dist;              // 1 - cos in float64
  uint32_t id;              // 1-based object id, 0 = empty
  uint32_t rank;            // bucket rank (tie-break: lower first)
};

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace lmi
