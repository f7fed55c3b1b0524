// store.cuh -- bucket store build: per-object categories -> CSR layout of fp16 rows.
//
// Replaces what the reference redoes on every rank of every search: ``groupby('category')`` over
// all N rows (search/li/LearnedIndex.py:143) and the ``data_search.loc[g.index]`` gather
// (LearnedIndex.py:152,168).  Done once here: a stable counting sort by category (objects stay in
// ascending id order inside a bucket, like ``g.index``), rows copied contiguous per bucket, plus
// 1/|x| per object (scikit-learn's ``normalize`` divides by the L2 norm, zero norm -> 1) and the
// reference's 1-based object id.
#pragma once
#include "common.cuh"

namespace lmi {

constexpr int kStoreBlockObjs = 2048;     // objects per histogram block
constexpr int kStoreThreads = 256;

// pass 1: per-block category histogram.  block_hist: int32 [n_blocks][C]
__global__ void __launch_bounds__(kStoreThreads)
store_hist_kernel(const int64_t* __restrict__ cat, int64_t n, int C, int32_t* __restrict__ block_hist) {
  extern __shared__ int s_hist[];
  for (int c = threadIdx.x; c < C; c += kStoreThreads) s_hist[c] = 0;
  __syncthreads();
  const int64_t lo = (int64_t)blockIdx.x * kStoreBlockObjs;
  const int64_t hi = min(n, lo + kStoreBlockObjs);
  for (int64_t i = lo + threadIdx.x; i < hi; i += kStoreThreads) {
    const int64_t c = cat[i];
    if (c >= 0 && c < C) atomicAdd(&s_hist[(int)c], 1);
  }
  __syncthreads();
  for (int c = threadIdx.x; c < C; c += kStoreThreads) block_hist[(size_t)blockIdx.x * C + c] = s_hist[c];
}

// pass 2: per category, exclusive scan over blocks (in place) and the global size.
__global__ void store_scan_blocks_kernel(int32_t* __restrict__ block_hist, int n_blocks, int C,
                                         int64_t* __restrict__ bucket_size) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  int64_t run = 0;
  for (int b = 0; b < n_blocks; ++b) {
    const int v = block_hist[(size_t)b * C + c];
    block_hist[(size_t)b * C + c] = (int32_t)run;      // objects of category c in earlier blocks (< 2^31)
    run += v;
  }
  bucket_size[c] = run;
}

// pass 3 (single thread per quantity, C is small): shard decision + local CSR.
//   policy 0 (buckets): owner[c] given; local bucket = whole bucket if owner==rank else empty
//   policy 1 (stripes): local bucket = rows [n*rank/world, n*(rank+1)/world) of the bucket
//   keep_lo/keep_hi: in-bucket rank range kept locally; local_ptr: int64 [C+1]
__global__ void store_layout_kernel(const int64_t* __restrict__ bucket_size, int C, int rank, int world,
                                    int policy, const int32_t* __restrict__ owner,
                                    int64_t* __restrict__ keep_lo, int64_t* __restrict__ keep_hi,
                                    int64_t* __restrict__ local_ptr) {
  if (blockIdx.x != 0 || threadIdx.x != 0) return;
  int64_t run = 0;
  for (int c = 0; c < C; ++c) {
    const int64_t n = bucket_size[c];
    int64_t lo = 0, hi = n;
    if (world > 1) {
      if (policy == 0) { if (owner[c] != rank) hi = 0; }
      else { lo = n * rank / world; hi = n * (rank + 1) / world; }
    }
    keep_lo[c] = lo; keep_hi[c] = hi;
    local_ptr[c] = run;
    run += hi - lo;
  }
  local_ptr[C] = run;
}

// pass 4: stable position of every object: pos[i] = local CSR row, or -1 if not kept here.
// Objects of one block are processed in 8 ordered sub-batches of 256; inside a sub-batch the
// in-category rank comes from __match_any_sync plus an ordered walk over the 8 warps.
__global__ void __launch_bounds__(kStoreThreads)
store_assign_kernel(const int64_t* __restrict__ cat, int64_t n, int C,
                    const int32_t* __restrict__ block_prefix, const int64_t* __restrict__ keep_lo,
                    const int64_t* __restrict__ keep_hi, const int64_t* __restrict__ local_ptr,
                    int64_t* __restrict__ pos) {
  extern __shared__ int s_cnt[];                 // running count per category inside this block
  for (int c = threadIdx.x; c < C; c += kStoreThreads) s_cnt[c] = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int64_t lo = (int64_t)blockIdx.x * kStoreBlockObjs;
  const int32_t* pre = block_prefix + (size_t)blockIdx.x * C;
  for (int sub = 0; sub < kStoreBlockObjs / kStoreThreads; ++sub) {
    const int64_t i = lo + (int64_t)sub * kStoreThreads + threadIdx.x;
    int c = -1;
    if (i < n) { const int64_t cc = cat[i]; if (cc >= 0 && cc < C) c = (int)cc; }
    const unsigned same = __match_any_sync(0xffffffffu, c);
    const int rank_in_warp = __popc(same & ((1u << lane) - 1u));
    const bool leader = rank_in_warp == 0;
    int base = 0;
    for (int w = 0; w < kStoreThreads / 32; ++w) {          // warps take turns, in order
      if (warp == w && c >= 0 && leader) {
        base = s_cnt[c];
        s_cnt[c] = base + __popc(same);
      }
      __syncthreads();
    }
    base = __shfl_sync(0xffffffffu, base, __ffs(same) - 1);
    if (i < n) {
      int64_t p = -1;
      if (c >= 0) {
        const int64_t r = (int64_t)pre[c] + base + rank_in_warp;     // rank inside the bucket
        if (r >= keep_lo[c] && r < keep_hi[c]) p = local_ptr[c] + (r - keep_lo[c]);
      }
      pos[i] = p;
    }
  }
}

// pass 5: scatter a chunk of rows into the store.  One warp per row: copy d halves, compute
// 1/|x| (float64 sum of squares of the fp16 values, stored fp32; zero norm -> 1), write the id.
__global__ void __launch_bounds__(256)
store_scatter_rows_kernel(const uint16_t* __restrict__ rows, int64_t first_row, int64_t n_rows, int d,
                          const int64_t* __restrict__ pos, int64_t id_base,
                          uint16_t* __restrict__ store, float* __restrict__ inv_norm,
                          uint32_t* __restrict__ ids) {
  const int lane = threadIdx.x & 31;
  const int vec = d / 8;
  for (int64_t r = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows; r += (int64_t)gridDim.x * 8) {
    const int64_t p = pos[first_row + r];
    if (p < 0) continue;                                   // warp-uniform
    const uint4* src = reinterpret_cast<const uint4*>(rows + (size_t)r * d);
    uint4* dst = reinterpret_cast<uint4*>(store + (size_t)p * d);
    double ss = 0.0;
    for (int v = lane; v < vec; v += 32) {
      const uint4 u = src[v];
      dst[v] = u;
      const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __half22float2(h[j]);
        ss += (double)f.x * (double)f.x + (double)f.y * (double)f.y;
      }
    }
    ss = warp_sum(ss);
    if (lane == 0) {
      const double nrm = sqrt(ss);
      inv_norm[p] = (nrm == 0.0) ? 1.0f : (float)(1.0 / nrm);
      ids[p] = (uint32_t)(id_base + first_row + r);
    }
  }
}

// [d][n] (feature-major: how a pandas DataFrame block of [n, d] is laid out) -> [n][d] row-major, 64 x 64 tiles
// through shared memory; both sides coalesced.  n_pitch = elements between consecutive features in `src`.
__global__ void __launch_bounds__(256)
store_transpose_kernel(const uint16_t* __restrict__ src, int64_t n, int d, int64_t n_pitch, uint16_t* __restrict__ dst) {
  __shared__ uint16_t tile[64][66];
  const int64_t r0 = (int64_t)blockIdx.x * 64;
  const int f0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;          // 64 x 4
  for (int f = ty; f < 64; f += 4) {
    const int64_t r = r0 + tx;
    tile[f][tx] = (r < n && f0 + f < d) ? src[(size_t)(f0 + f) * n_pitch + r] : (uint16_t)0;
  }
  __syncthreads();
  for (int rr = ty; rr < 64; rr += 4) {
    const int64_t r = r0 + rr;
    if (r < n && f0 + tx < d) dst[(size_t)r * d + f0 + tx] = tile[tx][rr];
  }
}

}  // namespace lmi
