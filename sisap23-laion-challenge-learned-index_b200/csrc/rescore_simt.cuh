// rescore_simt.cuh -- K3 (reference flavour): rescoring on CUDA cores.
//
// Same work list, same partial-list output and same row-list epilogue as the tcgen05 kernel in
// rescore_tc.cuh, but the 128 x 64 score tile is computed with fp32 FFMA on fp16 inputs.  It is
// the on-GPU cross-check for the tensor-core kernel (option "rescore" = LMI_RESCORE_SIMT) and is
// obviously-correct by construction; it is ~15x slower and never the default.
//
// Computes, per work item, score[r][c] = dot(q_r, x_c) * inv_norm(x_c) -- the fp32 image of the
// reference's ``cosine_similarity`` (search/li/utils.py:11,19) up to the per-row factor 1/|q_r|,
// which does not change a row's order and is applied in float64 by the finalize kernel.
#pragma once
#include "common.cuh"
#include "plan.cuh"
#include "rowlist.cuh"

namespace lmi {

constexpr int kSimtTN = 64;        // objects per tile
constexpr int kSimtTK = 32;        // k chunk
constexpr int kSimtAStride = kTileQ + 4;
constexpr int kSimtBStride = kSimtTN + 4;
constexpr int kSimtSStride = kTileQ + 4;

template <int KP>
constexpr size_t simt_smem_bytes() {
  return sizeof(float) * (kSimtTK * kSimtAStride + kSimtTK * kSimtBStride + kSimtTN * kSimtSStride + kSimtTN) +
         sizeof(int32_t) * kTileQ + (sizeof(float) + sizeof(int32_t)) * KP * kTileQ;
}

template <int KP>
__global__ void __launch_bounds__(kTileQ)
rescore_simt_kernel(const uint16_t* __restrict__ q,          // [nq, d] fp16 bits (original order)
                    const int32_t* __restrict__ slot_pair,   // slot -> pair index q*nb + r
                    int nb, int d,
                    const uint16_t* __restrict__ store,      // [n_local, d] fp16 bits, CSR by bucket
                    const float* __restrict__ inv_norm,      // [n_local]
                    const WorkItem* __restrict__ items, const PlanScalars* __restrict__ sc,
                    Cand* __restrict__ plists) {
  extern __shared__ __align__(16) unsigned char simt_smem_raw[];
  float* As = reinterpret_cast<float*>(simt_smem_raw);               // [TK][AStride]
  float* Bs = As + kSimtTK * kSimtAStride;                           // [TK][BStride]
  float* Ss = Bs + kSimtTK * kSimtBStride;                           // [TN][SStride]
  float* s_inv = Ss + kSimtTN * kSimtSStride;                        // [TN]
  int32_t* s_qi = reinterpret_cast<int32_t*>(s_inv + kSimtTN);       // [128] query row of each slot
  float* l_sc = reinterpret_cast<float*>(s_qi + kTileQ);             // [KP][128]
  int32_t* l_ps = reinterpret_cast<int32_t*>(l_sc + KP * kTileQ);    // [KP][128]

  const int tid = threadIdx.x;
  const int tx = tid & 7, ty = tid >> 3;          // 8 column groups x 16 row groups
  const int n_items = sc->n_items;
  RowList<KP> list;

  for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
    const WorkItem wi = items[it];
    __syncthreads();
    s_qi[tid] = (tid < wi.rows) ? slot_pair[wi.slot_base + tid] / nb : -1;
    list.reset(l_sc, l_ps, tid, tid < wi.rows);
    __syncthreads();
    const int my_q = s_qi[tid];

    for (int c0 = 0; c0 < wi.obj_count; c0 += kSimtTN) {
      const int ncols = min(kSimtTN, wi.obj_count - c0);
      float acc[8][8];
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

      for (int k0 = 0; k0 < d; k0 += kSimtTK) {
        // stage A: thread = row, 32 halves
        {
          uint4 u[4];
          if (my_q >= 0) {
            const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)my_q * d + k0);
#pragma unroll
            for (int v = 0; v < 4; ++v) u[v] = src[v];
          } else {
#pragma unroll
            for (int v = 0; v < 4; ++v) u[v] = make_uint4(0, 0, 0, 0);
          }
          const __half* h = reinterpret_cast<const __half*>(u);
#pragma unroll
          for (int k = 0; k < kSimtTK; ++k) As[k * kSimtAStride + tid] = __half2float(h[k]);
        }
        // stage B: thread = (object, half of the k chunk), 16 halves
        {
          const int o = tid >> 1, part = tid & 1;
          uint4 u[2];
          if (o < ncols) {
            const uint4* src = reinterpret_cast<const uint4*>(
                store + (size_t)(wi.obj_begin + c0 + o) * d + k0 + part * 16);
            u[0] = src[0]; u[1] = src[1];
          } else {
            u[0] = u[1] = make_uint4(0, 0, 0, 0);
          }
          const __half* h = reinterpret_cast<const __half*>(u);
#pragma unroll
          for (int k = 0; k < 16; ++k) Bs[(part * 16 + k) * kSimtBStride + o] = __half2float(h[k]);
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < kSimtTK; ++k) {
          const float4 a0 = *reinterpret_cast<const float4*>(As + k * kSimtAStride + ty * 8);
          const float4 a1 = *reinterpret_cast<const float4*>(As + k * kSimtAStride + ty * 8 + 4);
          const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
          float b[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) b[j] = Bs[k * kSimtBStride + tx + 8 * j];
#pragma unroll
          for (int i = 0; i < 8; ++i)
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        __syncthreads();
      }

      // score tile -> shared, [col][row]
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float* dst = Ss + (tx + 8 * j) * kSimtSStride + ty * 8;
        *reinterpret_cast<float4*>(dst) = make_float4(acc[0][j], acc[1][j], acc[2][j], acc[3][j]);
        *reinterpret_cast<float4*>(dst + 4) = make_float4(acc[4][j], acc[5][j], acc[6][j], acc[7][j]);
      }
      if (tid < kSimtTN) s_inv[tid] = (tid < ncols) ? inv_norm[wi.obj_begin + c0 + tid] : 0.f;
      __syncthreads();

      // epilogue: thread = row
      for (int c = 0; c < ncols; ++c) {
        const float s = Ss[c * kSimtSStride + tid] * s_inv[c];
        if (s > list.thr) list.insert(l_sc, l_ps, tid, s, wi.obj_begin + c0 + c);
      }
      __syncthreads();
    }

    if (tid < wi.rows)
      list.flush(l_sc, l_ps, tid, plists + (size_t)(wi.plist_base + (int64_t)tid * wi.plist_stride) * KP);
  }
}

}  // namespace lmi
