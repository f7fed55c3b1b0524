// nav.cuh -- K1: navigation.  fp32 MLP forward + softmax + per-query top-n bucket select.
//
// Replaces NeuralNetwork.predict_proba (search/li/model.py:214-229: model.eval(), forward,
// softmax(dim=1), prob.topk(C)) and NeuralNetwork.predict (model.py:201-212: argmax of the
// logits).  Arithmetic is plain fp32 FFMA (no TF32, no tensor cores) so bucket order matches
// the reference's torch fp32 model up to summation order.
//
// One CTA = 16 queries (128 threads).  Activations live in shared memory transposed ([feature][query]) so a
// thread that owns 8 queries x 4 outputs reads its 8 activations with two broadcast LDS.128;
// weights are stored transposed on the device ([in][out], rows padded to a multiple of 4 floats) and streamed
// through shared memory in tiles of 8 input rows x 256 outputs with 16-byte loads, the next tile's loads in flight
// while one is consumed.
#pragma once
#include "common.cuh"

namespace lmi {

constexpr int kNavTQ = 16;          // queries per CTA (10k queries = 625 CTAs, 5 per SM: all resident at once)
constexpr int kNavThreads = 128;
constexpr int kNavMaxLayers = 8;
constexpr int kNavKT = 8;           // input rows per weight tile
constexpr int kNavRegC = 8;          // classes per lane kept in registers by the top-n select (C <= 256)
constexpr int kNavOB = 256;         // outputs per block (4 per thread)

struct NavLayers {
  int n_layers;
  int in_dim[kNavMaxLayers];
  int out_dim[kNavMaxLayers];
  int pitch[kNavMaxLayers];          // floats per row of wt: out_dim rounded up to a multiple of 4, zero-filled (16-byte loads)
  const float* wt[kNavMaxLayers];    // [in][pitch]  (transposed nn.Linear.weight)
  const float* bias[kNavMaxLayers];  // [out]
  int max_width;                     // max over all in/out dims
  int width_a, width_b;              // activation buffers: A holds the input and every second layer's output, B the others
};

// One weight tile (rows x 256 outputs in shared memory) into a thread's 8 queries x NU*1 outputs.
// Straight-line code: NU (output groups of 64 in use) and FULL (all kNavKT rows) are compile-time.
template <int NU, bool FULL>
__device__ __forceinline__ void nav_tile_fma(float (&acc)[4][8], const float* __restrict__ w,
                                             const float4* __restrict__ a, int rows) {
#pragma unroll
  for (int r = 0; r < kNavKT; ++r) {
    if (FULL || r < rows) {
      const float4 lo = a[r * (kNavTQ / 4)];
      const float4 hi = a[r * (kNavTQ / 4) + 1];
#pragma unroll
      for (int u = 0; u < NU; ++u) {
        const float wv = w[r * kNavOB + 64 * u];
        acc[u][0] = fmaf(lo.x, wv, acc[u][0]); acc[u][1] = fmaf(lo.y, wv, acc[u][1]);
        acc[u][2] = fmaf(lo.z, wv, acc[u][2]); acc[u][3] = fmaf(lo.w, wv, acc[u][3]);
        acc[u][4] = fmaf(hi.x, wv, acc[u][4]); acc[u][5] = fmaf(hi.y, wv, acc[u][5]);
        acc[u][6] = fmaf(hi.z, wv, acc[u][6]); acc[u][7] = fmaf(hi.w, wv, acc[u][7]);
      }
    }
  }
}

// x: [n, d_nav] fp32.  out_classes: int32 [n, n_top]; out_probs: fp32 [n, n_top] or nullptr.
__global__ void __launch_bounds__(kNavThreads, 5)
nav_kernel(const float* __restrict__ x, int64_t n, NavLayers L, int n_top,
           int32_t* __restrict__ out_classes, float* __restrict__ out_probs) {
  extern __shared__ float nav_smem[];
  float* act_a = nav_smem;
  float* act_b = nav_smem + (size_t)L.width_a * kNavTQ;
  const int tid = threadIdx.x;
  const int64_t q0 = (int64_t)blockIdx.x * kNavTQ;
  const int nq_here = (int)min((int64_t)kNavTQ, n - q0);

  // stage the input tile transposed: act_a[i][q]
  const int d0 = L.in_dim[0];
  for (int idx = tid; idx < kNavTQ * d0; idx += kNavThreads) {
    int q = idx / d0, i = idx - q * d0;
    act_a[i * kNavTQ + q] = (q < nq_here) ? x[(q0 + q) * d0 + i] : 0.f;
  }
  __syncthreads();

  // Layers.  Weights are streamed through shared memory in tiles of kNavKT input rows x 256 outputs
  // (next tile's global loads in flight while the current one is consumed): reading them straight from
  // L2 inside the FFMA loop left the kernel waiting ~1,900 cycles per 4 input features (ncu, round 1).
  // A thread owns 8 queries x 4 outputs (og, og+64, og+128, og+192 of the current 256-output block);
  // every accumulator starts from the bias and adds input features in ascending order.
  const int og = tid & 63;           // output lane
  const int qg = tid >> 6;           // query group: queries qg*8 .. qg*8+7
  float* cur = act_a;
  float* nxt = act_b;
  float* wtile = nav_smem + (size_t)(L.width_a + L.width_b) * kNavTQ;   // [2][kNavKT][kNavOB]
  for (int l = 0; l < L.n_layers; ++l) {
    const int I = L.in_dim[l], O = L.out_dim[l];
    const float* __restrict__ wt = L.wt[l];
    const float* __restrict__ bias = L.bias[l];
    const bool relu = (l + 1 < L.n_layers);
    const int n_tiles = (I + kNavKT - 1) / kNavKT;
    for (int ob = 0; ob < O; ob += kNavOB) {
      const int ncols = min(kNavOB, O - ob);
      const int nu = (ncols + 63) / 64;                    // output groups of 64 in use (warp-uniform)
      float acc[4][8];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int o = ob + og + 64 * u;
        const float bv = (o < O) ? bias[o] : 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[u][j] = bv;
      }
      // a tile is kNavKT rows x 64 float4: thread `tid` moves column tid & 63 of rows (tid >> 6) + 2j, j = 0..3, with
      // 16-byte loads and stores (rows of wt are padded to a multiple of 4 floats with zeros)
      constexpr int kVec = kNavKT * (kNavOB / 4) / kNavThreads;          // float4 per thread and tile
      const int P = L.pitch[l];
      const int c4 = (tid & 63) * 4;
      const bool col_ok = ob + c4 < P;
      const float* __restrict__ wcol = wt + ob + c4;
      float4 wreg[kVec];
      auto fetch = [&](int t) {
#pragma unroll
        for (int j = 0; j < kVec; ++j) {
          const int i = t * kNavKT + (tid >> 6) + 2 * j;
          wreg[j] = (col_ok && i < I) ? __ldg(reinterpret_cast<const float4*>(wcol + (size_t)i * P)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      };
      auto stash = [&](int buf) {
#pragma unroll
        for (int j = 0; j < kVec; ++j)
          *reinterpret_cast<float4*>(wtile + (buf * kNavKT + (tid >> 6) + 2 * j) * kNavOB + c4) = wreg[j];
      };
      fetch(0);
      stash(0);
      __syncthreads();
      const float4* a4 = reinterpret_cast<const float4*>(cur + qg * 8);
      for (int t = 0; t < n_tiles; ++t) {
        if (t + 1 < n_tiles) fetch(t + 1);
        const float* w = wtile + (size_t)(t & 1) * kNavKT * kNavOB + og;
        const float4* a = a4 + (size_t)t * kNavKT * (kNavTQ / 4);
        const int rows = min(kNavKT, I - t * kNavKT);
        if (rows == kNavKT) {
          switch (nu) {
            case 1: nav_tile_fma<1, true>(acc, w, a, rows); break;
            case 2: nav_tile_fma<2, true>(acc, w, a, rows); break;
            case 3: nav_tile_fma<3, true>(acc, w, a, rows); break;
            default: nav_tile_fma<4, true>(acc, w, a, rows); break;
          }
        } else {
          switch (nu) {
            case 1: nav_tile_fma<1, false>(acc, w, a, rows); break;
            case 2: nav_tile_fma<2, false>(acc, w, a, rows); break;
            case 3: nav_tile_fma<3, false>(acc, w, a, rows); break;
            default: nav_tile_fma<4, false>(acc, w, a, rows); break;
          }
        }
        if (t + 1 < n_tiles) stash((t + 1) & 1);      // the other buffer: last read two barriers ago
        __syncthreads();
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int o = ob + og + 64 * u;
        if (o < O) {
          if (relu) {
#pragma unroll
            for (int j = 0; j < 8; ++j) acc[u][j] = fmaxf(acc[u][j], 0.f);
          }
          float4* dp = reinterpret_cast<float4*>(nxt + (size_t)o * kNavTQ + qg * 8);
          dp[0] = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
          dp[1] = make_float4(acc[u][4], acc[u][5], acc[u][6], acc[u][7]);
        }
      }
    }
    __syncthreads();
    float* tmp = cur; cur = nxt; nxt = tmp;
  }

  // cur = logits [C][TQ].  softmax statistics + n_top rounds of warp arg-max per query.
  const int C = L.out_dim[L.n_layers - 1];
  const int warp = tid >> 5, lane = tid & 31;
  for (int qi = 0; qi < kNavTQ / (kNavThreads / 32); ++qi) {
    const int q = warp * (kNavTQ / (kNavThreads / 32)) + qi;
    if (q >= nq_here) break;                       // warp-uniform
    // a lane's logits (classes lane, lane+32, ...) are read once into registers when C <= 256: the
    // [class][query] layout makes a warp's column read a 32-way bank conflict, so do it once, not per round
    const bool in_regs = C <= 32 * kNavRegC;
    float lv[kNavRegC];
#pragma unroll
    for (int u = 0; u < kNavRegC; ++u) {
      const int c = lane + 32 * u;
      lv[u] = (in_regs && c < C) ? cur[c * kNavTQ + q] : kNegInf * 2.f;
    }
    float m = kNegInf;
    if (in_regs) {
#pragma unroll
      for (int u = 0; u < kNavRegC; ++u) if (lane + 32 * u < C) m = fmaxf(m, lv[u]);
    } else {
      for (int c = lane; c < C; c += 32) m = fmaxf(m, cur[c * kNavTQ + q]);
    }
    m = warp_max(m);
    float s = 0.f;
    if (in_regs) {
#pragma unroll
      for (int u = 0; u < kNavRegC; ++u) if (lane + 32 * u < C) s += expf(lv[u] - m);
    } else {
      for (int c = lane; c < C; c += 32) s += expf(cur[c * kNavTQ + q] - m);
    }
    s = warp_sum(s);
    for (int t = 0; t < n_top; ++t) {
      float bv = kNegInf; int bc = 0x7fffffff;
      if (in_regs) {
#pragma unroll
        for (int u = 0; u < kNavRegC; ++u) {
          if (lv[u] > bv) { bv = lv[u]; bc = lane + 32 * u; }          // ascending in c: lowest c wins ties
        }
      } else {
        for (int c = lane; c < C; c += 32) {
          const float v = cur[c * kNavTQ + q];
          if (v > bv) { bv = v; bc = c; }            // strided scan is ascending in c: lowest c wins ties
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
        const int oc = __shfl_xor_sync(0xffffffffu, bc, o);
        if (ov > bv || (ov == bv && oc < bc)) { bv = ov; bc = oc; }
      }
      if (bc == 0x7fffffff) bc = 0;                // degenerate (all -inf / NaN): keep indices valid
      if (lane == 0) {
        out_classes[(q0 + q) * n_top + t] = bc;
        if (out_probs) out_probs[(q0 + q) * n_top + t] = expf(bv - m) / s;
      }
      if ((bc & 31) == lane) {                                       // mark taken (-inf)
        if (in_regs) {
#pragma unroll
          for (int u = 0; u < kNavRegC; ++u) if ((bc >> 5) == u) lv[u] = kNegInf * 2.f;
        } else {
          cur[bc * kNavTQ + q] = kNegInf * 2.f;
        }
      }
      __syncwarp();
    }
  }
}

inline size_t nav_smem_bytes(const NavLayers& L) {
  return ((size_t)(L.width_a + L.width_b) * kNavTQ + (size_t)2 * kNavKT * kNavOB) * sizeof(float);
}

}  // namespace lmi
