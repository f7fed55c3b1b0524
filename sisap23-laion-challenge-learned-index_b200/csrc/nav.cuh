// nav.cuh -- K1: navigation.  fp32 MLP forward + softmax + per-query top-n bucket select.
//
// Replaces NeuralNetwork.predict_proba (search/li/model.py:214-229: model.eval(), forward,
// softmax(dim=1), prob.topk(C)) and NeuralNetwork.predict (model.py:201-212: argmax of the
// logits).  Arithmetic is plain fp32 FFMA (no TF32, no tensor cores) so bucket order matches
// the reference's torch fp32 model up to summation order.
//
// One CTA = 32 queries.  Activations live in shared memory transposed ([feature][query]) so a
// thread that owns 8 queries x 2 outputs reads its 8 activations with two broadcast LDS.128;
// weights are stored transposed on the device ([in][out]) so a warp reads 32 consecutive
// outputs of one input feature as a single 128-byte line (L2 resident: <= a few 100 KB).
#pragma once
#include "common.cuh"

namespace lmi {

constexpr int kNavTQ = 32;          // queries per CTA
constexpr int kNavThreads = 256;
constexpr int kNavMaxLayers = 8;

struct NavLayers {
  int n_layers;
  int in_dim[kNavMaxLayers];
  int out_dim[kNavMaxLayers];
  const float* wt[kNavMaxLayers];    // [in][out]  (transposed nn.Linear.weight)
  const float* bias[kNavMaxLayers];  // [out]
  int max_width;                     // max over all in/out dims
};

// x: [n, d_nav] fp32.  out_classes: int32 [n, n_top]; out_probs: fp32 [n, n_top] or nullptr.
__global__ void __launch_bounds__(kNavThreads, 2)
nav_kernel(const float* __restrict__ x, int64_t n, NavLayers L, int n_top,
           int32_t* __restrict__ out_classes, float* __restrict__ out_probs) {
  extern __shared__ float nav_smem[];
  float* act_a = nav_smem;
  float* act_b = nav_smem + (size_t)L.max_width * kNavTQ;
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

  const int og = tid & 63;           // output lane
  const int qg = tid >> 6;           // query group: queries qg*8 .. qg*8+7
  float* cur = act_a;
  float* nxt = act_b;
  for (int l = 0; l < L.n_layers; ++l) {
    const int I = L.in_dim[l], O = L.out_dim[l];
    const float* __restrict__ wt = L.wt[l];
    const float* __restrict__ bias = L.bias[l];
    const bool relu = (l + 1 < L.n_layers);
    for (int o = og; o < O; o += 128) {
      const int o1 = o + 64;
      const bool has1 = o1 < O;
      float acc0[8], acc1[8];
      const float b0 = bias[o], b1 = has1 ? bias[o1] : 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) { acc0[j] = b0; acc1[j] = b1; }
      const float4* a4 = reinterpret_cast<const float4*>(cur + qg * 8);
#pragma unroll 4
      for (int i = 0; i < I; ++i) {
        const float w0 = __ldg(wt + (size_t)i * O + o);
        const float w1 = has1 ? __ldg(wt + (size_t)i * O + o1) : 0.f;
        const float4 lo = a4[i * (kNavTQ / 4)];
        const float4 hi = a4[i * (kNavTQ / 4) + 1];
        acc0[0] = fmaf(lo.x, w0, acc0[0]); acc0[1] = fmaf(lo.y, w0, acc0[1]);
        acc0[2] = fmaf(lo.z, w0, acc0[2]); acc0[3] = fmaf(lo.w, w0, acc0[3]);
        acc0[4] = fmaf(hi.x, w0, acc0[4]); acc0[5] = fmaf(hi.y, w0, acc0[5]);
        acc0[6] = fmaf(hi.z, w0, acc0[6]); acc0[7] = fmaf(hi.w, w0, acc0[7]);
        acc1[0] = fmaf(lo.x, w1, acc1[0]); acc1[1] = fmaf(lo.y, w1, acc1[1]);
        acc1[2] = fmaf(lo.z, w1, acc1[2]); acc1[3] = fmaf(lo.w, w1, acc1[3]);
        acc1[4] = fmaf(hi.x, w1, acc1[4]); acc1[5] = fmaf(hi.y, w1, acc1[5]);
        acc1[6] = fmaf(hi.z, w1, acc1[6]); acc1[7] = fmaf(hi.w, w1, acc1[7]);
      }
      float4* d0p = reinterpret_cast<float4*>(nxt + (size_t)o * kNavTQ + qg * 8);
      if (relu) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc0[j] = fmaxf(acc0[j], 0.f); acc1[j] = fmaxf(acc1[j], 0.f); }
      }
      d0p[0] = make_float4(acc0[0], acc0[1], acc0[2], acc0[3]);
      d0p[1] = make_float4(acc0[4], acc0[5], acc0[6], acc0[7]);
      if (has1) {
        float4* d1p = reinterpret_cast<float4*>(nxt + (size_t)o1 * kNavTQ + qg * 8);
        d1p[0] = make_float4(acc1[0], acc1[1], acc1[2], acc1[3]);
        d1p[1] = make_float4(acc1[4], acc1[5], acc1[6], acc1[7]);
      }
    }
    __syncthreads();
    float* t = cur; cur = nxt; nxt = t;
  }

  // cur = logits [C][TQ].  softmax statistics + n_top rounds of warp arg-max per query.
  const int C = L.out_dim[L.n_layers - 1];
  const int warp = tid >> 5, lane = tid & 31;
  for (int qi = 0; qi < kNavTQ / (kNavThreads / 32); ++qi) {
    const int q = warp * (kNavTQ / (kNavThreads / 32)) + qi;
    if (q >= nq_here) break;                       // warp-uniform
    float m = kNegInf;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, cur[c * kNavTQ + q]);
    m = warp_max(m);
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s += expf(cur[c * kNavTQ + q] - m);
    s = warp_sum(s);
    for (int t = 0; t < n_top; ++t) {
      float bv = kNegInf; int bc = 0x7fffffff;
      for (int c = lane; c < C; c += 32) {
        const float v = cur[c * kNavTQ + q];
        if (v > bv) { bv = v; bc = c; }            // strided scan is ascending in c: lowest c wins ties
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
      if ((bc & 31) == lane) cur[bc * kNavTQ + q] = kNegInf * 2.f;   // mark taken (-inf)
      __syncwarp();
    }
  }
}

inline size_t nav_smem_bytes(const NavLayers& L) {
  return (size_t)2 * L.max_width * kNavTQ * sizeof(float);
}

}  // namespace lmi
