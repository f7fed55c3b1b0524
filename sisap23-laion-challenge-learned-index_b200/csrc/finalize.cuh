// finalize.cuh -- K4 + K5: merge the partial top-k' lists of a query, re-score the survivors in
// float64 the way scikit-learn does, and emit the reference's result rows.
//
// Replaces the rank merge of LearnedIndex.search (search/li/LearnedIndex.py:88-97: hstack +
// stable argsort + fancy index) and pins the arithmetic of the returned distances to the
// reference's: ``1 - cosine_similarity`` in float64 on the float16 inputs (utils.py:10-11;
// scikit-learn upcasts to float64, divides by the L2 norm with zero norms -> 1, then a dot).
// Order: ascending distance; ties -> lower bucket rank (strict '<' at utils.py:23 + stable merge at
// LearnedIndex.py:91), then ascending id.  Unfilled slots are (id 0, dist 10000.0)
// (LearnedIndex.py:137-138).
//
// One warp per query.  The partial lists of one (query, rank) are contiguous in memory (list id =
// plist_base[b] + slot_in_bucket * chunks_b + chunk), a few KB per query, so the R selection
// rounds re-read them from L1.
#pragma once
#include "common.cuh"
#include "plan.cuh"

namespace lmi {

constexpr int kFinWarps = 8;
constexpr double kCertEps = 2e-5;       // bound used by the exactness certificate

__device__ __forceinline__ bool key_less(double da, uint32_t ra, uint32_t ia, int ga,
                                         double db, uint32_t rb, uint32_t ib, int gb) {
  if (da != db) return da < db;
  if (ra != rb) return ra < rb;
  if (ia != ib) return ia < ib;
  return ga < gb;
}

// mode 0: write out_dists/out_ids [nq,k];  mode 1: write Partial [nq,k]
__global__ void __launch_bounds__(kFinWarps * 32)
finalize_kernel(int64_t nq, int nb, int k, int KP, int d,
                const int32_t* __restrict__ top_cls, const int32_t* __restrict__ pair_slot,
                const int32_t* __restrict__ qoff, const int32_t* __restrict__ chunks_b,
                const int64_t* __restrict__ plist_base, const Cand* __restrict__ plists,
                const uint16_t* __restrict__ q, const uint16_t* __restrict__ store,
                const uint32_t* __restrict__ ids, const double* __restrict__ threshold,
                PlanScalars* __restrict__ sc, int mode,
                double* __restrict__ out_dists, uint32_t* __restrict__ out_ids,
                Partial* __restrict__ out_partial) {
  __shared__ int32_t s_pos[kFinWarps][kMaxKPrime];
  __shared__ uint32_t s_rank[kFinWarps][kMaxKPrime];
  __shared__ uint32_t s_id[kFinWarps][kMaxKPrime];
  __shared__ double s_dist[kFinWarps][kMaxKPrime];
  __shared__ int64_t s_rbeg[kFinWarps][32];     // per rank: first entry (global Cand index)
  __shared__ int32_t s_rcum[kFinWarps][33];     // per rank: cumulative entry count

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t qi = (int64_t)blockIdx.x * kFinWarps + warp;
  if (qi >= nq) return;
  const int R = KP;

  // ---- entry ranges of this query ---------------------------------------------------------
  // (nb may exceed 32: ranks are processed in groups of 32; selection state carries over)
  // For simplicity and because nb <= C is small, ranges are recomputed per group.
  int n_sel = 0;
  float prev_s = 3.0e38f;
  long long prev_g = -1;
  const int n_groups = (nb + 31) / 32;

  auto load_ranges = [&](int grp) {
    __syncwarp();
    const int r = grp * 32 + lane;
    int cnt = 0;
    int64_t beg = 0;
    if (r < nb) {
      const int64_t e = qi * nb + r;
      const int slot = pair_slot[e];
      if (slot >= 0) {
        const int b = top_cls[e];
        const int ch = chunks_b[b];
        beg = (plist_base[b] + (int64_t)(slot - qoff[b]) * ch) * KP;
        cnt = ch * KP;
      }
    }
    int inc = cnt;                                     // inclusive scan of cnt over the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
    s_rbeg[warp][lane] = beg;
    s_rcum[warp][lane + 1] = inc;
    if (lane == 0) s_rcum[warp][0] = 0;
    __syncwarp();
  };

  for (int t = 0; t < R; ++t) {
    float best_s = kNegInf * 2.f;      // -inf
    long long best_g = 0x7fffffffffffffffLL;
    int32_t best_pos = -1;
    uint32_t best_rank = 0;
    long long gbase = 0;
    for (int grp = 0; grp < n_groups; ++grp) {
      if (n_groups > 1 || t == 0) load_ranges(grp);     // one group (nb <= 32): ranges loaded once
      __syncwarp();
      const int nr = min(32, nb - grp * 32);
      const int total = s_rcum[warp][nr];
      for (int r = 0; r < nr; ++r) {
        const int c0 = s_rcum[warp][r], c1 = s_rcum[warp][r + 1];
        const Cand* base = plists + s_rbeg[warp][r];
        for (int j = lane; j < c1 - c0; j += 32) {
          const Cand c = base[j];
          if (c.pos < 0) continue;
          const long long g = gbase + c0 + j;
          const bool after_prev = (c.score < prev_s) || (c.score == prev_s && g > prev_g);
          if (!after_prev) continue;
          if (c.score > best_s || (c.score == best_s && g < best_g)) {
            best_s = c.score; best_g = g; best_pos = c.pos; best_rank = (uint32_t)(grp * 32 + r);
          }
        }
      }
      gbase += total;
    }
    // warp arg-max by (score desc, g asc)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float os = __shfl_xor_sync(0xffffffffu, best_s, o);
      const long long og = __shfl_xor_sync(0xffffffffu, best_g, o);
      const int32_t op = __shfl_xor_sync(0xffffffffu, best_pos, o);
      const uint32_t orr = __shfl_xor_sync(0xffffffffu, best_rank, o);
      if (os > best_s || (os == best_s && og < best_g)) { best_s = os; best_g = og; best_pos = op; best_rank = orr; }
    }
    if (best_pos < 0) break;           // no more valid entries (warp-uniform)
    if (lane == 0) {
      s_pos[warp][t] = best_pos;
      s_rank[warp][t] = best_rank;
      s_dist[warp][t] = (double)best_s;   // temporarily the fp32 score
    }
    prev_s = best_s; prev_g = best_g;
    n_sel = t + 1;
  }
  __syncwarp();

  // ---- float64 re-scoring of the survivors ---------------------------------------------------
  const int vec = d / 8;                                  // uint4 per row
  const uint4* qrow = reinterpret_cast<const uint4*>(q + (size_t)qi * d);
  double qq = 0.0;
  for (int v = lane; v < vec; v += 32) {
    const uint4 u = qrow[v];
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) { const float2 f = __half22float2(h[j]); qq += (double)f.x * f.x + (double)f.y * f.y; }
  }
  qq = warp_sum(qq);
  const double qn = (qq == 0.0) ? 1.0 : sqrt(qq);
  double max_err = 0.0;
  float score_R = 0.f;
  for (int t = 0; t < n_sel; ++t) {
    const int32_t pos = s_pos[warp][t];
    const uint4* xrow = reinterpret_cast<const uint4*>(store + (size_t)pos * d);
    double dot = 0.0, xx = 0.0;
    for (int v = lane; v < vec; v += 32) {
      const uint4 ux = xrow[v];
      const uint4 uq = qrow[v];
      const __half2* hx = reinterpret_cast<const __half2*>(&ux);
      const __half2* hq = reinterpret_cast<const __half2*>(&uq);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 fx = __half22float2(hx[j]);
        const float2 fq = __half22float2(hq[j]);
        dot += (double)fx.x * (double)fq.x + (double)fx.y * (double)fq.y;
        xx += (double)fx.x * (double)fx.x + (double)fx.y * (double)fx.y;
      }
    }
    dot = warp_sum(dot);
    xx = warp_sum(xx);
    const double xn = (xx == 0.0) ? 1.0 : sqrt(xx);
    const double cos64 = dot / (qn * xn);
    const float sc32 = (float)s_dist[warp][t];
    const double cos32 = (double)sc32 / qn;
    max_err = fmax(max_err, fabs(cos32 - cos64));
    if (t == R - 1) score_R = sc32;
    __syncwarp();
    if (lane == 0) { s_dist[warp][t] = 1.0 - cos64; s_id[warp][t] = ids[pos]; }
  }
  __syncwarp();

  // ---- order by (dist, rank, id) and emit -----------------------------------------------------
  const double thr = threshold ? threshold[qi] : 1.0e300;
  double kth_dist = kEmptyDist;
  for (int t = lane; t < n_sel; t += 32) {
    const double dt = s_dist[warp][t];
    const uint32_t rt = s_rank[warp][t], it = s_id[warp][t];
    int ord = 0;
    for (int u = 0; u < n_sel; ++u)
      ord += key_less(s_dist[warp][u], s_rank[warp][u], s_id[warp][u], u, dt, rt, it, t) ? 1 : 0;
    if (ord < k) {
      const bool keep = dt < thr;
      if (mode == 0) {
        out_dists[qi * k + ord] = keep ? dt : kEmptyDist;
        out_ids[qi * k + ord] = keep ? it : 0u;
      } else {
        Partial p; p.dist = keep ? dt : kEmptyDist; p.id = keep ? it : 0u; p.rank = keep ? rt : 0xffffffffu;
        out_partial[qi * k + ord] = p;
      }
    }
    if (ord == k - 1) kth_dist = dt;
  }
  for (int o = n_sel + lane; o < k; o += 32) {
    if (mode == 0) { out_dists[qi * k + o] = kEmptyDist; out_ids[qi * k + o] = 0u; }
    else { Partial p; p.dist = kEmptyDist; p.id = 0u; p.rank = 0xffffffffu; out_partial[qi * k + o] = p; }
  }

  // ---- exactness certificate + error statistic -------------------------------------------------
  // Everything dropped (by a full k' list in K3 or by the top-R cut above) has fp32 score <= score_R.
  // If even that, plus the error bound, stays below the k-th float64 cosine, no dropped object can
  // belong to the true top-k.
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    kth_dist = fmin(kth_dist, __shfl_xor_sync(0xffffffffu, kth_dist, o));
    max_err = fmax(max_err, __shfl_xor_sync(0xffffffffu, max_err, o));
  }
  if (lane == 0) {
    if (n_sel == R) {
      const double cos_dropped = (double)score_R / qn + kCertEps;
      const double cos_kth = 1.0 - kth_dist;     // kth_dist = 10000 if fewer than k results -> uncertified
      if (!(cos_dropped < cos_kth)) atomicAdd(&sc->n_uncertified, 1ull);
    }
    atomicMax(&sc->max_err_bits, (unsigned long long)__double_as_longlong(max_err));
  }
}

// Multi-GPU: k best of world*k gathered entries per query.  One warp per query.
__global__ void __launch_bounds__(kFinWarps * 32)
merge_partials_kernel(const Partial* __restrict__ gathered, int world, int64_t nq, int k,
                      double* __restrict__ out_dists, uint32_t* __restrict__ out_ids) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t qi = (int64_t)blockIdx.x * kFinWarps + warp;
  if (qi >= nq) return;
  const int E = world * k;
  for (int t = lane; t < E; t += 32) {
    const Partial pt = gathered[((int64_t)(t / k) * nq + qi) * k + (t % k)];
    int ord = 0;
    for (int u = 0; u < E; ++u) {
      const Partial pu = gathered[((int64_t)(u / k) * nq + qi) * k + (u % k)];
      ord += key_less(pu.dist, pu.rank, pu.id, u, pt.dist, pt.rank, pt.id, t) ? 1 : 0;
    }
    if (ord < k) { out_dists[qi * k + ord] = pt.dist; out_ids[qi * k + ord] = pt.id; }
  }
}

}  // namespace lmi
