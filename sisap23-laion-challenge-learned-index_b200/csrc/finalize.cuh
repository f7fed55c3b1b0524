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
constexpr int kFinRegSmall = 8;      // entries per lane of the register-resident selection: <= 256 entries per query,
constexpr int kFinRegE = 24;          // <= 768,
constexpr int kFinRegBig = 48;        // <= 1536 (k' >= 64 instantiation only)
constexpr int kFinMaxVec = 128;       // uint4 per row, d <= 1024
// Error bound of the exactness certificate, per dimension: |cos32 - cos64| <= d * kCertEpsPerDim.  K3's score is
// fl32(acc * inv_norm) with acc the tensor-core sum of d exact fp16 x fp16 products: every addend of a K = 16 step (16
// products + the running sum) is aligned to the largest of them and truncated below 2^-23 of it, so a step loses at most
// 17 * 2^-23 * max|addend| and the whole sum at most (d + d/16) * 2^-23 * sum|q_i x_i| <= 1.0625 * d * 2^-23 * |q||x|
// (Cauchy-Schwarz), i.e. 1.27e-7 * d in cosine units; the fp32 inverse norm and the final product add 2 * 2^-24.
// 1.3e-7 * d = 1.0e-4 at d = 768 (observed maximum on the bench data: 3.7e-6).  A query that fails the test is not wrong,
// it takes the float64 fallback (exact_fixup_kernel).
constexpr double kCertEpsPerDim = 1.3e-7;

__device__ __forceinline__ bool key_less(double da, uint32_t ra, uint32_t ia, int ga,
                                         double db, uint32_t rb, uint32_t ib, int gb) {
  if (da != db) return da < db;
  if (ra != rb) return ra < rb;
  if (ia != ib) return ia < ib;
  return ga < gb;
}

// float64 dot(q, x) and |x|^2 of one candidate row by a group of 8 lanes (sub = lane & 7): lane `sub` takes the
// 16-byte vectors sub, sub + 8, ... in ascending order, then the 8 partial sums are folded with xor-shuffles 4, 2, 1.
// The operation sequence is spelled out with intrinsics so that every caller (K5 and the exact fallback) produces
// bit-identical distances for the same (query, object) pair -- which keeps results independent of the path taken
// and of the number of GPUs.  sq: the query row in shared memory; returns the reduced sums in every lane of the group.
__device__ __forceinline__ void fin_dot_group(const uint4* __restrict__ xrow, const uint4* sq, int vec, int sub,
                                              double& dot_out, double& xx_out) {
  constexpr int kUn = 6;                       // 6 x 16 bytes per lane in flight (the rows are random 1.5 KB HBM reads)
  double dot = 0.0, xx = 0.0;
  for (int v0 = sub; v0 < vec; v0 += 8 * kUn) {
    uint4 ux[kUn];
#pragma unroll
    for (int u = 0; u < kUn; ++u) { const int v = v0 + 8 * u; ux[u] = (v < vec) ? __ldg(xrow + v) : make_uint4(0, 0, 0, 0); }
#pragma unroll
    for (int u = 0; u < kUn; ++u) {
      const int v = v0 + 8 * u;
      if (v < vec) {
        const uint4 uq = sq[v];
        const __half2* hx = reinterpret_cast<const __half2*>(&ux[u]);
        const __half2* hq = reinterpret_cast<const __half2*>(&uq);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 fx = __half22float2(hx[j]);
          const float2 fq = __half22float2(hq[j]);
          const double ax = (double)fx.x, ay = (double)fx.y;
          dot = __dadd_rn(dot, __fma_rn(ay, (double)fq.y, __dmul_rn(ax, (double)fq.x)));
          xx = __dadd_rn(xx, __fma_rn(ay, ay, __dmul_rn(ax, ax)));
        }
      }
    }
  }
#pragma unroll
  for (int o = 4; o > 0; o >>= 1) {                     // sum over the 8 lanes of the group
    dot = __dadd_rn(dot, __shfl_xor_sync(0xffffffffu, dot, o));
    xx = __dadd_rn(xx, __shfl_xor_sync(0xffffffffu, xx, o));
  }
  dot_out = dot; xx_out = xx;
}

// |q| in float64 by one warp (lane takes vectors lane, lane + 32, ...; xor-shuffle fold 16..1) and the row into shared memory
__device__ __forceinline__ double fin_load_query(const uint4* __restrict__ qrow, uint4* sq, int vec, int lane) {
  double qq = 0.0;
  for (int v = lane; v < vec; v += 32) {
    const uint4 u = qrow[v];
    sq[v] = u;
    const __half2* h = reinterpret_cast<const __half2*>(&u);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float2 f = __half22float2(h[j]);
      const double ax = (double)f.x, ay = (double)f.y;
      qq = __dadd_rn(qq, __fma_rn(ay, ay, __dmul_rn(ax, ax)));
    }
  }
  qq = warp_sum(qq);
  return (qq == 0.0) ? 1.0 : sqrt(qq);
}

// order-preserving uint32 image of an fp32 score (0 = invalid entry)
__device__ __forceinline__ uint32_t fin_key(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float fin_unkey(uint32_t u) {
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// Selection of the R best of up to 32 x NE register-resident entries by BISECTION on the score (k' >= 64, i.e.
// k = 100: R rounds of warp arg-max cost R x (NE selects + 10 shuffles); a 32-step binary search for the R-th largest
// key costs 32 x (NE compares + 5 shuffles), then one pass of ballots compacts the survivors).  Ties at the R-th
// score go to the lowest entry number (rank, list, position order), like the round-based selection.
// Returns the number selected; survivors land in s_pos/s_rank/s_dist in entry order; score_R = R-th best score.
template <int NE>
__device__ __forceinline__ int fin_select_bisect(int lane, int R, int total, const int32_t* s_rcum, const int64_t* s_rbeg,
                                                 const Cand* __restrict__ plists, int32_t* s_pos, uint32_t* s_rank,
                                                 double* s_dist, float& score_R) {
  uint32_t key[NE]; int32_t ep[NE];
  int valid = 0;
#pragma unroll
  for (int u = 0; u < NE; ++u) {
    const int g = lane + 32 * u;                   // entry number in (rank, list, position) order
    key[u] = 0u; ep[u] = -1;
    if (g < total) {
      int r = 0;
      while (g >= s_rcum[r + 1]) ++r;
      const Cand c = plists[s_rbeg[r] + (g - s_rcum[r])];
      if (c.pos >= 0) { key[u] = fin_key(c.score); ep[u] = c.pos; ++valid; }
    }
  }
  valid = __reduce_add_sync(0xffffffffu, valid);
  uint32_t T = 1u;                                 // everything valid has key >= 1
  int need_eq = 0x7fffffff;
  if (valid >= R) {                                // (valid == R: T = the smallest key, which the certificate needs)
    T = 0u;
    for (int bit = 31; bit >= 0; --bit) {          // largest T with count(key >= T) >= R
      const uint32_t cand = T | (1u << bit);
      int c = 0;
#pragma unroll
      for (int u = 0; u < NE; ++u) c += (key[u] >= cand) ? 1 : 0;
      c = __reduce_add_sync(0xffffffffu, c);
      if (c >= R) T = cand;
    }
    int gt = 0;
#pragma unroll
    for (int u = 0; u < NE; ++u) gt += (key[u] > T) ? 1 : 0;
    gt = __reduce_add_sync(0xffffffffu, gt);
    need_eq = R - gt;                              // >= 1 entries equal to T complete the selection
  }
  const uint32_t lt_mask = (1u << lane) - 1u;
  int out = 0, eq_seen = 0;
#pragma unroll
  for (int u = 0; u < NE; ++u) {
    const bool is_eq = key[u] == T && key[u] != 0u;
    const uint32_t m_eq = __ballot_sync(0xffffffffu, is_eq);
    const bool sel = key[u] != 0u && (key[u] > T || (is_eq && eq_seen + __popc(m_eq & lt_mask) < need_eq));
    const uint32_t m_sel = __ballot_sync(0xffffffffu, sel);
    if (sel) {
      const int o = out + __popc(m_sel & lt_mask);
      const int g = lane + 32 * u;
      int r = 0;
      while (g >= s_rcum[r + 1]) ++r;
      s_pos[o] = ep[u]; s_rank[o] = (uint32_t)r; s_dist[o] = (double)fin_unkey(key[u]);
    }
    out += __popc(m_sel);
    eq_seen += __popc(m_eq);
  }
  score_R = fin_unkey(T);
  return out;
}

// mode 0: write out_dists/out_ids [nq,k];  mode 1: write Partial [nq,k]
// kBig: instantiation for k' >= 64 (up to 48 register-resident entries per lane in the selection); the k' <= 32
// instantiation stops at 24 entries per lane and is compiled for three CTAs per SM.
template <bool kBig>
__global__ void __launch_bounds__(kFinWarps * 32, kBig ? 2 : 4)
finalize_kernel(int64_t nq, int nb, int k, int KP, int d,
                const int32_t* __restrict__ top_cls, const int32_t* __restrict__ pair_slot,
                const int32_t* __restrict__ qoff, const int32_t* __restrict__ chunks_b,
                const int64_t* __restrict__ plist_base, const Cand* __restrict__ plists,
                const uint16_t* __restrict__ q, const uint16_t* __restrict__ store,
                const uint32_t* __restrict__ ids, const double* __restrict__ threshold,
                PlanScalars* __restrict__ sc, int32_t* __restrict__ uncert_list, int mode,
                double* __restrict__ out_dists, uint32_t* __restrict__ out_ids,
                Partial* __restrict__ out_partial) {
  constexpr int kSel = kBig ? kMaxKPrime : 32;  // survivors per query (R = k')
  __shared__ int32_t s_pos[kFinWarps][kSel];
  __shared__ uint32_t s_rank[kFinWarps][kSel];
  __shared__ uint32_t s_id[kFinWarps][kSel];
  __shared__ double s_dist[kFinWarps][kSel];
  __shared__ int64_t s_rbeg[kFinWarps][32];     // per rank: first entry (global Cand index)
  __shared__ int32_t s_rcum[kFinWarps][33];     // per rank: cumulative entry count
  __shared__ uint4 s_q[kFinWarps][kFinMaxVec];  // the query row (fp16 bits)

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

  // Usual case (one rank group, entries fit the registers of a warp -- e.g. 4 ranks x 3 chunks x k' = 16 = 192
  // entries, 6 per lane): every lane reads its entries once and the R best are found by bisection on the score
  // (32 steps of NE compares + one warp reduction; R rounds of warp arg-max over NE entries -- the round-1 form --
  // were 45 % of this kernel's instructions at k' = 16).  NE is the smallest of 8 / 24 / 48 that holds the entries.
  int sel_ne = 0;                                      // 0: general path below
  float score_R_sel = 0.f;                             // R-th best fp32 score (bisection paths)
  if (n_groups == 1) {
    load_ranges(0);
    const int total = s_rcum[warp][nb];
    if (total <= 32 * kFinRegSmall) sel_ne = kFinRegSmall;
    else if (total <= 32 * kFinRegE) sel_ne = kFinRegE;
    else if (kBig && total <= 32 * kFinRegBig) sel_ne = kFinRegBig;
  }
  if (sel_ne == kFinRegSmall) {
    n_sel = fin_select_bisect<kFinRegSmall>(lane, R, s_rcum[warp][nb], s_rcum[warp], s_rbeg[warp], plists,
                                            s_pos[warp], s_rank[warp], s_dist[warp], score_R_sel);
  } else if (sel_ne == kFinRegE) {
    n_sel = fin_select_bisect<kFinRegE>(lane, R, s_rcum[warp][nb], s_rcum[warp], s_rbeg[warp], plists,
                                        s_pos[warp], s_rank[warp], s_dist[warp], score_R_sel);
  } else if (kBig && sel_ne == kFinRegBig) {
    n_sel = fin_select_bisect<kFinRegBig>(lane, R, s_rcum[warp][nb], s_rcum[warp], s_rbeg[warp], plists,
                                          s_pos[warp], s_rank[warp], s_dist[warp], score_R_sel);
  } else {
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
}
  __syncwarp();

  // ---- float64 re-scoring of the survivors ---------------------------------------------------
  // The candidate rows are random 1.5 KB reads from HBM: 4 candidates are in flight per warp (8 lanes
  // each, 6 x 16 bytes per lane outstanding) instead of one after the other.
  if (n_sel == 0) {
    // nothing to re-score (on a shard most queries probe no local bucket): emit the empty row without touching
    // the query vector or the store
    for (int o = lane; o < k; o += 32) {
      if (mode == 0) { out_dists[qi * k + o] = kEmptyDist; out_ids[qi * k + o] = 0u; }
      else { Partial p; p.dist = kEmptyDist; p.id = 0u; p.rank = 0xffffffffu; out_partial[qi * k + o] = p; }
    }
    return;
  }
  const int vec = d / 8;                                  // uint4 per row (d is a multiple of 64)
  const uint4* qrow = reinterpret_cast<const uint4*>(q + (size_t)qi * d);
  uint4* sq = s_q[warp];
  const double qn = fin_load_query(qrow, sq, vec, lane);
  double max_err = 0.0;
  // (the round-based selections leave the scores in descending order; read before they are overwritten)
  const float score_R = (n_sel == R) ? (sel_ne != 0 ? score_R_sel : (float)s_dist[warp][R - 1]) : 0.f;
  const bool list_full = (n_sel == R);
  __syncwarp();
  // Only candidates that can still reach the top k are re-scored.  With a_k the k-th largest fp32 cosine among the
  // selected and eps the error bound of the certificate, an entry below a_k - 2 eps has a float64 cosine below
  // a_k - eps, which the k entries at or above a_k all reach: it cannot be among the k best.  k' = 16 selected for
  // k = 10 shrink to 10-11 rows read from HBM and re-scored (k' = 128 for k = 100 to ~101); the result is unchanged.
  if (n_sel > k) {
    constexpr int kPer = kSel / 32;                      // selected entries per lane
    uint32_t key[kPer];
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      const int t = lane + 32 * u;
      key[u] = (t < n_sel) ? fin_key((float)s_dist[warp][t]) : 0u;
    }
    uint32_t T = 0u;                                     // k-th largest key: the largest T with count(key >= T) >= k
    for (int bit = 31; bit >= 0; --bit) {
      const uint32_t cand = T | (1u << bit);
      int c = 0;
#pragma unroll
      for (int u = 0; u < kPer; ++u) c += (key[u] >= cand) ? 1 : 0;
      c = __reduce_add_sync(0xffffffffu, c);
      if (c >= k) T = cand;
    }
    const double cut = (double)fin_unkey(T) - 2.0 * kCertEpsPerDim * (double)d * qn;      // scores are cosine x |q|
    int32_t e_pos[kPer]; uint32_t e_rank[kPer]; double e_sc[kPer]; bool e_keep[kPer];
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      const int t = lane + 32 * u;
      e_keep[u] = false; e_pos[u] = 0; e_rank[u] = 0u; e_sc[u] = 0.0;
      if (t < n_sel) {
        e_pos[u] = s_pos[warp][t]; e_rank[u] = s_rank[warp][t]; e_sc[u] = s_dist[warp][t];
        e_keep[u] = e_sc[u] >= cut;
      }
    }
    __syncwarp();
    const uint32_t lt_mask = (1u << lane) - 1u;
    int out = 0;
#pragma unroll
    for (int u = 0; u < kPer; ++u) {
      const uint32_t m = __ballot_sync(0xffffffffu, e_keep[u]);
      if (e_keep[u]) {
        const int o = out + __popc(m & lt_mask);
        s_pos[warp][o] = e_pos[u]; s_rank[warp][o] = e_rank[u]; s_dist[warp][o] = e_sc[u];
      }
      out += __popc(m);
    }
    __syncwarp();
    n_sel = out;                                         // >= k: the k entries at or above a_k are kept
  }
  const int grp = lane >> 3, sub = lane & 7;
  for (int t0 = 0; t0 < n_sel; t0 += 4) {
    const int t = t0 + grp;
    const bool valid = t < n_sel;
    const int32_t pos = s_pos[warp][valid ? t : t0];
    const uint4* xrow = reinterpret_cast<const uint4*>(store + (size_t)pos * d);
    double dot, xx;
    fin_dot_group(xrow, sq, vec, sub, dot, xx);
    if (valid && sub == 0) {
      const double xn = (xx == 0.0) ? 1.0 : sqrt(xx);
      const double cos64 = dot / (qn * xn);
      const double cos32 = (double)(float)s_dist[warp][t] / qn;
      max_err = fmax(max_err, fabs(cos32 - cos64));
      s_dist[warp][t] = 1.0 - cos64;
      s_id[warp][t] = ids[pos];
    }
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
    if (list_full) {
      const double cos_dropped = (double)score_R / qn + kCertEpsPerDim * (double)d;
      const double cos_kth = 1.0 - kth_dist;     // kth_dist = 10000 if fewer than k results -> uncertified
      if (!(cos_dropped < cos_kth)) {
        // not provably exact: listed for exact_fixup_kernel, which re-derives the row from every probed object
        const unsigned long long u = atomicAdd(&sc->n_uncertified, 1ull);
        uncert_list[u] = (int32_t)qi;                 // capacity nq: every query is listed at most once
      }
    }
    // 10k same-address atomics serialise in L2 (~0.15 ms per search): only a new maximum goes there
    const unsigned long long bits = (unsigned long long)__double_as_longlong(max_err);
    if (bits > __ldcg(&sc->max_err_bits)) atomicMax(&sc->max_err_bits, bits);
  }
}

// Exactness fallback.  K3 selects by fp32 scores and K5 certifies, per query, that nothing it dropped can belong to the
// float64 top-k; a query whose certificate fails (near-ties denser than k' - k at the k-th place, or fp32 error above the
// bound) is re-derived here from EVERY object of its probed buckets in float64 -- the reference's own arithmetic
// (utils.py:10-11) over the reference's own candidate set (LearnedIndex.py:143-172) -- so "ids bit-exact" does not rest
// on the certificate's error bound.  Launched after every finalize; exits at once when nothing is listed (the usual
// case).  One CTA per listed query, a group of 8 lanes per object (same arithmetic as K5: fin_dot_group), per-warp
// k-best lists in shared memory, merged by rank counting.
constexpr int kFixWarps = 8;
__global__ void __launch_bounds__(kFixWarps * 32)
exact_fixup_kernel(int64_t nq, int nb, int k, int d, const int32_t* __restrict__ top_cls,
                   const int32_t* __restrict__ pair_slot, const int64_t* __restrict__ bucket_ptr,
                   const uint16_t* __restrict__ q, const uint16_t* __restrict__ store, const uint32_t* __restrict__ ids,
                   const double* __restrict__ threshold, const PlanScalars* __restrict__ sc,
                   const int32_t* __restrict__ uncert_list, int mode,
                   double* __restrict__ out_dists, uint32_t* __restrict__ out_ids, Partial* __restrict__ out_partial) {
  __shared__ double s_d[kFixWarps][kMaxKPrime];
  __shared__ uint32_t s_r[kFixWarps][kMaxKPrime];
  __shared__ uint32_t s_i[kFixWarps][kMaxKPrime];
  __shared__ int s_cnt[kFixWarps];
  __shared__ uint4 s_q[kFinMaxVec];
  __shared__ double s_qn;
  const int n_list = (int)min((unsigned long long)nq, sc->n_uncertified);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = lane >> 3, sub = lane & 7;
  const int vec = d / 8;
  for (int u = blockIdx.x; u < n_list; u += gridDim.x) {
    const int64_t qi = uncert_list[u];
    __syncthreads();                                       // previous query's shared state is no longer read
    if (warp == 0) {
      const double qn = fin_load_query(reinterpret_cast<const uint4*>(q + (size_t)qi * d), s_q, vec, lane);
      if (lane == 0) s_qn = qn;
    }
    __syncthreads();
    const double qn = s_qn;
    int cnt = 0;                                           // entries in this warp's list (warp-uniform)
    int worst = 0;                                         // entry holding the largest key once the list is full
    for (int r = 0; r < nb; ++r) {
      const int64_t e = qi * nb + r;
      if (pair_slot[e] < 0) continue;                      // empty / out-of-range category: contributes nothing
      const int b = top_cls[e];
      const int64_t o0 = bucket_ptr[b], o1 = bucket_ptr[b + 1];
      for (int64_t base = o0 + (int64_t)warp * 4; base < o1; base += kFixWarps * 4) {
        const int64_t pos = base + grp;
        const bool valid = pos < o1;
        double dot, xx;
        fin_dot_group(reinterpret_cast<const uint4*>(store + (size_t)(valid ? pos : o0) * d), s_q, vec, sub, dot, xx);
        const double xn = (xx == 0.0) ? 1.0 : sqrt(xx);
        const double dist = 1.0 - dot / (qn * xn);
        const uint32_t id = valid ? ids[pos] : 0u;
        for (int g4 = 0; g4 < 4; ++g4) {                   // the warp inserts its (up to) 4 candidates one after the other
          if (!__shfl_sync(0xffffffffu, (int)valid, g4 * 8)) continue;
          const double cd = __shfl_sync(0xffffffffu, dist, g4 * 8);
          const uint32_t ci = __shfl_sync(0xffffffffu, id, g4 * 8);
          int slot = -1;
          if (cnt < k) slot = cnt++;
          else if (key_less(cd, (uint32_t)r, ci, 0, s_d[warp][worst], s_r[warp][worst], s_i[warp][worst], 0)) slot = worst;
          if (slot < 0) continue;                          // warp-uniform
          __syncwarp();
          if (lane == 0) { s_d[warp][slot] = cd; s_r[warp][slot] = (uint32_t)r; s_i[warp][slot] = ci; }
          __syncwarp();
          if (cnt == k) {                                  // (re)locate the largest key of the full list
            double wd = -1.0; uint32_t wr = 0, wi = 0; int wj = -1;
            for (int j = lane; j < k; j += 32)
              if (wj < 0 || key_less(wd, wr, wi, 0, s_d[warp][j], s_r[warp][j], s_i[warp][j], 0)) { wd = s_d[warp][j]; wr = s_r[warp][j]; wi = s_i[warp][j]; wj = j; }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
              const double od = __shfl_xor_sync(0xffffffffu, wd, o);
              const uint32_t orr = __shfl_xor_sync(0xffffffffu, wr, o), oi = __shfl_xor_sync(0xffffffffu, wi, o);
              const int oj = __shfl_xor_sync(0xffffffffu, wj, o);
              if (oj >= 0 && (wj < 0 || key_less(wd, wr, wi, 0, od, orr, oi, 0))) { wd = od; wr = orr; wi = oi; wj = oj; }
            }
            worst = wj;
          }
        }
      }
    }
    if (lane == 0) s_cnt[warp] = cnt;
    __syncthreads();
    // ---- merge the warps' lists: the position of an entry = number of entries with a smaller key ----
    const double thr = threshold ? threshold[qi] : 1.0e300;
    int total = 0;
    for (int w = 0; w < kFixWarps; ++w) total += s_cnt[w];
    for (int t = threadIdx.x; t < kFixWarps * k; t += blockDim.x) {
      const int w = t / k, j = t - w * k;
      if (j >= s_cnt[w]) continue;
      const double dt = s_d[w][j]; const uint32_t rt = s_r[w][j], it = s_i[w][j];
      int ord = 0;
      for (int w2 = 0; w2 < kFixWarps; ++w2)
        for (int j2 = 0; j2 < s_cnt[w2]; ++j2)
          ord += key_less(s_d[w2][j2], s_r[w2][j2], s_i[w2][j2], w2 * k + j2, dt, rt, it, t) ? 1 : 0;
      if (ord < k) {
        const bool keep = dt < thr;
        if (mode == 0) { out_dists[qi * k + ord] = keep ? dt : kEmptyDist; out_ids[qi * k + ord] = keep ? it : 0u; }
        else { Partial p; p.dist = keep ? dt : kEmptyDist; p.id = keep ? it : 0u; p.rank = keep ? rt : 0xffffffffu; out_partial[qi * k + ord] = p; }
      }
    }
    for (int o = total + threadIdx.x; o < k; o += blockDim.x) {
      if (mode == 0) { out_dists[qi * k + o] = kEmptyDist; out_ids[qi * k + o] = 0u; }
      else { Partial p; p.dist = kEmptyDist; p.id = 0u; p.rank = 0xffffffffu; out_partial[qi * k + o] = p; }
    }
  }
}

// Multi-GPU: k best of world*k gathered entries per query.  One warp per query: the entries are staged in
// shared memory (up to kMergeSmemE per query) and every lane ranks its own entries against all of them
// (broadcast reads); larger inputs (world * k > kMergeSmemE) rank straight from global memory.
// kSmall: instantiation for world * k <= kMergeSmemE (k = 10 on up to 12 ranks): only the shared-memory path, few
// registers, so every CTA of a 10k-query batch is resident at once -- the merge runs beside the next batch's navigation
// on the exchange stream and should be gone before that batch's rescoring kernel wants every SM.
constexpr int kMergeSmemE = 128;
template <bool kSmall>
__global__ void __launch_bounds__(kFinWarps * 32, kSmall ? 8 : 2)
merge_partials_kernel(const Partial* __restrict__ gathered, int world, int64_t nq, int k,
                      double* __restrict__ out_dists, uint32_t* __restrict__ out_ids) {
  __shared__ Partial s_e[kFinWarps][kMergeSmemE];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t qi = (int64_t)blockIdx.x * kFinWarps + warp;
  if (qi >= nq) return;
  const int E = world * k;
  if (kSmall || E <= kMergeSmemE) {
    int n_valid = 0;
    for (int t0 = 0; t0 < E; t0 += 32) {
      const int t = t0 + lane;
      bool valid = false;
      if (t < E) {
        const int w = t / k;
        const Partial p = gathered[((int64_t)w * nq + qi) * k + (t - w * k)];
        s_e[warp][t] = p;
        valid = p.dist != kEmptyDist;
      }
      n_valid += __popc(__ballot_sync(0xffffffffu, valid));
    }
    __syncwarp();
    // empty slots (10000.0, id 0) sort behind every real entry: when k real entries exist they are not ranked at all
    // (on a bucket shard at most n_buckets of the `world` lists of a query hold anything)
    const bool skip_empty = n_valid >= k;
    for (int t = lane; t < E; t += 32) {
      const Partial pt = s_e[warp][t];
      if (skip_empty && pt.dist == kEmptyDist) continue;
      int ord = 0;
      for (int u = 0; u < E; ++u) {
        const Partial pu = s_e[warp][u];
        ord += key_less(pu.dist, pu.rank, pu.id, u, pt.dist, pt.rank, pt.id, t) ? 1 : 0;
      }
      if (ord < k) { out_dists[qi * k + ord] = pt.dist; out_ids[qi * k + ord] = pt.id; }
    }
    return;
  }
  if (kSmall) return;                              // (not reached: the host picks this instantiation by world * k)
  // Large inputs (k = 100 on 8 ranks: 800 entries).  Up to 1,024 entries: every lane keeps the distance bits of its
  // <= 32 entries in registers, a bisection over the differing bits finds the k-th smallest distance (non-negative
  // doubles order like their bit patterns), one ballot pass compacts the k survivors into shared memory and they
  // are ranked among themselves by the full key -- O(E x bits + k^2) instead of a binary search per entry and list
  // through global memory (0.54 ms per search at k = 100 on 8 ranks).
  constexpr int kMergeRegE = 32;
  if (E <= 32 * kMergeRegE && k <= kMergeSmemE) {
    unsigned long long db[kMergeRegE];
    unsigned long long mn = ~0ull, mx = 0ull;
#pragma unroll
    for (int u = 0; u < kMergeRegE; ++u) {
      const int t = lane + 32 * u;
      db[u] = ~0ull;
      if (t < E) {
        const int w = t / k;
        db[u] = (unsigned long long)__double_as_longlong(gathered[((int64_t)w * nq + qi) * k + (t - w * k)].dist);
        mn = min(mn, db[u]); mx = max(mx, db[u]);
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      mn = min(mn, (unsigned long long)__shfl_xor_sync(0xffffffffu, mn, o));
      mx = max(mx, (unsigned long long)__shfl_xor_sync(0xffffffffu, mx, o));
    }
    // T = k-th smallest distance: the largest T with count(d < T) < k; only bits below the common prefix are searched
    const int top = (mn == mx) ? -1 : 63 - __clzll((long long)(mn ^ mx));
    unsigned long long T = (top >= 63) ? 0ull : (top < 0 ? mn : (mn >> (top + 1)) << (top + 1));
    for (int bit = top; bit >= 0; --bit) {
      const unsigned long long cand = T | (1ull << bit);
      int c = 0;
#pragma unroll
      for (int u = 0; u < kMergeRegE; ++u) c += (db[u] < cand) ? 1 : 0;
      c = __reduce_add_sync(0xffffffffu, c);
      if (c < k) T = cand;
    }
    int less = 0, eq = 0;
#pragma unroll
    for (int u = 0; u < kMergeRegE; ++u) { less += (db[u] < T) ? 1 : 0; eq += (db[u] == T) ? 1 : 0; }
    less = __reduce_add_sync(0xffffffffu, less);
    eq = __reduce_add_sync(0xffffffffu, eq);
    const int need = k - less;
    // equal distances at the cut are resolved by (rank, id): rare for real entries (then the general path below runs);
    // empty slots (10000.0, ~0, 0) are interchangeable, the first `need` in entry order are taken
    const bool cut_is_empty = __longlong_as_double((long long)T) == kEmptyDist;
    if (eq <= need || cut_is_empty) {
      const uint32_t lt_mask = (1u << lane) - 1u;
      int out = 0, eq_seen = 0;
#pragma unroll
      for (int u = 0; u < kMergeRegE; ++u) {
        const bool is_eq = db[u] == T;
        const uint32_t m_eq = __ballot_sync(0xffffffffu, is_eq);
        const bool sel = db[u] < T || (is_eq && eq_seen + __popc(m_eq & lt_mask) < need);
        const uint32_t m_sel = __ballot_sync(0xffffffffu, sel);
        if (sel) {
          const int t = lane + 32 * u, w = t / k;
          s_e[warp][out + __popc(m_sel & lt_mask)] = gathered[((int64_t)w * nq + qi) * k + (t - w * k)];
        }
        out += __popc(m_sel);
        eq_seen += __popc(m_eq);
      }
      __syncwarp();
      const int n_out = min(out, k);               // == k whenever E >= k
      for (int t = lane; t < n_out; t += 32) {
        const Partial pt = s_e[warp][t];
        int ord = 0;
        for (int u = 0; u < n_out; ++u) {
          const Partial pu = s_e[warp][u];
          ord += key_less(pu.dist, pu.rank, pu.id, u, pt.dist, pt.rank, pt.id, t) ? 1 : 0;
        }
        out_dists[qi * k + ord] = pt.dist; out_ids[qi * k + ord] = pt.id;
      }
      return;
    }
  }
  // General path: every rank's list is already sorted by the merge key (finalize emits ascending (dist, rank, id);
  // empty slots (10000, ~0, 0) last), so the position of an entry in the union is its own position plus, per other
  // list, a binary search -- O(E * world * log k).
  for (int t = lane; t < E; t += 32) {
    const int wt = t / k, pt_pos = t - wt * k;
    const Partial pt = gathered[((int64_t)wt * nq + qi) * k + pt_pos];
    int ord = pt_pos;
    for (int w = 0; w < world; ++w) {
      if (w == wt) continue;
      const Partial* lst = gathered + ((int64_t)w * nq + qi) * k;
      int lo = 0, hi = k;                          // first position whose entry is NOT less than pt
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const Partial pu = lst[mid];
        if (key_less(pu.dist, pu.rank, pu.id, w * k + mid, pt.dist, pt.rank, pt.id, t)) lo = mid + 1; else hi = mid;
      }
      ord += lo;
    }
    if (ord < k) { out_dists[qi * k + ord] = pt.dist; out_ids[qi * k + ord] = pt.id; }
  }
}

}  // namespace lmi
