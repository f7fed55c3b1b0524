// plan.cuh -- K2: invert per-query bucket lists into per-bucket query lists and build the
// rescoring work list, entirely on the device (no host round trip inside a search).
//
// Replaces the scan ``cat_idxs = np.where(pred_categories == cat)`` that the reference repeats
// for every (rank, category) (search/li/LearnedIndex.py:143-144): one counting sort of the
// nq*nb (query, rank) pairs by bucket.  The order inside a bucket is ascending (query, rank),
// so the layout -- and with it every downstream tie-break -- is deterministic.
#pragma once
#include "common.cuh"

namespace lmi {

struct PlanScalars {          // device-resident scalars of one search
  int32_t n_items;            // work items
  int32_t n_slots;            // = sum_b m_b (pairs whose bucket is non-empty on this shard may be fewer)
  int64_t n_plists;           // partial lists
  int64_t pairs;              // sum_b m_b * n_b
  int64_t bucket_objs;        // sum_{b: m_b>0} n_b
  unsigned long long n_uncertified;
  unsigned long long max_err_bits;   // float64 bits of max |cos32-cos64| (atomicMax on non-negative doubles)
  int32_t work_counter;       // dynamic scheduler cursor
  int32_t pad;
  int64_t padded_pairs;       // sum over items of 128 x ceil64(objects): what the 128 x 64 MMA tiles really compute
  unsigned long long k3_clk;  // rescoring kernel, CTA 0: SM cycles (clock64) and nanoseconds (globaltimer) from
  unsigned long long k3_ns;   //   start to end -> the SM clock the kernel actually ran at
};

// m_b histogram.  top_cls: int32 [nq*nb]; entries outside [0,C) are ignored.
__global__ void plan_count_kernel(const int32_t* __restrict__ top_cls, int64_t n_pairs, int C,
                                  int32_t* __restrict__ m_b) {
  for (int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; e < n_pairs;
       e += (int64_t)gridDim.x * blockDim.x) {
    const int c = top_cls[e];
    if (c >= 0 && c < C) atomicAdd(&m_b[c], 1);
  }
}

// Single CTA: prefix sums over buckets + work item expansion.
//   bucket_ptr: int64 [C+1] CSR of the local store;  chunk: objects per work item.
//   outputs: qoff int32 [C+1]; chunks_b int32 [C]; plist_base int64 [C]; items; scalars.
__global__ void __launch_bounds__(1024)
plan_build_kernel(const int32_t* __restrict__ m_b, const int64_t* __restrict__ bucket_ptr, int C,
                  int chunk, int32_t* __restrict__ qoff, int32_t* __restrict__ chunks_b,
                  int64_t* __restrict__ plist_base, WorkItem* __restrict__ items, int max_items,
                  PlanScalars* __restrict__ sc, int pair_mode = 0) {
  __shared__ long long s_scan[1024];
  __shared__ long long s_carry[4];
  __shared__ long long s_total_full;
  const int tid = threadIdx.x;
  if (tid < 4) s_carry[tid] = 0;
  long long tot_pairs = 0, tot_objs = 0, tot_padded = 0;
  // Pair mode lists every FULL pair (two 128-row tiles, 1,536 tensor cycles per object tile) before any HALF pair
  // (an odd last tile split in two, ~half the time per tile): the persistent grid takes entries in list order, and
  // long items first / short items last packs it better than bucket order (list-scheduling simulation on the bench's
  // bucket sizes: makespan 1.17 -> 1.10 x ideal on a 1/8 shard, 1.06 -> 1.03 unsharded).  Full pairs of one chunk
  // stay adjacent (they stream it in lock-step and share it through L2); a half pair runs twice as fast as its
  // chunk's full pairs and drifts out of their L2 window after a few dozen tiles wherever it is listed.
  // Pass 1: number of full-pair entries over all buckets = where the half pairs start.
  {
    long long my_full = 0;
    if (pair_mode == 1 || pair_mode == 3) {   // (pair_mode 2 = bucket-major order, kept for A/B runs: option "item_order")
      for (int b = tid; b < C; b += 1024) {
        const long long nobj = bucket_ptr[b + 1] - bucket_ptr[b];
        const long long m = (nobj > 0) ? m_b[b] : 0;
        const long long ch = (nobj + chunk - 1) / chunk;
        const long long qt = (m + kTileQ - 1) / kTileQ;
        if (!(pair_mode == 3 && qt == 3)) my_full += 2 * ch * (qt / 2);
      }
    }
    s_scan[tid] = my_full;
    __syncthreads();
    for (int off = 512; off > 0; off >>= 1) { if (tid < off) s_scan[tid] += s_scan[tid + off]; __syncthreads(); }
    if (tid == 0) s_total_full = s_scan[0];
  }
  __syncthreads();
  // three running exclusive scans, processed in blocks of 1024 buckets:
  //   0: slots (m_b)   1: items (qtiles*chunks)   2: plists (m_b*chunks)
  for (int base = 0; base < C; base += 1024) {
    const int b = base + tid;
    long long m = 0, nobj = 0, ch = 0, qt = 0, csize = chunk;
    if (b < C) {
      nobj = bucket_ptr[b + 1] - bucket_ptr[b];
      m = (nobj > 0) ? m_b[b] : 0;                 // empty bucket: contributes nothing (LearnedIndex.py:147)
      ch = (nobj + chunk - 1) / chunk;
      // equal-sized chunks (multiples of 64 objects) instead of `chunk`-sized ones plus a short tail:
      // uniform work items, and no near-empty item that still pays for loading its query tile
      if (ch > 0) csize = min((long long)chunk, ((nobj + ch - 1) / ch + 63) / 64 * 64);
      qt = (m + kTileQ - 1) / kTileQ;
      chunks_b[b] = (int)ch;
      tot_pairs += m * nobj;
      if (m > 0) tot_objs += nobj;
      if (ch > 0) {
        const long long last = nobj - (ch - 1) * csize;
        // (pair mode: an odd last tile runs as a half-tile pair with M = 128 MMAs = the tensor time of one tile)
        tot_padded += qt * kTileQ * ((ch - 1) * csize + (last + 63) / 64 * 64);
      }
    }
    // work list entries come in PAIRS (entries 2i, 2i+1 = one item of a 2-CTA cluster of the
    // TMEM-resident rescoring kernel; the other kernels treat every entry as an independent item):
    //   shared pair : query tiles 2p and 2p+1 of the same object chunk  (flag kItemShared: the
    //                 cluster loads the chunk once and multicasts it to both CTAs)
    //   plain pair  : the odd last query tile of chunks c and c+1 (or a dummy entry)
    // pair mode (rescore_ts2.cuh, CTA-pair MMAs: a cluster can only run shared pairs): ceil(qt / 2) shared
    //   pairs per chunk; an odd last query tile is split into two half tiles (<= 64 rows each, kItemHalf) of the same chunk
    const long long pairs_per_chunk = pair_mode ? (qt + 1) / 2 : qt / 2;
    const long long n_pair = pair_mode ? ch * pairs_per_chunk : ch * pairs_per_chunk + ((qt & 1) ? (ch + 1) / 2 : 0);
    // scans: 0 slots, 1 entries (pair mode: full-pair entries), 2 partial lists, 3 half-pair entries (pair mode)
    // pair_mode 3 (experiment, option item_order = 2): as 1, but a bucket with exactly three query tiles runs as THREE half
    // pairs per chunk (all M = 128 MMAs, equal speed, adjacent in the list -> one DRAM read of the chunk shared through L2)
    // instead of one full pair + one half pair that run at 1 : 2 and read the chunk twice
    const bool all_half = (pair_mode == 3 && qt == 3);
    const bool split_order = (pair_mode == 1 || pair_mode == 3);
    const long long n_full_e = all_half ? 0 : split_order ? 2 * ch * (qt / 2) : 2 * n_pair;
    const long long n_half_e = all_half ? 2 * ch * qt : split_order ? 2 * ch * (qt & 1) : 0;
    long long vals[4] = {m, n_full_e, m * ch, n_half_e};
    long long excl[4];
    for (int w = 0; w < 4; ++w) {
      s_scan[tid] = vals[w];
      __syncthreads();
      for (int off = 1; off < 1024; off <<= 1) {
        long long t = (tid >= off) ? s_scan[tid - off] : 0;
        __syncthreads();
        s_scan[tid] += t;
        __syncthreads();
      }
      excl[w] = s_carry[w] + s_scan[tid] - vals[w];
      __syncthreads();
      if (tid == 1023) s_carry[w] += s_scan[1023];
      __syncthreads();
    }
    if (b < C) {
      qoff[b] = (int)excl[0];
      plist_base[b] = excl[2];
      const long long obj0 = bucket_ptr[b];
      long long it = excl[1];
      long long it_half = s_total_full + excl[3];
      // order: for every two chunks, their shared pairs first, then the plain pair of their odd
      // tiles -- consecutive items are picked up by clusters that become free one after the other,
      // so everything that streams one chunk runs at about the same time and all but the first
      // reader find it in L2.
      auto emit = [&](long long c, long long t, int flags) {
        if (it < max_items) {
          WorkItem wi;
          if (c < 0) {                                     // dummy
            wi.slot_base = 0; wi.rows = 0; wi.obj_begin = (int)obj0; wi.obj_count = 0;
            wi.plist_base = 0; wi.plist_stride = 0; wi.bucket = b; wi.flags = 0;
          } else {
            wi.slot_base = (int)excl[0] + (int)t * kTileQ;
            wi.rows = (int)min((long long)kTileQ, m - t * kTileQ);
            wi.obj_begin = (int)(obj0 + c * csize);
            wi.obj_count = (int)min(csize, nobj - c * csize);
            wi.plist_base = excl[2] + t * kTileQ * ch + c;
            wi.plist_stride = (int)ch;
            wi.bucket = b;
            wi.flags = flags;
          }
          items[it] = wi;
        }
        ++it;
      };
      // one entry with an explicit row range [row0, row0 + rows) of the bucket's query list
      auto emit_rows = [&](long long& it, long long c, long long row0, long long rows, int flags) {
        if (it < max_items) {
          WorkItem wi;
          wi.slot_base = (int)excl[0] + (int)row0;
          wi.rows = (int)rows;
          wi.obj_begin = (int)(obj0 + c * csize);
          wi.obj_count = (int)min(csize, nobj - c * csize);
          wi.plist_base = excl[2] + row0 * ch + c;
          wi.plist_stride = (int)ch;
          wi.bucket = b;
          wi.flags = flags;
          wi.pad = 0;
          items[it] = wi;
        }
        ++it;
      };
      if (all_half) {
        for (long long c = 0; c < ch; ++c)
          for (long long t = 0; t < qt; ++t) {
            const long long r0 = t * kTileQ, r = min((long long)kTileQ, m - r0), h = (r + 1) / 2;
            const int lastf = (t == qt - 1) ? kItemLastReader : 0;          // the three run at one speed and share the chunk
            emit_rows(it_half, c, r0, h, kItemShared | kItemHalf | lastf);
            emit_rows(it_half, c, r0 + h, r - h, kItemShared | kItemHalf | lastf);
          }
      } else if (pair_mode) {
        for (long long c = 0; c < ch; ++c)
          for (long long p = 0; p < pairs_per_chunk; ++p) {
            const long long r0 = 2 * p * kTileQ;
            if (2 * p + 1 < qt) {
              // kItemLastReader: the last full pair of a chunk (the earlier ones run at the same speed just ahead of it and
              // leave the chunk in L2; nothing that could still find it there comes after)
              // bucket-major order: the half pair of an odd bucket is the next pair of the list (lock-step option of the kernel)
              const int lastf = (p == qt / 2 - 1) ? (kItemLastReader | ((!split_order && (qt & 1)) ? kItemLeads : 0)) : 0;
              emit_rows(it, c, r0, kTileQ, kItemShared | lastf);
              emit_rows(it, c, r0 + kTileQ, min((long long)kTileQ, m - r0 - kTileQ), kItemShared | lastf);
            } else {                                           // odd last tile: two halves (the second may be empty)
              const long long r = m - r0, h = (r + 1) / 2;
              long long& ih = split_order ? it_half : it;      // (runs at twice the speed of the full pairs unless it follows in lock-step)
              const int fol = (!split_order && qt >= 3) ? kItemFollows : 0;
              emit_rows(ih, c, r0, h, kItemShared | kItemHalf | kItemLastReader | fol);
              emit_rows(ih, c, r0 + h, r - h, kItemShared | kItemHalf | kItemLastReader | fol);
            }
          }
      } else {
        for (long long c0 = 0; c0 < ch; c0 += 2) {
          for (long long c = c0; c < min(ch, c0 + 2); ++c)
            for (long long p = 0; p < pairs_per_chunk; ++p) { emit(c, 2 * p, kItemShared); emit(c, 2 * p + 1, kItemShared); }
          if (qt & 1) { emit(c0, qt - 1, 0); emit(c0 + 1 < ch ? c0 + 1 : -1, qt - 1, 0); }
        }
      }
    }
    __syncthreads();
  }
  // reduce the two totals
  s_scan[tid] = tot_pairs;
  __syncthreads();
  for (int off = 512; off > 0; off >>= 1) { if (tid < off) s_scan[tid] += s_scan[tid + off]; __syncthreads(); }
  const long long pairs = s_scan[0];
  __syncthreads();
  s_scan[tid] = tot_padded;
  __syncthreads();
  for (int off = 512; off > 0; off >>= 1) { if (tid < off) s_scan[tid] += s_scan[tid + off]; __syncthreads(); }
  const long long padded = s_scan[0];
  __syncthreads();
  s_scan[tid] = tot_objs;
  __syncthreads();
  for (int off = 512; off > 0; off >>= 1) { if (tid < off) s_scan[tid] += s_scan[tid + off]; __syncthreads(); }
  if (tid == 0) {
    qoff[C] = (int)s_carry[0];
    sc->n_slots = (int)s_carry[0];
    sc->n_items = (int)min((long long)max_items, s_carry[1] + s_carry[3]);
    sc->n_plists = s_carry[2];
    sc->pairs = pairs;
    sc->bucket_objs = s_scan[0];
    sc->n_uncertified = 0ull;
    sc->max_err_bits = 0ull;
    sc->work_counter = 0;
    sc->padded_pairs = padded;
    sc->k3_clk = 0ull;
    sc->k3_ns = 0ull;
  }
}

// One CTA per bucket: ordered compaction of the (query, rank) pairs routed to it.
//   slot_pair int32 [n_slots]: pair index e = q*nb + r of every slot (bucket-major, e ascending)
//   pair_slot int32 [nq*nb]  : slot of every pair, -1 if its bucket is empty / out of range
constexpr int kScatterThreads = 1024;
constexpr int kScatterPer = 4;                 // consecutive pairs per thread and round
__global__ void __launch_bounds__(kScatterThreads)
plan_scatter_kernel(const int32_t* __restrict__ top_cls, int64_t n_pairs, int C,
                    const int32_t* __restrict__ qoff, const int64_t* __restrict__ bucket_ptr,
                    int32_t* __restrict__ slot_pair, int32_t* __restrict__ pair_slot) {
  // Every CTA scans the whole class list (160 KB at 10k x 4, L2 resident) and keeps the pairs of its
  // bucket in ascending pair order.  The scan is latency-bound, so a round covers 4096 pairs (4 per
  // thread, the next round's loads issued before this round's block scan) -- 10 rounds at 10k x 4.
  const int b = blockIdx.x;
  __shared__ int s_warp[2][kScatterThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool live = (bucket_ptr[b + 1] - bucket_ptr[b]) > 0;
  int base = qoff[b];
  auto load4 = [&](int64_t e0, int (&c)[kScatterPer]) {
#pragma unroll
    for (int i = 0; i < kScatterPer; ++i) c[i] = (e0 + i < n_pairs) ? top_cls[e0 + i] : -1;
  };
  constexpr int64_t kRound = (int64_t)kScatterThreads * kScatterPer;
  int cur[kScatterPer], nxt[kScatterPer];
  load4((int64_t)tid * kScatterPer, cur);
  int buf = 0;
  for (int64_t r0 = 0; r0 < n_pairs; r0 += kRound, buf ^= 1) {
    const int64_t e0 = r0 + (int64_t)tid * kScatterPer;
    if (r0 + kRound < n_pairs) load4(e0 + kRound, nxt);
    int mine = 0;
#pragma unroll
    for (int i = 0; i < kScatterPer; ++i) mine += (cur[i] == b) ? 1 : 0;
    int incl = mine;                                                   // inclusive scan over the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
    if (lane == 31) s_warp[buf][warp] = incl;
    __syncthreads();                                                   // (double-buffered: one barrier per round)
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kScatterThreads / 32; ++w) { const int v = s_warp[buf][w]; if (w < warp) before += v; total += v; }
    if (mine) {
      int slot = base + before + incl - mine;
#pragma unroll
      for (int i = 0; i < kScatterPer; ++i) {
        if (cur[i] == b) {
          if (live) { slot_pair[slot] = (int)(e0 + i); pair_slot[e0 + i] = slot; ++slot; }
          else pair_slot[e0 + i] = -1;
        }
      }
    }
    if (live) base += total;
#pragma unroll
    for (int i = 0; i < kScatterPer; ++i) cur[i] = nxt[i];
  }
}

// pairs whose class id is outside [0,C) never match a bucket CTA: mark them explicitly.
__global__ void plan_mark_invalid_kernel(const int32_t* __restrict__ top_cls, int64_t n_pairs, int C,
                                         int32_t* __restrict__ pair_slot) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (e < n_pairs) { const int c = top_cls[e]; if (c < 0 || c >= C) pair_slot[e] = -1; }
}

// caller-supplied categories (search_single path): int64 -> int32, anything unrepresentable -> -1
__global__ void cast_i64_to_i32_kernel(const int64_t* __restrict__ in, int64_t n, int32_t* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) { const int64_t v = in[i]; out[i] = (v < 0 || v > 0x7fffffffLL) ? -1 : (int32_t)v; }
}

// Bucket-ordered copy of the fp16 query rows (A operand of the rescoring GEMM):
//   qg[slot, :] = q[slot_pair[slot] / nb, :]      one warp per slot, 16-byte vectors
__global__ void __launch_bounds__(256)
gather_queries_kernel(const uint16_t* __restrict__ q, const int32_t* __restrict__ slot_pair,
                      const PlanScalars* __restrict__ sc, int nb, int d, uint16_t* __restrict__ qg) {
  const int n_slots = sc->n_slots;
  const int lane = threadIdx.x & 31;
  const int vec = d / 8;                           // uint4 per row
  for (int slot = blockIdx.x * 8 + (threadIdx.x >> 5); slot < n_slots; slot += gridDim.x * 8) {
    const int qi = slot_pair[slot] / nb;
    const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)qi * d);
    uint4* dst = reinterpret_cast<uint4*>(qg + (size_t)slot * d);
    for (int v = lane; v < vec; v += 32) dst[v] = src[v];
  }
}

}  // namespace lmi
