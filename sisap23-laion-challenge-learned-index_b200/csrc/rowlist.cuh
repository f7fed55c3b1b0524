// rowlist.cuh -- thread-private running top-k' list used by both rescoring kernels.
//
// One thread owns one query row (= one TMEM lane in the tcgen05 kernel).  The list lives in
// shared memory as [k'][128] (entry-major, row-minor) so the 32 lanes of a warp touch 32
// consecutive banks.  Strategy: keep the list unsorted, remember its minimum; a candidate that
// beats the minimum overwrites it and the k' entries are rescanned for the new minimum.  After
// warm-up an insertion happens about k'*ln(n/k') times per row per n objects, i.e. rarely, while
// the common path is one FMUL + one compare per score.
//
// Replaces the per-row ``argsort(kind='quicksort')[:, :k]`` over the full distance matrix
// (search/li/LearnedIndex.py:170-172) and the running k-th-distance threshold of the shipped
// search (LearnedIndex.py:71-72, utils.py:22-23): the threshold here is per row and updated
// after every insertion instead of once per rank.
#pragma once
#include "common.cuh"

namespace lmi {

template <int KP>
struct RowList {
  float thr;      // insertion threshold = max(lmin, seed)
  float lmin;     // current minimum of the list (kNegInf while it still has empty entries)
  float seed;     // external lower bound on the row's final k'-th best score (kNegInf = none)
  int minpos;     // entry holding the list minimum

  __device__ __forceinline__ void reset(float* sc, int32_t* ps, int row, bool live, float seed_ = kNegInf) {
#pragma unroll 4
    for (int j = 0; j < KP; ++j) { sc[j * kTileQ + row] = kNegInf; ps[j * kTileQ + row] = -1; }
    lmin = kNegInf;
    seed = live ? seed_ : 3.0e38f;     // dead rows never insert
    thr = seed;
    minpos = 0;
  }

  // a better external bound became known
  __device__ __forceinline__ void raise_seed(float s) {
    seed = fmaxf(seed, s);
    thr = fmaxf(thr, seed);
  }

  __device__ __forceinline__ void insert(float* sc, int32_t* ps, int row, float s, int32_t pos) {
    sc[minpos * kTileQ + row] = s;
    ps[minpos * kTileQ + row] = pos;
    float m = sc[row];
    int mp = 0;
#pragma unroll 4
    for (int j = 1; j < KP; ++j) {
      const float v = sc[j * kTileQ + row];
      if (v < m) { m = v; mp = j; }
    }
    lmin = m;
    thr = fmaxf(m, seed);
    minpos = mp;
  }

  // write the row's list (unsorted) to global memory
  __device__ __forceinline__ void flush(const float* sc, const int32_t* ps, int row, Cand* out) const {
#pragma unroll 4
    for (int j = 0; j < KP; ++j) {
      Cand c;
      c.score = sc[j * kTileQ + row];
      c.pos = ps[j * kTileQ + row];
      out[j] = c;
    }
  }
};

// RowBuf: the same top-k' list with DEFERRED insertion, used by the TMEM-resident tcgen05 kernel.
//
// In a warp of 32 rows a candidate is rare per row but not per warp, and an immediate insertion
// (write + k' rescan, ~300 cycles of dependent shared-memory latency) stalls all 32 rows each time
// any one of them has a hit.  Here a hit is only appended to the row's pending area (two stores);
// the expensive merge ("prune") runs for all 32 rows of the warp at once, when some row's pending
// area is about to overflow or when the list minimum is needed (threshold exchange, flush).
// Layout: [k' + kPend][128] entries; pending entries live at k'..k'+cnt-1.  The filter threshold
// `thr` is the list minimum as of the last prune (stale = lower = conservative).
constexpr int kPend = 16;
constexpr int kGrp = 16;          // entries per group of the two-level minimum (k' >= 64)

// Replacing the minimum of a k'-entry list costs a rescan of all k' entries to find the new one: at k' = 128
// (k = 100) with ~700 replacements per row and chunk that rescan IS the kernel.  For k' >= 64 the list is therefore
// split into groups of 16 entries whose minima (and positions) live in registers: a replacement rescans the 16
// entries of one group and then the k'/16 group minima.
template <int KP>
struct RowBuf {
  static constexpr int kG = (KP >= 64) ? KP / kGrp : 1;
  float thr;      // filter: max(lmin at the last prune, seed)
  float lmin;     // minimum of the k' list (kNegInf while it has empty entries)
  float seed;     // external lower bound on the row's final k'-th best score
  int minpos;     // entry holding lmin
  int cnt;        // pending entries
  float gmin[kG]; // (k' >= 64) minimum of every group of 16 entries ...
  int gpos[kG];   // ... and the entry holding it

  __device__ __forceinline__ void reset(float* sc, int32_t* ps, int row, bool live, float seed_) {
#pragma unroll 4
    for (int j = 0; j < KP; ++j) { sc[j * kTileQ + row] = kNegInf; ps[j * kTileQ + row] = -1; }
    lmin = kNegInf;
    seed = live ? seed_ : 3.0e38f;     // dead rows never append
    thr = seed;
    minpos = 0;
    cnt = 0;
    if constexpr (kG > 1) {
#pragma unroll
      for (int g = 0; g < kG; ++g) { gmin[g] = kNegInf; gpos[g] = g * kGrp; }
    }
  }
  __device__ __forceinline__ void raise_seed(float s) {
    seed = fmaxf(seed, s);
    thr = fmaxf(thr, seed);
  }
  __device__ __forceinline__ void append(float* sc, int32_t* ps, int row, float s, int32_t pos) {
    sc[(KP + cnt) * kTileQ + row] = s;
    ps[(KP + cnt) * kTileQ + row] = pos;
    ++cnt;
  }
  // overwrite the list minimum with (s, pos) and find the new minimum
  __device__ __forceinline__ void replace_min(float* sc, int32_t* ps, int row, float s, int32_t pos) {
    sc[minpos * kTileQ + row] = s;
    ps[minpos * kTileQ + row] = pos;
    if constexpr (kG == 1) {
      float m = sc[row];
      int mp = 0;
#pragma unroll
      for (int j = 1; j < KP; ++j) {
        const float v = sc[j * kTileQ + row];
        if (v < m) { m = v; mp = j; }
      }
      lmin = m;
      minpos = mp;
    } else {
      const int g = minpos / kGrp;                       // the group that held the minimum: rescan its 16 entries
      const int base = g * kGrp;
      float m = sc[base * kTileQ + row];
      int mp = base;
#pragma unroll
      for (int j = 1; j < kGrp; ++j) {
        const float v = sc[(base + j) * kTileQ + row];
        if (v < m) { m = v; mp = base + j; }
      }
      float bm = 3.0e38f; int bp = 0;
#pragma unroll
      for (int gg = 0; gg < kG; ++gg) {                  // (select chains: the arrays stay in registers)
        gmin[gg] = (gg == g) ? m : gmin[gg];
        gpos[gg] = (gg == g) ? mp : gpos[gg];
        if (gmin[gg] < bm) { bm = gmin[gg]; bp = gpos[gg]; }
      }
      lmin = bm;
      minpos = bp;
    }
  }
  // warp-collective: merge every row's pending entries into its list
  __device__ __forceinline__ void prune(float* sc, int32_t* ps, int row) {
    const int mx = __reduce_max_sync(0xffffffffu, cnt);
    for (int i = 0; i < mx; ++i) {
      if (i < cnt) {
        const float s = sc[(KP + i) * kTileQ + row];
        if (s > lmin) replace_min(sc, ps, row, s, ps[(KP + i) * kTileQ + row]);
      }
    }
    cnt = 0;
    thr = fmaxf(lmin, seed);
  }
  // immediate insertion of one candidate that beats the list minimum (no pending entries; per-thread)
  __device__ __forceinline__ void insert_now(float* sc, int32_t* ps, int row, float s, int32_t pos) {
    replace_min(sc, ps, row, s, pos);
    thr = fmaxf(lmin, seed);
  }
  __device__ __forceinline__ void flush(const float* sc, const int32_t* ps, int row, Cand* out) const {
#pragma unroll 4
    for (int j = 0; j < KP; ++j) {
      Cand c;
      c.score = sc[j * kTileQ + row];
      c.pos = ps[j * kTileQ + row];
      out[j] = c;
    }
  }
};

}  // namespace lmi
