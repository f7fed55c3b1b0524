// rescore_tc.cuh -- K3: candidate rescoring on the 5th-generation tensor cores.
//
// For every work item (128 query slots of one bucket x one chunk of that bucket's objects) the
// kernel computes score[r][c] = dot(q_r, x_c) * inv_norm(x_c) with fp16 inputs and fp32
// accumulation and keeps, per query row, the k' best scores in a thread-private list -- no score
// matrix ever reaches HBM.  It replaces, for all ranks and categories at once, the reference's
// per-(rank, category) float64 ``cosine_similarity`` GEMM + full-row argsort
// (search/li/LearnedIndex.py:143-172, search/li/utils.py:10-11).
//
// Structure (persistent, one CTA per SM, warp-specialised):
//   warp 0   : TMA producer   -- 128B-swizzled K-major tiles of queries (A, 128 x 64) and
//                                objects (B, 256 x 64) into a ring of smem stages
//   warp 1   : MMA issuer     -- one elected thread issues tcgen05.mma (M=128, N=256, K=16),
//                                accumulators double-buffered in TMEM (2 x 256 columns)
//   warps 2-5: epilogue       -- tcgen05.ld of the finished accumulator (thread = TMEM lane =
//                                query row), scale by inv_norm, compare against the row's
//                                running threshold, rare insertion into the row list
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <-> epilogue).
#pragma once
#include "common.cuh"
#include "plan.cuh"
#include "ptx_sm100.cuh"
#include "rowlist.cuh"

namespace lmi {

constexpr int kTcBlockN = 256;                    // objects per MMA tile (UMMA N)
constexpr int kTcBlockK = 64;                     // halves per smem row = 128 bytes (one swizzle span)
constexpr int kTcThreads = 192;
constexpr int kTcABytes = kTileQ * kTcBlockK * 2;       // 16 KB
constexpr int kTcBBytes = kTcBlockN * kTcBlockK * 2;    // 32 KB
constexpr int kTcStageBytes = kTcABytes + kTcBBytes;

template <int KP, int STAGES>
constexpr size_t tc_smem_bytes() {
  return 1024 /*alignment slack*/ + (size_t)STAGES * kTcStageBytes + (sizeof(float) + sizeof(int32_t)) * KP * kTileQ +
         2 * kTcBlockN * sizeof(float) + 256 /*barriers + tmem ptr*/;
}

template <int KP, int STAGES>
__global__ void __launch_bounds__(kTcThreads, 1)
rescore_tc_kernel(const __grid_constant__ CUtensorMap tmap_q,      // gathered queries [slots, d]
                  const __grid_constant__ CUtensorMap tmap_x,      // bucket store     [n_local, d]
                  int d, const float* __restrict__ inv_norm,
                  const WorkItem* __restrict__ items, const PlanScalars* __restrict__ sc,
                  Cand* __restrict__ plists) {
  extern __shared__ unsigned char tc_smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atom (8 rows x 128 B)
  unsigned char* smem = reinterpret_cast<unsigned char*>(
      (reinterpret_cast<uintptr_t>(tc_smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  unsigned char* stage_base = smem;
  float* l_sc = reinterpret_cast<float*>(smem + (size_t)STAGES * kTcStageBytes);    // [KP][128]
  int32_t* l_ps = reinterpret_cast<int32_t*>(l_sc + KP * kTileQ);                   // [KP][128]
  float* s_inv = reinterpret_cast<float*>(l_ps + KP * kTileQ);                      // [2][256]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_inv + 2 * kTcBlockN);
  uint64_t* full_bar = bars;                       // [STAGES]
  uint64_t* empty_bar = bars + STAGES;             // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;         // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;    // [2]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = sc->n_items;
  const int k_chunks = d / kTcBlockK;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_q);
    ptx::prefetch_tensormap(&tmap_x);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty_bar[s]), 4); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_ptr));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const WorkItem wi = items[it];
        const int n_tiles = (wi.obj_count + kTcBlockN - 1) / kTcBlockN;
        for (int t = 0; t < n_tiles; ++t) {
          const int obj0 = wi.obj_begin + t * kTcBlockN;
          for (int kc = 0; kc < k_chunks; ++kc) {
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);
            const uint32_t fb = ptx::smem_u32(&full_bar[stage]);
            ptx::mbar_arrive_expect_tx(fb, kTcStageBytes);
            const uint32_t sa = ptx::smem_u32(stage_base + (size_t)stage * kTcStageBytes);
            ptx::tma_load_2d(sa, &tmap_q, fb, kc * kTcBlockK, wi.slot_base);
            ptx::tma_load_2d(sa + kTcABytes, &tmap_x, fb, kc * kTcBlockK, obj0);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = ptx::umma_idesc_f16(kTileQ, kTcBlockN);
      int stage = 0; uint32_t phase = 0;
      int acc = 0; uint32_t acc_phase = 0;
      for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
        const WorkItem wi = items[it];
        const int n_tiles = (wi.obj_count + kTcBlockN - 1) / kTcBlockN;
        for (int t = 0; t < n_tiles; ++t) {
          ptx::mbar_wait(ptx::smem_u32(&tempty_bar[acc]), acc_phase ^ 1);
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)(acc * kTcBlockN);
          for (int kc = 0; kc < k_chunks; ++kc) {
            ptx::mbar_wait(ptx::smem_u32(&full_bar[stage]), phase);
            ptx::tc_fence_after();
            const uint32_t sa = ptx::smem_u32(stage_base + (size_t)stage * kTcStageBytes);
            const uint64_t a_desc = ptx::umma_desc_k_sw128(sa);
            const uint64_t b_desc = ptx::umma_desc_k_sw128(sa + kTcABytes);
#pragma unroll
            for (int k4 = 0; k4 < kTcBlockK / 16; ++k4) {
              // advance 16 halves = 32 bytes inside the swizzle span: +2 in the (addr >> 4) field
              ptx::umma_f16_ss(d_tmem, a_desc + (uint64_t)(k4 * 2), b_desc + (uint64_t)(k4 * 2), idesc,
                               (kc | k4) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(ptx::smem_u32(&empty_bar[stage]));      // smem stage reusable once MMAs retire
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          ptx::umma_commit(ptx::smem_u32(&tfull_bar[acc]));          // accumulator complete
          if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue (4 warps, thread = query row) =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const int et = threadIdx.x - 64;              // 0..127 among epilogue threads
    RowList<KP> list;
    int acc = 0; uint32_t acc_phase = 0;
    for (int it = blockIdx.x; it < n_items; it += gridDim.x) {
      const WorkItem wi = items[it];
      list.reset(l_sc, l_ps, row, row < wi.rows);
      const int n_tiles = (wi.obj_count + kTcBlockN - 1) / kTcBlockN;
      for (int t = 0; t < n_tiles; ++t) {
        const int col0 = t * kTcBlockN;
        const int ncols = min(kTcBlockN, wi.obj_count - col0);
        float* inv = s_inv + acc * kTcBlockN;
        // stage inverse norms of this tile's objects (overlaps the MMA still in flight)
        for (int c = et; c < kTcBlockN; c += 128) inv[c] = (c < ncols) ? __ldg(inv_norm + wi.obj_begin + col0 + c) : 0.f;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        ptx::mbar_wait(ptx::smem_u32(&tfull_bar[acc]), acc_phase);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(acc * kTcBlockN);
        const int pos0 = wi.obj_begin + col0;
        for (int cb = 0; cb < kTcBlockN; cb += 32) {
          if (cb >= ncols) break;                 // warp-uniform
          uint32_t v[32];
          ptx::tmem_ld_32x32b_x32(taddr + cb, v);
          ptx::tmem_ld_wait();
          if (cb + 32 <= ncols) {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float s = __uint_as_float(v[j]) * inv[cb + j];
              if (s > list.thr) list.insert(l_sc, l_ps, row, s, pos0 + cb + j);
            }
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              if (cb + j < ncols) {
                const float s = __uint_as_float(v[j]) * inv[cb + j];
                if (s > list.thr) list.insert(l_sc, l_ps, row, s, pos0 + cb + j);
              }
            }
          }
        }
        // accumulator stage drained -> hand it back to the MMA warp
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty_bar[acc]));
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (row < wi.rows)
        list.flush(l_sc, l_ps, row, plists + (size_t)(wi.plist_base + (int64_t)row * wi.plist_stride) * KP);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem_base); }
}

}  // namespace lmi
