// rescore_ts.cuh -- K3 (default): candidate rescoring on the 5th-generation tensor cores with the
// QUERY tile resident in tensor memory.
//
// For every work item (128 query slots of one bucket x one chunk of that bucket's objects) the
// kernel computes score[r][c] = dot(q_r, x_c) * inv_norm(x_c) with fp16 inputs and fp32
// accumulation and keeps, per query row, the k' best scores in a thread-private list -- no score
// matrix ever reaches HBM.  It replaces, for all ranks and categories at once, the reference's
// per-(rank, category) float64 ``cosine_similarity`` GEMM + full-row argsort
// (search/li/LearnedIndex.py:143-172, search/li/utils.py:10-11).
//
// Why TMEM-resident queries: a 128-query tile of clip768 is 192 KB -- it does not fit in shared
// memory next to a pipeline, and re-streaming it for every object tile (rescore_tc.cuh, SS mode)
// makes the kernel L2->SM bound at ~85 FLOP/B.  TMEM has 512 columns x 128 lanes x 32 bit; the
// fp16 query tile takes d/2 = 384 columns (lane = query row, column c = halves 2c,2c+1), which
// leaves 128 columns = two 128 x 64 fp32 accumulators.  tcgen05.mma reads A from TMEM, so shared
// memory only carries the object stream: 128 FLOP per byte that crosses L2->SM.
//
// Structure (persistent, one CTA per SM, warp-specialised):
//   warp 0   : TMA producer   -- 128B-swizzled K-major boxes of 64 objects x 64 halves; one
//                                stage = 64 objects x 256 halves (4 boxes, 32 KB)
//   warp 1   : MMA issuer     -- one elected thread issues tcgen05.mma (M=128, N=64, K=16),
//                                A from TMEM, B from shared memory, D double-buffered in TMEM
//   warps 2-5: epilogue       -- per item: write the item's 128 query rows into TMEM
//                                (tcgen05.st); per tile: tcgen05.ld of the finished accumulator
//                                (thread = TMEM lane = query row), scale by inv_norm, compare
//                                against the row's running threshold, rare insertion
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <->
// epilogue), a_full (epilogue -> MMA: the item's queries are in TMEM).
//
// Running threshold: besides its own list minimum a row uses a per-QUERY lower bound shared by
// all CTAs (qthr, order-preserving uint32 image of an fp32 score, atomicMax): once any list of
// that query holds k' entries >= t, nothing below t can reach the query's final top-k', whatever
// bucket or chunk it comes from.  Rows publish/refresh it every kTsRefresh tiles.
#pragma once
#include "common.cuh"
#include "plan.cuh"
#include "ptx_sm100.cuh"
#include "rowlist.cuh"

namespace lmi {

constexpr int kTsBlockN = 64;                       // objects per MMA tile (UMMA N)
constexpr int kTsAtomK = 64;                        // halves per swizzle atom row (128 bytes)
constexpr int kTsAtomsPerStage = 4;                 // 256 halves of K per stage
constexpr int kTsAtomBytes = kTsBlockN * kTsAtomK * 2;            // 8 KB
constexpr int kTsStageBytes = kTsAtomsPerStage * kTsAtomBytes;    // 32 KB
constexpr int kTsThreads = 192;
constexpr int kTsACols = 384;                       // TMEM columns reserved for the query tile (d <= 768)
constexpr int kTsItemRing = 2;                      // depth of the producer -> consumers work item ring
constexpr int kTsRefresh = 8;                       // tiles between refreshes of the shared threshold

template <int KP, int STAGES>
constexpr size_t ts_smem_bytes() {
  return 1024 /*alignment slack*/ + (size_t)STAGES * kTsStageBytes + (sizeof(float) + sizeof(int32_t)) * (KP + kPend) * kTileQ +
         4 * 2 * kTsBlockN * sizeof(float) + 256 /*barriers + tmem ptr*/;
}

// order-preserving float <-> uint32 (0 = "nothing published yet")
__device__ __forceinline__ uint32_t thr_encode(float f) {
  const uint32_t b = __float_as_uint(f);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float thr_decode(uint32_t u) {
  if (u == 0u) return kNegInf;
  return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

template <int KP, int STAGES>
__global__ void __launch_bounds__(kTsThreads, 1)
rescore_ts_kernel(const __grid_constant__ CUtensorMap tmap_x,      // bucket store [n_local, d], box 64 x 64
                  const uint16_t* __restrict__ q,                  // [nq, d] fp16 bits (original order)
                  const int32_t* __restrict__ slot_pair,           // slot -> pair index q*nb + r
                  int nb, int d, const float* __restrict__ inv_norm,
                  const WorkItem* __restrict__ items, PlanScalars* __restrict__ sc,
                  uint32_t* __restrict__ qthr,                     // [nq] shared per-query threshold
                  Cand* __restrict__ plists, int dbg_flags) {
  extern __shared__ unsigned char ts_smem_raw[];
  // 1024-byte alignment: required by the 128B swizzle atom (8 rows x 128 B)
  // (offset arithmetic on the shared array keeps the shared address space visible to the compiler)
  unsigned char* smem = ts_smem_raw + ((1024u - (ptx::smem_u32(ts_smem_raw) & 1023u)) & 1023u);
  unsigned char* stage_base = smem;
  float* l_sc = reinterpret_cast<float*>(smem + (size_t)STAGES * kTsStageBytes);    // [KP + kPend][128]
  int32_t* l_ps = reinterpret_cast<int32_t*>(l_sc + (KP + kPend) * kTileQ);         // [KP + kPend][128]
  float* s_inv = reinterpret_cast<float*>(l_ps + (KP + kPend) * kTileQ);            // [4 warps][2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_inv + 4 * 2 * kTsBlockN);
  uint64_t* full_bar = bars;                       // [STAGES]
  uint64_t* empty_bar = bars + STAGES;             // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;         // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;    // [2]
  uint64_t* afull_bar = bars + 2 * STAGES + 4;     // [1]
  uint64_t* ifull_bar = bars + 2 * STAGES + 5;     // [kTsItemRing] work item published by the producer
  uint64_t* iempty_bar = bars + 2 * STAGES + 5 + kTsItemRing;   // [kTsItemRing] ... read by MMA + 4 epilogue warps
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 5 + 2 * kTsItemRing);
  int32_t* s_item = reinterpret_cast<int32_t*>(tmem_ptr + 1);   // [kTsItemRing]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_items = sc->n_items;
  const int k_atoms = d / kTsAtomK;                                  // 12 for clip768
  const int k_stages = (k_atoms + kTsAtomsPerStage - 1) / kTsAtomsPerStage;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_x);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty_bar[s]), 4); }
    ptx::mbar_init(ptx::smem_u32(afull_bar), 4);
    for (int s = 0; s < kTsItemRing; ++s) { ptx::mbar_init(ptx::smem_u32(&ifull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&iempty_bar[s]), 5); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_ptr));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_d = tmem_base + kTsACols;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int islot = 0; uint32_t iphase = 0;
      for (;;) {
        // dynamic scheduling: CTAs that become free one after the other pick up consecutive items =
        // the query tiles of one object chunk, so those stream the chunk nearly in lock-step and
        // all but the first read it from L2
        ptx::mbar_wait(ptx::smem_u32(&iempty_bar[islot]), iphase ^ 1);
        const int it = atomicAdd(&sc->work_counter, 1);
        s_item[islot] = it;
        ptx::mbar_arrive(ptx::smem_u32(&ifull_bar[islot]));
        if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
        if (it >= n_items) break;
        const WorkItem wi = items[it];
        const int n_tiles = (wi.obj_count + kTsBlockN - 1) / kTsBlockN;
        for (int t = 0; t < n_tiles; ++t) {
          const int obj0 = wi.obj_begin + t * kTsBlockN;
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);
            const uint32_t fb = ptx::smem_u32(&full_bar[stage]);
            ptx::mbar_arrive_expect_tx(fb, (uint32_t)(atoms * kTsAtomBytes));
            const uint32_t sb = ptx::smem_u32(stage_base + (size_t)stage * kTsStageBytes);
            for (int a = 0; a < atoms; ++a)
              ptx::tma_load_2d(sb + a * kTsAtomBytes, &tmap_x, fb, (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp walks the loops (warp-uniform control flow); one elected lane issues.  With
    // N = 64 a tcgen05.mma is 32 tensor-pipe cycles, so the issue sequence must stay short:
    // full stages are unrolled with compile-time operand offsets and accumulate flags.
    constexpr uint32_t idesc = ptx::umma_idesc_f16(kTileQ, kTsBlockN);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    uint32_t a_phase = 0;
    int islot = 0; uint32_t iphase = 0;
    for (;;) {
      ptx::mbar_wait(ptx::smem_u32(&ifull_bar[islot]), iphase);
      const int it = s_item[islot];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&iempty_bar[islot]));
      if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
      if (it >= n_items) break;
      const int n_tiles = (items[it].obj_count + kTsBlockN - 1) / kTsBlockN;
      ptx::mbar_wait(ptx::smem_u32(afull_bar), a_phase);             // this item's queries are in TMEM
      a_phase ^= 1;
      ptx::tc_fence_after();
      for (int t = 0; t < n_tiles; ++t) {
        ptx::mbar_wait(ptx::smem_u32(&tempty_bar[acc]), acc_phase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_d + (uint32_t)(acc * kTsBlockN);
        for (int ks = 0; ks < k_stages; ++ks) {
          const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
          ptx::mbar_wait(ptx::smem_u32(&full_bar[stage]), phase);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint64_t b0 = ptx::umma_desc_k_sw128(ptx::smem_u32(stage_base + (size_t)stage * kTsStageBytes));
            const uint32_t a0 = tmem_base + (uint32_t)(ks * kTsAtomsPerStage * 32);   // 32 columns per atom
            if (atoms == kTsAtomsPerStage) {
              // A: 16 halves = 8 TMEM columns per K=16 slice; B: +32 bytes inside the swizzle span
              // (+2 in the addr>>4 field), next atom +8 KB (+512)
              if (ks == 0) ptx::umma_f16_ts_c<false>(d_tmem, a0, b0, idesc);
              else         ptx::umma_f16_ts_c<true>(d_tmem, a0, b0, idesc);
#pragma unroll
              for (int i = 1; i < 4 * kTsAtomsPerStage; ++i)
                ptx::umma_f16_ts_c<true>(d_tmem, a0 + (uint32_t)(i * 8),
                                         b0 + (uint64_t)((i >> 2) * (kTsAtomBytes >> 4) + (i & 3) * 2), idesc);
            } else {
              for (int i = 0; i < 4 * atoms; ++i)
                ptx::umma_f16_ts(d_tmem, a0 + (uint32_t)(i * 8),
                                 b0 + (uint64_t)((i >> 2) * (kTsAtomBytes >> 4) + (i & 3) * 2), idesc,
                                 (ks | i) != 0 ? 1u : 0u);
            }
            ptx::umma_commit(ptx::smem_u32(&empty_bar[stage]));      // smem stage reusable once MMAs retire
            if (ks == k_stages - 1) ptx::umma_commit(ptx::smem_u32(&tfull_bar[acc]));   // accumulator complete
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue (4 warps, thread = query row) =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    RowBuf<KP> list;
    float* inv_w = s_inv + quarter * 2 * kTsBlockN;    // this warp's private copy of the tile's inverse norms
    int acc = 0; uint32_t acc_phase = 0;
    int islot = 0; uint32_t iphase = 0;
    for (;;) {
      ptx::mbar_wait(ptx::smem_u32(&ifull_bar[islot]), iphase);
      const int it = s_item[islot];
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&iempty_bar[islot]));
      if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
      if (it >= n_items) break;
      const WorkItem wi = items[it];
      const bool live = row < wi.rows && !(dbg_flags & 1);
      const int qi = live ? slot_pair[wi.slot_base + row] / nb : -1;
      // ---- the item's query rows -> TMEM (all MMAs of the previous item have completed: this
      //      thread has seen the tfull barrier of its last tile) ----
      {
        const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(live ? qi : 0) * d);
        const int n_blk = d / 64;                 // 64 halves = 32 columns per block
        uint4 cur[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) cur[v] = live ? src[v] : make_uint4(0, 0, 0, 0);
        for (int blk = 0; blk < n_blk; ++blk) {
          uint4 nxt[8];
          if (blk + 1 < n_blk) {
#pragma unroll
            for (int v = 0; v < 8; ++v) nxt[v] = live ? src[(blk + 1) * 8 + v] : make_uint4(0, 0, 0, 0);
          }
          uint32_t w[32];
#pragma unroll
          for (int v = 0; v < 8; ++v) { w[4 * v] = cur[v].x; w[4 * v + 1] = cur[v].y; w[4 * v + 2] = cur[v].z; w[4 * v + 3] = cur[v].w; }
          ptx::tmem_st_32x32b_x32(tmem_base + lane_addr + (uint32_t)(blk * 32), w);
#pragma unroll
          for (int v = 0; v < 8; ++v) cur[v] = nxt[v];
        }
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(afull_bar));
      }
      list.reset(l_sc, l_ps, row, live, live ? thr_decode(__ldcg(qthr + qi)) : kNegInf);
      const int n_tiles = (wi.obj_count + kTsBlockN - 1) / kTsBlockN;
      // inverse norms are fetched one tile ahead (each lane owns two columns of its warp's copy)
      float inv_n0 = (lane < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane) : 0.f;
      float inv_n1 = (lane + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane + 32) : 0.f;
      uint32_t pending = 0u;                       // result of the last threshold exchange, consumed one tile later
      for (int t = 0; t < n_tiles; ++t) {
        const int col0 = t * kTsBlockN;
        const int ncols = min(kTsBlockN, wi.obj_count - col0);
        float* inv = inv_w + acc * kTsBlockN;
        inv[lane] = inv_n0;
        inv[lane + 32] = inv_n1;
        {
          const int c_next = col0 + kTsBlockN + lane;
          inv_n0 = (c_next < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_next) : 0.f;
          inv_n1 = (c_next + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_next + 32) : 0.f;
        }
        list.raise_seed(thr_decode(pending));
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&tfull_bar[acc]), acc_phase);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_d + lane_addr + (uint32_t)(acc * kTsBlockN);
        const int pos0 = wi.obj_begin + col0;
        uint32_t v0[32], v1[32];
        ptx::tmem_ld_32x32b_x32(taddr, v0);
        ptx::tmem_ld_32x32b_x32(taddr + 32, v1);
        ptx::tmem_ld_wait();
        // accumulator stage is in registers -> hand it back to the MMA warp right away
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty_bar[acc]));
        // groups of 8 columns: one max + one warp vote decide whether anybody has a candidate at all
        // (a real, warp-uniform branch: the insertion path must not be if-converted into the hot loop)
        const bool full_tile = (ncols == kTsBlockN);
#pragma unroll
        for (int g = 0; g < kTsBlockN / 8; ++g) {
          float s[8];
          const float4 i0 = *reinterpret_cast<const float4*>(inv + g * 8);
          const float4 i1 = *reinterpret_cast<const float4*>(inv + g * 8 + 4);
          const float iv[8] = {i0.x, i0.y, i0.z, i0.w, i1.x, i1.y, i1.z, i1.w};
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int c = g * 8 + j;
            s[j] = __uint_as_float(c < 32 ? v0[c & 31] : v1[c & 31]) * iv[j];
          }
          if (!full_tile) {
#pragma unroll
            for (int j = 0; j < 8; ++j) if (g * 8 + j >= ncols) s[j] = kNegInf;
          }
          const float m = fmaxf(fmaxf(fmaxf(s[0], s[1]), fmaxf(s[2], s[3])), fmaxf(fmaxf(s[4], s[5]), fmaxf(s[6], s[7])));
          if (__any_sync(0xffffffffu, m > list.thr)) {
#pragma unroll
            for (int j = 0; j < 8; ++j)
              if (s[j] > list.thr) list.append(l_sc, l_ps, row, s[j], pos0 + g * 8 + j);
            // keep room for the next group's (up to 8) appends
            if (__any_sync(0xffffffffu, list.cnt > kPend - 8)) list.prune(l_sc, l_ps, row);
          }
        }
        // publish this row's bound / pick up the other CTAs' (the atomic's old value is read next tile)
        if ((t + 1) % kTsRefresh == 0 || t + 1 == n_tiles) {
          list.prune(l_sc, l_ps, row);
          if (live) pending = atomicMax(qthr + qi, list.lmin > kNegInf ? thr_encode(list.lmin) : 0u);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (live)
        list.flush(l_sc, l_ps, row, plists + (size_t)(wi.plist_base + (int64_t)row * wi.plist_stride) * KP);
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem_base); }
}

}  // namespace lmi
