// rescore_ts2.cuh -- K3 (CTA-pair variant): the TMEM-resident rescoring kernel of rescore_ts.cuh with
// tcgen05 `cta_group::2` MMAs.
//
// Same computation as rescore_ts_kernel (score[r][c] = dot(q_r, x_c) * inv_norm(x_c), fp16 inputs,
// fp32 accumulation, per-row top-k' lists fused into the epilogue; replaces the reference's
// per-(rank, category) float64 cosine GEMM + argsort, search/li/LearnedIndex.py:143-172,
// search/li/utils.py:10-11).  What changes is who holds the object tile:
//
//   rescore_ts  : every CTA receives the whole 64-object tile (96 KB per 128 x 64 x 768 tile).
//   rescore_ts2 : a work item is a PAIR of 128-query tiles of the same object chunk, one per CTA of a
//                 2-CTA cluster.  One `tcgen05.mma.cta_group::2` (M = 256, N = 64, K = 16), issued by
//                 the leader CTA, multiplies both query tiles (A: 128 rows in each CTA's tensor
//                 memory) with a 64-object tile of which each CTA holds only ITS 32 objects in shared
//                 memory.  Per SM and tile the MMA time is unchanged (32 cycles per instruction) but
//                 the bytes that must arrive are halved (48 KB per tile), and one thread issues for
//                 both SMs: MMA + handshake alone run at ~1,620 cycles per tile (floor 1,536) against
//                 ~2,150 for the two lock-stepped CTAs of rescore_ts.
//
// Every tcgen05.alloc/mma/commit of a kernel must use one cta_group width, so this kernel cannot also
// run single-CTA items: the planner's pair mode (plan.cuh, `pair_mode`) emits shared pairs only.  An odd
// last query tile becomes two HALF tiles (<= 64 rows per CTA, kItemHalf) scored with M = 128 pair MMAs
// (64 rows per CTA, half the tensor time): lanes l and l + 64 of a CTA's tensor memory hold query row
// l & 63, lanes 0-63 receive the scores of the tile's objects 0-31, lanes 64-127 those of objects 32-63.
//
// Cross-CTA protocol (leader = cluster rank 0):
//   full_bar[s]   (leader's)  : leader's producer expects the bytes of BOTH halves; the peer's TMA
//                               signals the leader's barrier (cp.async.bulk.tensor ... cta_group::2)
//   empty_bar[s]  (both)      : tcgen05.commit.cta_group::2 multicast from the leader's issuing warp
//   tfull_bar[a]  (both)      : same (accumulator a complete in both CTAs' tensor memory)
//   tempty_bar[a] (leader's)  : 4 epilogue warps x 2 CTAs arrive (the peer's remotely)
//   afull_bar     (leader's)  : same, "this item's query rows are in tensor memory"
//   item ring                 : as in rescore_ts (leader's producer publishes into both CTAs)
#pragma once
#include "rescore_ts.cuh"

namespace lmi {

constexpr int kT2HalfN = kTsBlockN / 2;                            // objects per CTA and tile
constexpr int kT2AtomBytes = kT2HalfN * kTsAtomK * 2;              // 4 KB: 32 objects x 64 halves, 128B swizzle
constexpr int kT2StageBytes = kTsAtomsPerStage * kT2AtomBytes;     // 16 KB per CTA

template <int KP, int STAGES>
constexpr size_t ts2_smem_bytes() {
  return 1024 /*alignment slack*/ + (size_t)STAGES * kT2StageBytes + (sizeof(float) + sizeof(int32_t)) * (KP + kPend) * kTileQ +
         4 * 2 * kTsBlockN * sizeof(float) + 512 /*barriers + tmem ptr + item ring*/;
}

template <int KP, int STAGES>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTsThreads, 1)
rescore_ts2_kernel(const __grid_constant__ CUtensorMap tmap_xh,     // bucket store [n_local, d], box 32 x 64
                   const __grid_constant__ CUtensorMap tmap_x3h,    // 3-D view, box 64 halves x 32 objects x 4 K-atoms = one stage
                   const uint16_t* __restrict__ q,                  // [nq, d] fp16 bits (original order)
                   const int32_t* __restrict__ slot_pair,           // slot -> pair index q*nb + r
                   int nb, int d, const float* __restrict__ inv_norm,
                   const WorkItem* __restrict__ items, PlanScalars* __restrict__ sc,
                   uint32_t* __restrict__ qthr,                     // [nq] shared per-query threshold
                   Cand* __restrict__ plists, int dbg_flags) {
  static_assert(2 * STAGES * 8 + 9 * 8 + 4 * kTsItemRing * 4 + 16 <= 512, "barrier area too small");
  extern __shared__ unsigned char ts2_smem_raw[];
  // 1024-byte alignment (128B swizzle atom); the offset is the same in both CTAs, so one shared-memory
  // descriptor addresses the leader's and the peer's half of a tile
  unsigned char* smem = ts2_smem_raw + ((1024u - (ptx::smem_u32(ts2_smem_raw) & 1023u)) & 1023u);
  unsigned char* stage_base = smem;
  float* l_sc = reinterpret_cast<float*>(smem + (size_t)STAGES * kT2StageBytes);    // [KP + kPend][128]
  int32_t* l_ps = reinterpret_cast<int32_t*>(l_sc + (KP + kPend) * kTileQ);         // [KP + kPend][128]
  float* s_inv = reinterpret_cast<float*>(l_ps + (KP + kPend) * kTileQ);            // [4 warps][2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_inv + 4 * 2 * kTsBlockN);
  uint64_t* full_bar = bars;                       // [STAGES]  (the leader's are used)
  uint64_t* empty_bar = bars + STAGES;             // [STAGES]
  uint64_t* tfull_bar = bars + 2 * STAGES;         // [2]
  uint64_t* tempty_bar = bars + 2 * STAGES + 2;    // [2]       (the leader's are used)
  uint64_t* afull_bar = bars + 2 * STAGES + 4;     // [1]       (the leader's is used)
  uint64_t* ifull_bar = bars + 2 * STAGES + 5;     // [kTsItemRing]
  uint64_t* iempty_bar = bars + 2 * STAGES + 5 + kTsItemRing;   // [kTsItemRing]
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 5 + 2 * kTsItemRing);
  int32_t* s_item = reinterpret_cast<int32_t*>(tmem_ptr + 1);   // [kTsItemRing]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t crank = ptx::cluster_ctarank();                     // 0 = leader: fetches work, issues the MMAs
  const int n_pairs = sc->n_items >> 1;
  const int k_atoms = d / kTsAtomK;                                  // 12 for clip768
  const int k_stages = (k_atoms + kTsAtomsPerStage - 1) / kTsAtomsPerStage;
  const bool relay = (dbg_flags & 16) != 0;
  const bool cl_scope = (dbg_flags & 32) != 0;      // experiment: cluster-scope acquire/release on the per-tile barriers
  auto wait_x = [&](uint32_t bar, uint32_t parity) { if (cl_scope) ptx::mbar_wait_cluster(bar, parity); else ptx::mbar_wait(bar, parity); };
  auto arrive_x = [&](uint32_t cluster_addr) { if (cl_scope) ptx::mbar_arrive_cluster(cluster_addr); else ptx::mbar_arrive_remote(cluster_addr); };

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_xh);
    // relay mode (dbg_flags & 16, experiment): every CTA loads with the plain TMA form onto its OWN full barrier and
    // the peer's otherwise idle warp 1 forwards each completed stage to the leader's barrier (second arrival)
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full_bar[s]), (relay && crank == 0) ? 2 : 1); ptx::mbar_init(ptx::smem_u32(&empty_bar[s]), 1); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty_bar[s]), 8); }
    ptx::mbar_init(ptx::smem_u32(afull_bar), 8);
    // item ring: the leader's iempty collects its 2 MMA + 4 epilogue warps and the peer's producer + 4
    // epilogue warps (the peer's MMA warps idle: only the leader issues)
    for (int s = 0; s < kTsItemRing; ++s) { ptx::mbar_init(ptx::smem_u32(&ifull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&iempty_bar[s]), relay ? 12 : 11); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc2<512>(ptx::smem_u32(tmem_ptr));     // warp 1 of BOTH CTAs
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();                               // the peer's barriers and tensor memory exist too
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_d = tmem_base + kTsACols;
  long long clk0 = 0; unsigned long long ns0 = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) { clk0 = clock64(); ns0 = ptx::globaltimer_ns(); }

  const uint32_t iempty_leader = ptx::mapa(ptx::smem_u32(iempty_bar), 0);
  auto next_pair = [&](int& islot, uint32_t& iphase) -> int {
    ptx::mbar_wait_cluster(ptx::smem_u32(&ifull_bar[islot]), iphase);
    const int pr = s_item[islot];
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive_cluster(iempty_leader + (uint32_t)islot * 8u);
    if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
    return pr;
  };

  if (warp == 0) {
    // ===================== TMA producer (both CTAs; each loads its 32 objects of every tile) =====================
    if (lane == 0) {
      const uint32_t full_leader = ptx::mapa(ptx::smem_u32(full_bar), 0);
      int stage = 0; uint32_t phase = 0;
      int islot = 0; uint32_t iphase = 0;
      for (;;) {
        int pr;
        if (crank == 0) {
          ptx::mbar_wait_cluster(ptx::smem_u32(&iempty_bar[islot]), iphase ^ 1);
          pr = atomicAdd(&sc->work_counter, 1);
          s_item[islot] = pr;
          ptx::st_cluster_u32(ptx::mapa(ptx::smem_u32(&s_item[islot]), 1), (uint32_t)pr);
          ptx::mbar_arrive(ptx::smem_u32(&ifull_bar[islot]));
          ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&ifull_bar[islot]), 1));
          if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
        } else {
          ptx::mbar_wait_cluster(ptx::smem_u32(&ifull_bar[islot]), iphase);
          pr = s_item[islot];
          ptx::mbar_arrive_cluster(iempty_leader + (uint32_t)islot * 8u);
          if (++islot == kTsItemRing) { islot = 0; iphase ^= 1; }
        }
        if (pr >= n_pairs) break;
        const WorkItem wi = items[2 * pr];            // both entries of a pair score the same object chunk
        const int n_tiles = (wi.obj_count + kTsBlockN - 1) / kTsBlockN;
        for (int t = 0; t < n_tiles; ++t) {
          const int obj0 = wi.obj_begin + t * kTsBlockN + (int)crank * kT2HalfN;
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);    // leader's commit, multicast to both CTAs
            if (dbg_flags & 4) {                                     // experiment: no loads at all
              if (crank == 0) ptx::mbar_arrive(ptx::smem_u32(&full_bar[stage]));
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
              continue;
            }
            const uint32_t sb = ptx::smem_u32(stage_base + (size_t)stage * kT2StageBytes);
            const bool stage_box = (dbg_flags & 1024) != 0;     // one 3-D TMA per stage (needs d % 256 == 0)
            if (relay) {
              const uint32_t fbl = ptx::smem_u32(&full_bar[stage]);
              ptx::mbar_arrive_expect_tx(fbl, (uint32_t)(atoms * kT2AtomBytes));
              if (stage_box) ptx::tma_load_3d(sb, &tmap_x3h, fbl, 0, obj0, ks * kTsAtomsPerStage);
              else
                for (int a = 0; a < atoms; ++a)
                  ptx::tma_load_2d(sb + a * kT2AtomBytes, &tmap_xh, fbl, (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0);
            } else {
              if (crank == 0) ptx::mbar_arrive_expect_tx(ptx::smem_u32(&full_bar[stage]), (uint32_t)(2 * atoms * kT2AtomBytes));
              const uint32_t fb = full_leader + (uint32_t)stage * 8u;
              if (stage_box) ptx::tma_load_3d_cg2(sb, &tmap_x3h, fb, 0, obj0, ks * kTsAtomsPerStage);
              else
                for (int a = 0; a < atoms; ++a)
                  ptx::tma_load_2d_cg2(sb + a * kT2AtomBytes, &tmap_xh, fb, (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
      // Producer tail: no CTA exits while commit arrivals for its stages may still be in flight
      for (int s = 0; s < STAGES; ++s) {
        ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 || warp == 6) {
    // ===================== MMA issuers (leader CTA only; two warps alternate tiles) =====================
    if (crank == 0) {
      constexpr uint32_t idesc_full = ptx::umma_idesc_f16(2 * kTileQ, kTsBlockN);  // M = 256 over the pair, N = 64
      constexpr uint32_t idesc_half = ptx::umma_idesc_f16(kTileQ, kTsBlockN);      // M = 128 over the pair (64 rows per CTA)
      const int parity = (warp == 6) ? 1 : 0;
      const uint32_t d_tmem = tmem_d + (uint32_t)(parity * kTsBlockN);
      uint32_t g = 0;                                // running tile number; g & 1 = accumulator = issuing warp
      int stage = 0; uint32_t phase = 0;
      uint32_t a_phase = 0;
      int islot = 0; uint32_t iphase = 0;
      for (;;) {
        const int pr = next_pair(islot, iphase);
        if (pr >= n_pairs) break;
        const int n_tiles = (items[2 * pr].obj_count + kTsBlockN - 1) / kTsBlockN;
        const uint32_t idesc = (items[2 * pr].flags & kItemHalf) ? idesc_half : idesc_full;
        wait_x(ptx::smem_u32(afull_bar), a_phase);     // both CTAs' query rows are in tensor memory
        a_phase ^= 1;
        for (int t = 0; t < n_tiles; ++t, ++g) {
          if ((int)(g & 1u) != parity) {
            // the other warp's tile: observe each of its stages becoming full (see rescore_ts.cuh)
            for (int ks = 0; ks < k_stages; ++ks) {
              ptx::mbar_wait(ptx::smem_u32(&full_bar[stage]), phase);
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
            }
            continue;
          }
          wait_x(ptx::smem_u32(&tempty_bar[parity]), ((g >> 1) & 1u) ^ 1u);   // drained in BOTH CTAs
          ptx::tc_fence_after();
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            wait_x(ptx::smem_u32(&full_bar[stage]), phase);         // both halves have landed
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              const uint64_t b0 = ptx::umma_desc_k_sw128(ptx::smem_u32(stage_base + (size_t)stage * kT2StageBytes));
              const uint32_t a0 = tmem_base + (uint32_t)(ks * kTsAtomsPerStage * 32);   // 32 columns per atom
              // A: 8 TMEM columns per K=16 slice; B: +32 bytes inside the swizzle span, next atom +4 KB
              if (dbg_flags & 8) {
              } else if (atoms == kTsAtomsPerStage) {
                if (ks == 0) ptx::umma2_f16_ts_c<false>(d_tmem, a0, b0, idesc);
                else         ptx::umma2_f16_ts_c<true>(d_tmem, a0, b0, idesc);
#pragma unroll
                for (int i = 1; i < 4 * kTsAtomsPerStage; ++i)
                  ptx::umma2_f16_ts_c<true>(d_tmem, a0 + (uint32_t)(i * 8),
                                            b0 + (uint64_t)((i >> 2) * (kT2AtomBytes >> 4) + (i & 3) * 2), idesc);
              } else {
                for (int i = 0; i < 4 * atoms; ++i)
                  ptx::umma2_f16_ts(d_tmem, a0 + (uint32_t)(i * 8),
                                    b0 + (uint64_t)((i >> 2) * (kT2AtomBytes >> 4) + (i & 3) * 2), idesc,
                                    (ks | i) != 0 ? 1u : 0u);
              }
              ptx::umma2_commit_mc(ptx::smem_u32(&empty_bar[stage]), (uint16_t)3);           // stage retired in both CTAs
              if (ks == k_stages - 1) ptx::umma2_commit_mc(ptx::smem_u32(&tfull_bar[parity]), (uint16_t)3);   // accumulator complete in both
            }
            __syncwarp();
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
    } else if (relay && warp == 1) {
      // peer CTA, relay mode: forward "my half of stage s has landed" to the leader's full barrier
      const uint32_t full_leader = ptx::mapa(ptx::smem_u32(full_bar), 0);
      int stage = 0; uint32_t phase = 0;
      int islot = 0; uint32_t iphase = 0;
      for (;;) {
        const int pr = next_pair(islot, iphase);
        if (pr >= n_pairs) break;
        const int n_st = (items[2 * pr].obj_count + kTsBlockN - 1) / kTsBlockN * k_stages;
        for (int i = 0; i < n_st; ++i) {
          ptx::mbar_wait(ptx::smem_u32(&full_bar[stage]), phase);
          if (lane == 0) arrive_x(full_leader + (uint32_t)stage * 8u);
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else {
    // ===================== epilogue (4 warps per CTA, thread = query row of this CTA's tile) =====================
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t afull_leader = ptx::mapa(ptx::smem_u32(afull_bar), 0);
    const uint32_t tempty_leader = ptx::mapa(ptx::smem_u32(tempty_bar), 0);
    RowBuf<KP> list;
    float* inv_w = s_inv + quarter * 2 * kTsBlockN;    // this warp's private copy of the tile's inverse norms
    int acc = 0; uint32_t acc_phase = 0;
    int islot = 0; uint32_t iphase = 0;
    for (;;) {
      const int pr = next_pair(islot, iphase);
      if (pr >= n_pairs) break;
      const WorkItem wi = items[2 * pr + crank];
      // Half tile (kItemHalf, <= 64 rows per CTA, M = 128 pair MMAs): tensor memory lanes l and l + 64 both hold
      // query row l & 63 ("duplicated" A layout); lanes 0-63 receive the scores of the tile's objects 0-31,
      // lanes 64-127 those of objects 32-63 (accumulator columns 0-31 either way).  Threads l and l + 64 keep
      // separate candidate lists over their halves of the objects, merged when the item ends.
      const bool half = (wi.flags & kItemHalf) != 0;
      const int qrow = half ? (row & 63) : row;
      const int hoff = half ? (row >> 6) * (kTsBlockN / 2) : 0;
      const bool live = qrow < wi.rows && !(dbg_flags & 1);
      const int qi = live ? slot_pair[wi.slot_base + qrow] / nb : -1;
      // ---- this CTA's query rows -> its tensor memory (every MMA of the previous item has completed: this
      //      thread has seen the tfull barrier of its last tile, and that commit covers both CTAs) ----
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
        if (lane == 0) arrive_x(afull_leader);
      }
      list.reset(l_sc, l_ps, row, live, live ? thr_decode(__ldcg(qthr + qi)) : kNegInf);
      const int n_tiles = (wi.obj_count + kTsBlockN - 1) / kTsBlockN;
      // inverse norms are fetched one tile ahead (each lane owns two columns of its warp's copy);
      // accumulator column c = object obj_begin + 64 t + c (columns 0-31 from the leader's half, 32-63 the peer's)
      float inv_n0 = (lane < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane) : 0.f;
      float inv_n1 = (lane + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane + 32) : 0.f;
      uint32_t pending = 0u;                       // result of the last threshold exchange, consumed one tile later
      for (int t = 0; t < n_tiles; ++t) {
        const int col0 = t * kTsBlockN;
        const int ncols = max(0, min(kTsBlockN, wi.obj_count - col0));
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
        // accumulator stage is in registers -> hand it back to the leader's MMA warp right away
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) arrive_x(tempty_leader + (uint32_t)acc * 8u);
        if (dbg_flags & 2) { if (++acc == 2) { acc = 0; acc_phase ^= 1; } continue; }   // experiment: nothing scored
        float s[kTsBlockN];
#pragma unroll
        for (int c4 = 0; c4 < kTsBlockN / 4; ++c4) {
          const float4 iv = *reinterpret_cast<const float4*>(inv + ((hoff + c4 * 4) & (kTsBlockN - 1)));
          const uint32_t* v = (c4 < 8) ? (v0 + c4 * 4) : (v1 + (c4 - 8) * 4);
          s[c4 * 4 + 0] = __uint_as_float(v[0]) * iv.x;
          s[c4 * 4 + 1] = __uint_as_float(v[1]) * iv.y;
          s[c4 * 4 + 2] = __uint_as_float(v[2]) * iv.z;
          s[c4 * 4 + 3] = __uint_as_float(v[3]) * iv.w;
        }
        {
          // columns past the chunk's end; half tile: only accumulator columns 0-31 exist, object = hoff + column
          const int nvalid = half ? min(kTsBlockN / 2, ncols - hoff) : ncols;
          if (nvalid != kTsBlockN) {
#pragma unroll
            for (int c = 0; c < kTsBlockN; ++c) if (c >= nvalid) s[c] = kNegInf;
          }
        }
        uint32_t gmask = 0u;
#pragma unroll
        for (int g = 0; g < kTsBlockN / 8; ++g) {
          const float* p = s + g * 8;
          const float m = fmaxf(fmaxf(fmaxf(p[0], p[1]), fmaxf(p[2], p[3])), fmaxf(fmaxf(p[4], p[5]), fmaxf(p[6], p[7])));
          gmask |= __any_sync(0xffffffffu, m > list.thr) ? (1u << g) : 0u;
        }
        if (gmask != 0u) {
#pragma unroll
          for (int g = 0; g < kTsBlockN / 8; ++g) {
            if (gmask & (1u << g)) {
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (s[g * 8 + j] > list.thr) list.append(l_sc, l_ps, row, s[g * 8 + j], pos0 + hoff + g * 8 + j);
              if (__any_sync(0xffffffffu, list.cnt > kPend - 8)) list.prune(l_sc, l_ps, row);
            }
          }
        }
        if ((t + 1) % kTsRefresh == 0 || t + 1 == n_tiles) {
          list.prune(l_sc, l_ps, row);
          if (live) pending = atomicMax(qthr + qi, list.lmin > kNegInf ? thr_encode(list.lmin) : 0u);
        }
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      if (half) {
        // merge the list of thread l + 64 (objects 32-63 of every tile) into thread l's; named barrier over the
        // four epilogue warps on both sides (lists complete / merge done before the next item resets them)
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (row < 64 && live) {
#pragma unroll 1
          for (int j = 0; j < KP; ++j) {
            const float sv = l_sc[j * kTileQ + row + 64];
            if (sv > list.lmin) list.insert_now(l_sc, l_ps, row, sv, l_ps[j * kTileQ + row + 64]);
          }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (row < 64 && live)
          list.flush(l_sc, l_ps, row, plists + (size_t)(wi.plist_base + (int64_t)row * wi.plist_stride) * KP);
      } else if (live) {
        list.flush(l_sc, l_ps, row, plists + (size_t)(wi.plist_base + (int64_t)row * wi.plist_stride) * KP);
      }
    }
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->k3_clk = (unsigned long long)(clock64() - clk0);
    sc->k3_ns = ptx::globaltimer_ns() - ns0;
  }
  ptx::cluster_sync();                               // no CTA exits while the peer may still signal it or read its memory
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc2<512>(tmem_base); }
}

}  // namespace lmi
