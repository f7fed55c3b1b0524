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
//   warps 1,6: MMA issuers    -- an elected thread issues tcgen05.mma (M=128, N=64, K=16), A from
//                                TMEM, B from shared memory; D double-buffered in TMEM, warp 1
//                                owns the even tiles / accumulator 0, warp 6 the odd ones
//   warps 2-5: epilogue       -- per item: write the item's 128 query rows into TMEM
//                                (tcgen05.st); per tile: tcgen05.ld of the finished accumulator
//                                (thread = TMEM lane = query row), scale by inv_norm, compare
//                                against the row's running threshold, rare insertion
// Pipelines: smem full/empty mbarriers (TMA <-> MMA), TMEM full/empty mbarriers (MMA <->
// epilogue), a_full (epilogue -> MMA: the item's queries are in TMEM).
//
// 2-CTA clusters: the grid is launched as clusters of two CTAs; a cluster takes one PAIR of work
// list entries (plan.cuh).  For a shared pair (two query tiles of the same object chunk) each
// CTA loads half of every object box and TMA-multicasts it into both CTAs' shared memory, which
// halves the L2->SM traffic (the bound of the single-CTA version: 128 FLOP/B against ~12 TB/s).
// Plain pairs load unicast.  Either way the two CTAs consume the stages in lock-step: a stage is
// refilled only when both MMA warps have retired it (tcgen05.commit multicast on the empty barrier).
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
constexpr int kTsThreads = 224;                     // producer, MMA issuer 0, 4 epilogue warps, MMA issuer 1
constexpr int kTsACols = 384;                       // TMEM columns reserved for the query tile (d <= 768)
constexpr int kTsItemRing = 2;                      // depth of the producer -> consumers work item ring
constexpr int kTsPrefetch = 8;                      // tiles the L2 prefetch runs ahead of the loads
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
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kTsThreads, 1)
rescore_ts_kernel(const __grid_constant__ CUtensorMap tmap_x,      // bucket store [n_local, d], box 64 x 64
                  const __grid_constant__ CUtensorMap tmap_xh,     // same, box 32 x 64 (half of a shared tile)
                  const __grid_constant__ CUtensorMap tmap_pf,     // same, box 64 x 256 unswizzled (L2 prefetch)
                  const __grid_constant__ CUtensorMap tmap_x3,     // 3-D view, box 64 halves x 64 objects x 4 K-atoms = one stage
                  const uint16_t* __restrict__ q,                  // [nq, d] fp16 bits (original order)
                  const int32_t* __restrict__ slot_pair,           // slot -> pair index q*nb + r
                  int nb, int d, const float* __restrict__ inv_norm,
                  const WorkItem* __restrict__ items, PlanScalars* __restrict__ sc,
                  uint32_t* __restrict__ qthr,                     // [nq] shared per-query threshold
                  Cand* __restrict__ plists,
                  int stage_box_arg,                               // 1: one 3-D TMA box fills a stage (d % 256 == 0; unicast)
                  int dbg_flags) {                                 // experiments (LMI_EXPERIMENTS builds only)
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
  const uint32_t crank = ptx::cluster_ctarank();                     // 0 = leader (fetches the work items)
  const int n_pairs = sc->n_items >> 1;
  const int k_atoms = d / kTsAtomK;                                  // 12 for clip768
  const int k_stages = (k_atoms + kTsAtomsPerStage - 1) / kTsAtomsPerStage;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_x);
    ptx::prefetch_tensormap(&tmap_xh);
    // dbg_flags & 256 (experiment): no per-stage coupling between the two CTAs -- unicast loads only, a stage is
    // retired by this CTA's own commit; the pair still starts and ends its work item together
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&empty_bar[s]), LMI_DBG(dbg_flags, 256) ? 1 : 2); }
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty_bar[s]), 4); }
    ptx::mbar_init(ptx::smem_u32(afull_bar), 4);
    // item ring: the leader's producer publishes into both CTAs; its iempty collects the 6 local
    // consumer warps + the peer's 7 (the peer's producer is a consumer of the ring too)
    for (int s = 0; s < kTsItemRing; ++s) { ptx::mbar_init(ptx::smem_u32(&ifull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&iempty_bar[s]), 13); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(tmem_ptr));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();                               // the peer's barriers are initialised too
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_d = tmem_base + kTsACols;
  // CTA 0 reports the SM clock the kernel really ran at (cycles / nanoseconds between here and the end)
  long long clk0 = 0; unsigned long long ns0 = 0;
  if (blockIdx.x == 0 && threadIdx.x == 0) { clk0 = clock64(); ns0 = ptx::globaltimer_ns(); }

  // every consumer warp: next pair index from the ring, then release the slot to the leader
  // dbg_flags & 512 (experiment): only lane 0 of a warp polls a per-tile barrier, the other 31 lanes park at __syncwarp
  const bool poll_l0 = LMI_DBG(dbg_flags, 512);
  auto wait_w = [&](uint32_t bar, uint32_t parity) {
    if (poll_l0) { if (lane == 0) ptx::mbar_wait(bar, parity); __syncwarp(); }
    else ptx::mbar_wait(bar, parity);
  };
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
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0; uint32_t phase = 0;
      int islot = 0; uint32_t iphase = 0;
      for (;;) {
        // dynamic scheduling: clusters that become free one after the other pick up consecutive
        // pairs = everything that scores one object chunk, so those stream the chunk nearly in
        // lock-step and all but the first read it from L2
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
        const WorkItem wi = items[2 * pr + crank];
        const int peer_count = items[2 * pr + (crank ^ 1)].obj_count;
        const int n_tiles = (max(wi.obj_count, peer_count) + kTsBlockN - 1) / kTsBlockN;
        const bool stage_box = stage_box_arg != 0;
        const bool shared = (wi.flags & kItemShared) != 0 && !LMI_DBG(dbg_flags, 16 | 256) && !stage_box;
        // L2 prefetch kTsPrefetch tiles ahead: the loads that fill the smem ring then see L2 latency,
        // not HBM latency (the ring holds ~1.6 tiles: 160 KB / 1.8 us of loaded HBM latency is only
        // 45 B/clk per SM, the MMA consumes 64 B/clk).  One CTA of a shared pair prefetches for both.
        const bool do_pf = LMI_DBG(dbg_flags, 32) && (!shared || crank == 0);   // measured: no gain on B200 (kept as an experiment)
        auto prefetch_tile = [&](int t) {
          for (int ks = 0; ks < k_stages; ++ks)
            ptx::tma_prefetch_2d(&tmap_pf, ks * kTsAtomsPerStage * kTsAtomK, wi.obj_begin + t * kTsBlockN);
        };
        if (do_pf) for (int t = 0; t < min(kTsPrefetch, n_tiles); ++t) prefetch_tile(t);
        for (int t = 0; t < n_tiles; ++t) {
          const int obj0 = wi.obj_begin + t * kTsBlockN;
          if (do_pf && t + kTsPrefetch < n_tiles) prefetch_tile(t + kTsPrefetch);
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);    // retired by BOTH CTAs
            const uint32_t fb = ptx::smem_u32(&full_bar[stage]);
            if (LMI_DBG(dbg_flags, 4)) {                             // experiment: no loads at all
              ptx::mbar_arrive(fb);
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
              continue;
            }
            ptx::mbar_arrive_expect_tx(fb, (uint32_t)(atoms * kTsAtomBytes));
            const uint32_t sb = ptx::smem_u32(stage_base + (size_t)stage * kTsStageBytes);
            if (stage_box) {
              ptx::tma_load_3d(sb, &tmap_x3, fb, 0, obj0, ks * kTsAtomsPerStage);
            } else if (shared) {
              // this CTA fetches rows [32*crank, 32*crank+32) of the tile for both CTAs
              for (int a = 0; a < atoms; ++a)
                ptx::tma_load_2d_mc(sb + a * kTsAtomBytes + crank * (kTsAtomBytes / 2), &tmap_xh, fb,
                                    (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0 + (int)crank * (kTsBlockN / 2), (uint16_t)3);
            } else {
              for (int a = 0; a < atoms; ++a)
                ptx::tma_load_2d(sb + a * kTsAtomBytes, &tmap_x, fb, (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0);
            }
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
        }
      }
      // Producer tail: every stage must have been retired by BOTH CTAs before this CTA may exit.  The
      // peer's tcgen05.commit arrives on this CTA's empty barriers asynchronously (when its MMAs finish),
      // not ordered by the final cluster barrier; an arrive that lands in the shared memory of a CTA
      // that has already exited is a fault (seen as a rare "unspecified launch failure" with short items).
      for (int s = 0; s < STAGES; ++s) {
        ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 || warp == 6) {
    // ===================== MMA issuers (two warps) =====================
    // With N = 64 a tcgen05.mma is only 32 tensor-pipe cycles, and one thread cannot wait, compute
    // operands, issue 48 of them and commit inside the 1536 cycles of a tile (measured: ~2200).
    // So tiles alternate between two issuing warps: warp 1 takes the even tiles (accumulator 0),
    // warp 6 the odd ones (accumulator 1).  Their instruction streams are independent -- different
    // accumulators, disjoint smem stages, own barriers -- and the tensor pipe interleaves them.
    // Within a warp the control flow is warp-uniform and one elected lane issues.
    constexpr uint32_t idesc = ptx::umma_idesc_f16(kTileQ, kTsBlockN);
    const int parity = (warp == 6) ? 1 : 0;
    const uint32_t d_tmem = tmem_d + (uint32_t)(parity * kTsBlockN);
    uint32_t g = 0;                                // running tile number (all items); g & 1 = accumulator
    int stage = 0; uint32_t phase = 0;             // smem stage / phase of tile g's first K stage
    uint32_t a_phase = 0;
    int islot = 0; uint32_t iphase = 0;
    for (;;) {
      const int pr = next_pair(islot, iphase);
      if (pr >= n_pairs) break;
      const int n_tiles = (max(items[2 * pr].obj_count, items[2 * pr + 1].obj_count) + kTsBlockN - 1) / kTsBlockN;
      ptx::mbar_wait(ptx::smem_u32(afull_bar), a_phase);             // this item's queries are in TMEM
      a_phase ^= 1;
      for (int t = 0; t < n_tiles; ++t, ++g) {
        if ((int)(g & 1u) != parity) {
          // The other warp's tile: OBSERVE every stage of it becoming full instead of just counting past
          // them.  A parity wait for use n+1 of a stage returns at once while the barrier is still in
          // use n, so a warp may only wait for a phase after it has seen the previous one complete.  With
          // a ring shorter than two tiles a stage comes round again inside the other warp's tile (k' = 128
          // always; longer rings only if TMA completions were reordered by three stages) -- observing the
          // phases makes the scheme independent of both.  The stages are full long before this warp's own.
          for (int ks = 0; ks < k_stages; ++ks) {
            wait_w(ptx::smem_u32(&full_bar[stage]), phase);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          continue;
        }
        wait_w(ptx::smem_u32(&tempty_bar[parity]), ((g >> 1) & 1u) ^ 1u);
        ptx::tc_fence_after();
        for (int ks = 0; ks < k_stages; ++ks) {
          const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
          wait_w(ptx::smem_u32(&full_bar[stage]), phase);
          ptx::tc_fence_after();
          if (ptx::elect_one()) {
            const uint64_t b0 = ptx::umma_desc_k_sw128(ptx::smem_u32(stage_base + (size_t)stage * kTsStageBytes));
            const uint32_t a0 = tmem_base + (uint32_t)(ks * kTsAtomsPerStage * 32);   // 32 columns per atom
            // A: 16 halves = 8 TMEM columns per K=16 slice; B: +32 bytes inside the swizzle span
            // (+2 in the addr>>4 field), next atom +8 KB (+512)
            if (LMI_DBG(dbg_flags, 8)) {
            } else if (atoms == kTsAtomsPerStage) {
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
            if (LMI_DBG(dbg_flags, 256)) ptx::umma_commit(ptx::smem_u32(&empty_bar[stage]));
            else ptx::umma_commit_mc(ptx::smem_u32(&empty_bar[stage]), (uint16_t)3);   // stage retired here -> both CTAs' producers
            if (ks == k_stages - 1) ptx::umma_commit(ptx::smem_u32(&tfull_bar[parity]));   // accumulator complete
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
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
      const int pr = next_pair(islot, iphase);
      if (pr >= n_pairs) break;
      const WorkItem wi = items[2 * pr + crank];
      const int peer_count = items[2 * pr + (crank ^ 1)].obj_count;
      const bool live = row < wi.rows && !LMI_DBG(dbg_flags, 1);
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
      const int n_tiles = (max(wi.obj_count, peer_count) + kTsBlockN - 1) / kTsBlockN;   // the pair runs in lock-step
      // inverse norms are fetched one tile ahead (each lane owns two columns of its warp's copy)
      float inv_n0 = (lane < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane) : 0.f;
      float inv_n1 = (lane + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + lane + 32) : 0.f;
      uint32_t pending = 0u;                       // result of the last threshold exchange, consumed one tile later
      for (int t = 0; t < n_tiles; ++t) {
        const int col0 = t * kTsBlockN;
        const int ncols = max(0, min(kTsBlockN, wi.obj_count - col0));     // 0: padding tile of the shorter entry
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
        wait_w(ptx::smem_u32(&tfull_bar[acc]), acc_phase);
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
        // Straight-line part first (64 products, 8 group maxima, 8 votes: no branch, so the loads and
        // multiplies of all groups overlap), then the rare paths.  A group of 8 columns is entered only
        // if some row of the warp has a candidate in it (warp-uniform branch; the append path must
        // not be if-converted into the hot loop).
        if (LMI_DBG(dbg_flags, 2)) { if (++acc == 2) { acc = 0; acc_phase ^= 1; } continue; }   // experiment: accumulator drained, nothing scored
        float s[kTsBlockN];
#pragma unroll
        for (int c4 = 0; c4 < kTsBlockN / 4; ++c4) {
          const float4 iv = *reinterpret_cast<const float4*>(inv + c4 * 4);
          const uint32_t* v = (c4 < 8) ? (v0 + c4 * 4) : (v1 + (c4 - 8) * 4);
          s[c4 * 4 + 0] = __uint_as_float(v[0]) * iv.x;
          s[c4 * 4 + 1] = __uint_as_float(v[1]) * iv.y;
          s[c4 * 4 + 2] = __uint_as_float(v[2]) * iv.z;
          s[c4 * 4 + 3] = __uint_as_float(v[3]) * iv.w;
        }
        if (ncols != kTsBlockN) {
#pragma unroll
          for (int c = 0; c < kTsBlockN; ++c) if (c >= ncols) s[c] = kNegInf;
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
                if (s[g * 8 + j] > list.thr) list.append(l_sc, l_ps, row, s[g * 8 + j], pos0 + g * 8 + j);
              // keep room for the next group's (up to 8) appends
              if (__any_sync(0xffffffffu, list.cnt > kPend - 8)) list.prune(l_sc, l_ps, row);
            }
          }
        }
        // publish this row's bound / pick up the other CTAs' (the atomic's old value is read next tile)
        if (LMI_DBG(dbg_flags, 64)) { if (__any_sync(0xffffffffu, list.cnt >= 3)) list.prune(l_sc, l_ps, row); }
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
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    sc->k3_clk = (unsigned long long)(clock64() - clk0);
    sc->k3_ns = ptx::globaltimer_ns() - ns0;
  }
  ptx::cluster_sync();                               // no CTA exits while the peer may still signal it
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem_base); }
}

}  // namespace lmi
