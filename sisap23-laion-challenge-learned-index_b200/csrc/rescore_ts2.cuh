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
//                 both SMs.
//
// Every tcgen05.alloc/mma/commit of a kernel must use one cta_group width, so this kernel cannot also
// run single-CTA items: the planner's pair mode (plan.cuh, `pair_mode`) emits shared pairs only.  An odd
// last query tile becomes two HALF tiles (<= 64 rows per CTA, kItemHalf) scored with M = 128 pair MMAs
// (64 rows per CTA, half the tensor time): lanes l and l + 64 of a CTA's tensor memory hold query row
// l & 63, lanes 0-63 receive the scores of the tile's objects 0-31, lanes 64-127 those of objects 32-63
// (accumulator columns 0-31 either way, so the epilogue of a half tile reads and scores 32 columns).
//
// Warp roles (128 + 128 * SETS threads): warpgroup 0 = control (warp 0 TMA producer, warps 1 and 2 MMA
// issuers of the leader CTA -- even / odd tiles --, warp 3 idle); then SETS epilogue warpgroups.  With
// SETS = 2 (k' <= 32) epilogue set s owns accumulator s, i.e. every other tile: the epilogue of a tile
// (tcgen05.ld, 64 scores per row, filter, rare appends) is a latency chain of ~2,000 cycles on one warp
// per scheduler -- longer than the 1,536 (768 for half tiles) tensor-pipe cycles of the tile -- and a second
// warp per scheduler gives every tile two tile periods.  Each set keeps its own candidate list per row
// (they exchange their list minima through shared memory so both filter with the better bound) and the
// lists are merged in shared memory when the item ends, so one list per (row, chunk) reaches K4 as before.
// Registers move from the control warpgroup to the epilogue warpgroups with setmaxnreg.
//
// Cross-CTA protocol (leader = cluster rank 0):
//   full_bar[s]   (leader's)  : leader's producer expects the bytes of BOTH halves; the peer's TMA
//                               signals the leader's barrier (cp.async.bulk.tensor ... cta_group::2)
//   empty_bar[s]  (both)      : tcgen05.commit.cta_group::2 multicast from the leader's issuing warp
//   tfull_bar[a]  (both)      : same (accumulator a complete in both CTAs' tensor memory)
//   tempty_bar[a] (leader's)  : 4 epilogue warps x 2 CTAs arrive (the peer's remotely, RELAXED: the hand-over
//                               is "tcgen05.wait::ld has returned"; a release arrive on a shared::cluster
//                               address costs MEMBAR.ALL.GPU + CGAERRBAR per warp and tile)
//   afull_bar     (leader's)  : all epilogue warps of both CTAs, "this item's query rows are in tensor memory"
//   item ring                 : as in rescore_ts (leader's producer publishes into both CTAs)
#pragma once
#include "rescore_ts.cuh"

namespace lmi {

constexpr int kT2HalfN = kTsBlockN / 2;                            // objects per CTA and tile
constexpr int kT2AtomBytes = kT2HalfN * kTsAtomK * 2;              // 4 KB: 32 objects x 64 halves, 128B swizzle
constexpr int kT2StageBytes = kTsAtomsPerStage * kT2AtomBytes;     // 16 KB per CTA
constexpr int kT2CtrlRegs = 72, kT2EpiRegs = 216;                  // setmaxnreg split (SETS = 2): 128*72 + 256*216 = 64,512 of 65,536

// setmaxnreg moves registers inside the CTA's OWN allocation (SASS: USETMAXREG.TRY_ALLOC.CTAPOOL spins until the
// CTA pool has them): 384 threads launch with 168 registers each, so the split must not exceed 384 * 168 = 64,512 --
// a split that sums to the SM's 65,536 never gets its last warp served and the kernel hangs.
static_assert(128 * kT2CtrlRegs + 256 * kT2EpiRegs <= 384 * 168, "setmaxnreg split exceeds the CTA's register allocation");
static_assert(kT2CtrlRegs % 8 == 0 && kT2EpiRegs % 8 == 0, "setmaxnreg takes multiples of 8");

constexpr int ts2_threads(int sets) { return 128 + 128 * sets; }

template <int KP, int STAGES, int SETS>
constexpr size_t ts2_smem_bytes() {
  return 1024 /*alignment slack*/ + (size_t)STAGES * kT2StageBytes +
         (size_t)SETS * (sizeof(float) + sizeof(int32_t)) * (KP + kPend) * kTileQ /*candidate lists*/ +
         (size_t)SETS * 4 * 2 * kTsBlockN * sizeof(float) /*inverse norms, per epilogue warp, double-buffered*/ +
         (size_t)SETS * kTileQ * sizeof(float) /*list minima shared between the sets*/ + 512 /*barriers + tmem ptr + item ring*/;
}

// Scores NC accumulator columns of one row against the row's filter threshold and appends the survivors to
// the row's pending area.  v: raw dot products (fp32 bits) from tensor memory; inv: inverse norms of the
// NC objects (shared memory); nvalid: columns that exist; pos_base: CSR position of column 0.
template <int KP, int NC>
__device__ __forceinline__ void ts2_score_cols(RowBuf<KP>& list, float* l_sc, int32_t* l_ps, int row,
                                               const uint32_t (&v)[NC], const float* inv, int nvalid, int pos_base) {
  float s[NC];
#pragma unroll
  for (int c4 = 0; c4 < NC / 4; ++c4) {
    const float4 iv = *reinterpret_cast<const float4*>(inv + c4 * 4);
    s[c4 * 4 + 0] = __uint_as_float(v[c4 * 4 + 0]) * iv.x;
    s[c4 * 4 + 1] = __uint_as_float(v[c4 * 4 + 1]) * iv.y;
    s[c4 * 4 + 2] = __uint_as_float(v[c4 * 4 + 2]) * iv.z;
    s[c4 * 4 + 3] = __uint_as_float(v[c4 * 4 + 3]) * iv.w;
  }
  if (nvalid < NC) {                     // warp-uniform: last tile of a chunk
#pragma unroll
    for (int c = 0; c < NC; ++c) if (c >= nvalid) s[c] = kNegInf;
  }
  // straight-line part first (group maxima + votes), then the rare paths: a group of 8 columns is entered
  // only if some row of the warp has a candidate in it (warp-uniform branch)
  uint32_t gmask = 0u;
#pragma unroll
  for (int g = 0; g < NC / 8; ++g) {
    const float* p = s + g * 8;
    const float m = fmaxf(fmaxf(fmaxf(p[0], p[1]), fmaxf(p[2], p[3])), fmaxf(fmaxf(p[4], p[5]), fmaxf(p[6], p[7])));
    gmask |= __any_sync(0xffffffffu, m > list.thr) ? (1u << g) : 0u;
  }
  if (gmask != 0u) {
#pragma unroll
    for (int g = 0; g < NC / 8; ++g) {
      if (gmask & (1u << g)) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (s[g * 8 + j] > list.thr) list.append(l_sc, l_ps, row, s[g * 8 + j], pos_base + g * 8 + j);
        if (__any_sync(0xffffffffu, list.cnt > kPend - 8)) list.prune(l_sc, l_ps, row);   // room for the next group
      }
    }
  }
}

template <int KP, int STAGES, int SETS>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(128 + 128 * SETS, 1)
rescore_ts2_kernel(const __grid_constant__ CUtensorMap tmap_xh,     // bucket store [n_local, d], box 32 x 64
                   const __grid_constant__ CUtensorMap tmap_x3h,    // 3-D view, box 64 halves x 32 objects x 4 K-atoms = one stage
                   const uint16_t* __restrict__ q,                  // [nq, d] fp16 bits (original order)
                   const int32_t* __restrict__ slot_pair,           // slot -> pair index q*nb + r
                   int nb, int d, const float* __restrict__ inv_norm,
                   const WorkItem* __restrict__ items, PlanScalars* __restrict__ sc,
                   uint32_t* __restrict__ qthr,                     // [nq] shared per-query threshold
                   Cand* __restrict__ plists,
                   int stage_box,                                   // bit 0: one 3-D TMA box fills a stage (d % 256 == 0); bit 1: L2 eviction hints per item; bit 2: lock-step followers
                   int dbg_flags,                                   // experiments (LMI_EXPERIMENTS builds only)
                   int32_t* __restrict__ progress) {                // [pairs] lock-step option (stage_box bit 2): tiles requested so far by a leading pair
  static_assert(SETS == 1 || SETS == 2, "one or two epilogue warp sets");
  static_assert(STAGES >= 4, "the issuer warps' phase tracking needs a ring longer than one tile (3 stages at d = 768)");
  static_assert(2 * STAGES * 8 + 9 * 8 + 4 * kTsItemRing * 4 + 16 <= 512, "barrier area too small");
  constexpr int kEpiThreads = 128 * SETS;
  constexpr int kListFloats = (KP + kPend) * kTileQ;
  extern __shared__ unsigned char ts2_smem_raw[];
  // 1024-byte alignment (128B swizzle atom); the offset is the same in both CTAs, so one shared-memory
  // descriptor addresses the leader's and the peer's half of a tile
  unsigned char* smem = ts2_smem_raw + ((1024u - (ptx::smem_u32(ts2_smem_raw) & 1023u)) & 1023u);
  unsigned char* stage_base = smem;
  float* lists_sc = reinterpret_cast<float*>(smem + (size_t)STAGES * kT2StageBytes);       // [SETS][KP + kPend][128]
  int32_t* lists_ps = reinterpret_cast<int32_t*>(lists_sc + SETS * kListFloats);           // [SETS][KP + kPend][128]
  float* s_inv = reinterpret_cast<float*>(lists_ps + SETS * kListFloats);                   // [4 * SETS warps][2][64]
  float* s_rowthr = s_inv + SETS * 4 * 2 * kTsBlockN;                                       // [SETS][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(s_rowthr + SETS * kTileQ);
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

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_xh);
    ptx::prefetch_tensormap(&tmap_x3h);
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&empty_bar[s]), 1); }
    // an accumulator is drained by 4 epilogue warps in each CTA (SETS = 2: the set that owns it)
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty_bar[s]), 8); }
    ptx::mbar_init(ptx::smem_u32(afull_bar), 8 * SETS);
    // item ring: the leader's iempty collects its 2 MMA + 4 * SETS epilogue warps and the peer's producer +
    // 4 * SETS epilogue warps (the peer's MMA warps idle: only the leader issues)
    for (int s = 0; s < kTsItemRing; ++s) { ptx::mbar_init(ptx::smem_u32(&ifull_bar[s]), 1); ptx::mbar_init(ptx::smem_u32(&iempty_bar[s]), 3 + 8 * SETS); }
    ptx::fence_barrier_init();
  }
  if (warp == 1) ptx::tmem_alloc2<512>(ptx::smem_u32(tmem_ptr));     // warp 1 of BOTH CTAs
  ptx::tc_fence_before();
  __syncthreads();
  ptx::cluster_sync();                               // the peer's barriers and tensor memory exist too
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;
  const uint32_t tmem_d = tmem_base + kTsACols;
  // SETS = 2 runs 384 threads and the epilogue code needs more than 65536 / 384 registers: every role branch
  // below starts with its setmaxnreg (all four warps of a warpgroup execute one; ptxas allocates per branch)
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
    if constexpr (SETS == 2) ptx::setmaxnreg_dec<kT2CtrlRegs>();
    if (lane == 0) {
      const uint32_t full_leader = ptx::mapa(ptx::smem_u32(full_bar), 0);
      int stage = 0; uint32_t phase = 0;
      int islot = 0; uint32_t iphase = 0;
      // L2 eviction priority per item (option "l2_hint"): a chunk's last (or only) reader loads with evict_first, so the
      // streams nobody will read again do not push the chunks that a second, equal-speed reader is about to re-read out of L2
      const uint64_t pol_first = ptx::l2_policy_evict_first(), pol_normal = ptx::l2_policy_evict_normal();
      const bool l2_hint = (stage_box & 2) != 0;
      // Lock-step option: a half pair consumes its chunk twice as fast as the full pair listed just before it (same
      // chunk), runs ahead of it and both read the chunk from DRAM.  As a FOLLOWER it asks for tile t only once the
      // leader has asked for it, so its loads hit L2 (or merge with the leader's miss) and the chunk crosses the HBM
      // interface once; it then runs at the leader's pace with its tensor pipe half idle -- a win where HBM, not the
      // tensor pipe or the power cap, bounds the kernel (shards).  The counter is a timing hint, not a dependency: the
      // leader never waits, and a follower that sees no progress for kLockSpins polls runs free for the rest of its item.
      const bool lockstep = (stage_box & 4) != 0;
      constexpr int kLockSpins = 64;
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
        const bool leads = lockstep && (wi.flags & kItemLeads) && crank == 0;
        bool follows = lockstep && (wi.flags & kItemFollows) && pr > 0;
        volatile int32_t* lead_prog = progress + (follows ? pr - 1 : pr);
        int lead_seen = 0;
        for (int t = 0; t < n_tiles; ++t) {
          if (follows && lead_seen <= t) {
            int spins = 0;
            do { lead_seen = *lead_prog; } while (lead_seen <= t && ++spins < kLockSpins);
            if (lead_seen <= t) follows = false;
          }
          int obj0 = wi.obj_begin + t * kTsBlockN + (int)crank * kT2HalfN;
          // experiment 256: half-pair items re-read a 16-tile window (always an L2 hit) instead of streaming their chunk a
          // second time -- isolates what the second reader's DRAM traffic costs (results invalid)
          if (LMI_DBG(dbg_flags, 256) && (wi.flags & kItemHalf)) obj0 = wi.obj_begin + (t & 15) * kTsBlockN + (int)crank * kT2HalfN;
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);    // leader's commit, multicast to both CTAs
            if (LMI_DBG(dbg_flags, 4)) {                             // experiment: no loads at all
              if (crank == 0) ptx::mbar_arrive(ptx::smem_u32(&full_bar[stage]));
              if (++stage == STAGES) { stage = 0; phase ^= 1; }
              continue;
            }
            const uint32_t sb = ptx::smem_u32(stage_base + (size_t)stage * kT2StageBytes);
            if (crank == 0) ptx::mbar_arrive_expect_tx(ptx::smem_u32(&full_bar[stage]), (uint32_t)(2 * atoms * kT2AtomBytes));
            const uint32_t fb = full_leader + (uint32_t)stage * 8u;
            if ((stage_box & 1) && l2_hint) ptx::tma_load_3d_cg2_hint(sb, &tmap_x3h, fb, 0, obj0, ks * kTsAtomsPerStage, (wi.flags & kItemLastReader) ? pol_first : pol_normal);
            else if (stage_box & 1) ptx::tma_load_3d_cg2(sb, &tmap_x3h, fb, 0, obj0, ks * kTsAtomsPerStage);
            else
              for (int a = 0; a < atoms; ++a)
                ptx::tma_load_2d_cg2(sb + a * kT2AtomBytes, &tmap_xh, fb, (ks * kTsAtomsPerStage + a) * kTsAtomK, obj0);
            if (++stage == STAGES) { stage = 0; phase ^= 1; }
          }
          if (leads) *lead_prog = (t + 1 == n_tiles) ? 0x7fffffff : t + 1;
        }
      }
      // Producer tail: no CTA exits while commit arrivals for its stages may still be in flight
      for (int s = 0; s < STAGES; ++s) {
        ptx::mbar_wait(ptx::smem_u32(&empty_bar[stage]), phase ^ 1);
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 || warp == 2) {
    // ===================== MMA issuers (leader CTA only; two warps alternate tiles) =====================
    if constexpr (SETS == 2) ptx::setmaxnreg_dec<kT2CtrlRegs>();
    if (crank == 0) {
      constexpr uint32_t idesc_full = ptx::umma_idesc_f16(2 * kTileQ, kTsBlockN);  // M = 256 over the pair, N = 64
      constexpr uint32_t idesc_half = ptx::umma_idesc_f16(kTileQ, kTsBlockN);      // M = 128 over the pair (64 rows per CTA)
      const int parity = (warp == 2) ? 1 : 0;
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
        ptx::mbar_wait(ptx::smem_u32(afull_bar), a_phase);     // both CTAs' query rows are in tensor memory
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
          ptx::mbar_wait(ptx::smem_u32(&tempty_bar[parity]), ((g >> 1) & 1u) ^ 1u);   // drained in BOTH CTAs
          ptx::tc_fence_after();
          for (int ks = 0; ks < k_stages; ++ks) {
            const int atoms = min(kTsAtomsPerStage, k_atoms - ks * kTsAtomsPerStage);
            ptx::mbar_wait(ptx::smem_u32(&full_bar[stage]), phase);         // both halves have landed
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              const uint64_t b0 = ptx::umma_desc_k_sw128(ptx::smem_u32(stage_base + (size_t)stage * kT2StageBytes));
              const uint32_t a0 = tmem_base + (uint32_t)(ks * kTsAtomsPerStage * 32);   // 32 columns per atom
              // A: 8 TMEM columns per K=16 slice; B: +32 bytes inside the swizzle span, next atom +4 KB
              if (LMI_DBG(dbg_flags, 8)) {
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
    }
  } else if (warp == 3) {
    if constexpr (SETS == 2) ptx::setmaxnreg_dec<kT2CtrlRegs>();      // idle warp: completes the control warpgroup
  } else {
    // ===================== epilogue (SETS x 4 warps per CTA, thread = query row of this CTA's tile) =====================
    if constexpr (SETS == 2) ptx::setmaxnreg_inc<kT2EpiRegs>();
    const int set = (warp - 4) >> 2;              // epilogue warp set (SETS = 2: owns accumulator `set`)
    const int quarter = warp & 3;                 // TMEM lane quarter this warp may access
    const int row = quarter * 32 + lane;
    const uint32_t lane_addr = (uint32_t)(quarter * 32) << 16;
    const uint32_t afull_leader = ptx::mapa(ptx::smem_u32(afull_bar), 0);
    const uint32_t tempty_leader = ptx::mapa(ptx::smem_u32(tempty_bar), 0);
    float* l_sc = lists_sc + set * kListFloats;
    int32_t* l_ps = lists_ps + set * kListFloats;
    RowBuf<KP> list;
    float* inv_w = s_inv + (warp - 4) * 2 * kTsBlockN;   // this warp's private copy of the tile's inverse norms (2 tiles)
    volatile float* my_thr = s_rowthr + set * kTileQ + row;
    volatile float* other_thr = s_rowthr + (SETS - 1 - set) * kTileQ + row;
    int acc = (SETS == 2) ? set : 0; uint32_t acc_phase = 0;
    int ibuf = 0;
    uint32_t g0 = 0;                              // running tile number at the start of the item (SETS = 2: tile parity = owner set)
    int islot = 0; uint32_t iphase = 0;
    auto epi_bar = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(kEpiThreads) : "memory"); };
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
      const bool live = qrow < wi.rows && !LMI_DBG(dbg_flags, 1);
      const int qi = live ? slot_pair[wi.slot_base + qrow] / nb : -1;
      // ---- this CTA's query rows -> its tensor memory (every MMA of the previous item has completed: the set
      //      that drained the item's last tile saw its tfull barrier, and the merge barrier / the afull barrier
      //      order the other set behind it).  The sets split the 64-half blocks of a row. ----
      {
        const uint4* src = reinterpret_cast<const uint4*>(q + (size_t)(live ? qi : 0) * d);
        const int n_blk = d / 64;                 // 64 halves = 32 columns per block
        const int blk0 = set * n_blk / SETS, blk1 = (set + 1) * n_blk / SETS;
        uint4 cur[8];
#pragma unroll
        for (int v = 0; v < 8; ++v) cur[v] = live ? src[blk0 * 8 + v] : make_uint4(0, 0, 0, 0);
        for (int blk = blk0; blk < blk1; ++blk) {
          uint4 nxt[8];
          if (blk + 1 < blk1) {
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
        if constexpr (SETS == 2) *my_thr = kNegInf;
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive_remote(afull_leader);
      }
      list.reset(l_sc, l_ps, row, live, live ? thr_decode(__ldcg(qthr + qi)) : kNegInf);
      const int n_tiles = (wi.obj_count + kTsBlockN - 1) / kTsBlockN;
      // tiles of this item that this set scores: all (SETS = 1) or those whose running number has the set's parity
      const int t_first = (SETS == 2) ? (int)((set - g0) & 1u) : 0;
      // inverse norms are fetched one (own) tile ahead (each lane owns two columns of its warp's copy);
      // accumulator column c = object obj_begin + 64 t + c (columns 0-31 from the leader's half, 32-63 the peer's)
      const int c_first = t_first * kTsBlockN + lane;
      float inv_n0 = (c_first < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_first) : 0.f;
      float inv_n1 = (c_first + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_first + 32) : 0.f;
      uint32_t pending = 0u;                       // result of the last threshold exchange, consumed one tile later
      int n_done = 0;
      for (int t = t_first; t < n_tiles; t += SETS) {
        const int col0 = t * kTsBlockN;
        const int ncols = max(0, min(kTsBlockN, wi.obj_count - col0));
        float* inv = inv_w + ibuf * kTsBlockN;
        ibuf ^= 1;
        inv[lane] = inv_n0;
        inv[lane + 32] = inv_n1;
        {
          const int c_next = col0 + SETS * kTsBlockN + lane;
          inv_n0 = (c_next < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_next) : 0.f;
          inv_n1 = (c_next + 32 < wi.obj_count) ? __ldg(inv_norm + wi.obj_begin + c_next + 32) : 0.f;
        }
        list.raise_seed(thr_decode(pending));
        __syncwarp();
        ptx::mbar_wait(ptx::smem_u32(&tfull_bar[acc]), acc_phase);
        ptx::tc_fence_after();
        const uint32_t taddr = tmem_d + lane_addr + (uint32_t)(acc * kTsBlockN);
        const int pos0 = wi.obj_begin + col0;
        uint32_t v[kTsBlockN];
        ptx::tmem_ld_32x32b_x32(taddr, reinterpret_cast<uint32_t(&)[32]>(v[0]));
        if (!half) ptx::tmem_ld_32x32b_x32(taddr + 32, reinterpret_cast<uint32_t(&)[32]>(v[32]));   // half tile: only columns 0-31 exist
        ptx::tmem_ld_wait();
        // accumulator is in registers -> hand it back to the leader's MMA warp right away
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (LMI_DBG(dbg_flags, 64)) ptx::mbar_arrive_remote(tempty_leader + (uint32_t)acc * 8u);   // experiment: release arrive
          else ptx::mbar_arrive_remote_relaxed(tempty_leader + (uint32_t)acc * 8u);
        }
        if (!LMI_DBG(dbg_flags, 2)) {                              // (experiment 2: accumulator drained, nothing scored)
          if (half)
            ts2_score_cols<KP, kTsBlockN / 2>(list, l_sc, l_ps, row, reinterpret_cast<const uint32_t(&)[kTsBlockN / 2]>(v[0]),
                                             inv + hoff, min(kTsBlockN / 2, ncols - hoff), pos0 + hoff);
          else
            ts2_score_cols<KP, kTsBlockN>(list, l_sc, l_ps, row, v, inv, ncols, pos0);
        }
        ++n_done;
        if (n_done % kTsRefresh == 0 || t + SETS >= n_tiles) {
          // publish this row's bound / pick up the other CTAs' (the atomic's old value is read next tile) and the other set's
          list.prune(l_sc, l_ps, row);
          if (live) pending = atomicMax(qthr + qi, list.lmin > kNegInf ? thr_encode(list.lmin) : 0u);
          if constexpr (SETS == 2) { *my_thr = list.lmin; list.raise_seed(*other_thr); }
        }
        if constexpr (SETS == 2) { acc_phase ^= 1; }
        else { if (++acc == 2) { acc = 0; acc_phase ^= 1; } }
      }
      g0 += (uint32_t)n_tiles;
      // ---- one list per (row, chunk) leaves the kernel: fold the other set's list and, for a half tile, the list of
      //      thread l + 64 (objects 32-63 of every tile) into the list of set 0 / thread l ----
      if (SETS == 2 || half) {
        epi_bar();                                 // every list of this item is complete
        const bool owner = set == 0 && (!half || row < 64);
        if (owner && live) {
#pragma unroll 1
          for (int src_list = 1; src_list < SETS * 2; ++src_list) {
            const int s2 = src_list >> 1, h2 = src_list & 1;
            if (h2 && !half) continue;
            const float* o_sc = lists_sc + s2 * kListFloats;
            const int32_t* o_ps = lists_ps + s2 * kListFloats;
            const int o_row = row + 64 * h2;
#pragma unroll 1
            for (int j = 0; j < KP; ++j) {
              const float sv = o_sc[j * kTileQ + o_row];
              if (sv > list.lmin) list.insert_now(l_sc, l_ps, row, sv, o_ps[j * kTileQ + o_row]);
            }
          }
        }
        epi_bar();                                 // merged before any list is reset for the next item
        if (owner && live)
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
