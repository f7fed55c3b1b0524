// umma_tile.cu -- microbenchmark: cycles per 128x64x768 tile when one thread issues the exact MMA
// sequence of rescore_ts_kernel (A from 384 TMEM columns, B from 3 x 4 swizzled atoms in a 5-stage
// ring, two alternating accumulators) with NO barriers at all, and with one commit+wait per tile.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../sisap23-laion-challenge-learned-index_b200/csrc/ptx_sm100.cuh"
using namespace lmi;

template <int MODE>   // 0: free-running, commit once at the end; 1: commit per stage+tile (no waits); 2: + epilogue warps draining D
__global__ void __launch_bounds__(192, 1) tile_kernel(int tiles, long long* out) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = raw + ((1024u - (ptx::smem_u32(raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar, sbar[5], tfull[2], tempty[2];
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (160 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) {
    ptx::mbar_init(ptx::smem_u32(&bar), 1);
    for (int s = 0; s < 5; ++s) ptx::mbar_init(ptx::smem_u32(&sbar[s]), 1);
    for (int s = 0; s < 2; ++s) { ptx::mbar_init(ptx::smem_u32(&tfull[s]), 1); ptx::mbar_init(ptx::smem_u32(&tempty[s]), 4); }
    ptx::fence_barrier_init();
  }
  ptx::fence_proxy_async();
  if (warp == 1) ptx::tmem_alloc<512>(ptx::smem_u32(&tmem_ptr));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = tmem_ptr;
  constexpr uint32_t idesc = ptx::umma_idesc_f16(128, 64);
  if (warp == 1) {
    long long t0 = clock64();
    uint32_t leader;
    if (ptx::elect_one(leader)) {
      int stage = 0, acc = 0; uint32_t acc_phase = 0;
      for (int t = 0; t < tiles; ++t) {
        if (MODE == 2) { ptx::mbar_wait(ptx::smem_u32(&tempty[acc]), acc_phase ^ 1); ptx::tc_fence_after(); }
        const uint32_t d_tmem = tb + 384 + acc * 64;
        for (int ks = 0; ks < 3; ++ks) {
          const uint64_t b0 = ptx::umma_desc_k_sw128(ptx::smem_u32(smem + stage * 32768));
          const uint32_t a0 = tb + ks * 128;
          if (ks == 0) ptx::umma_f16_ts_c<false>(d_tmem, a0, b0, idesc);
          else         ptx::umma_f16_ts_c<true>(d_tmem, a0, b0, idesc);
#pragma unroll
          for (int i = 1; i < 16; ++i)
            ptx::umma_f16_ts_c<true>(d_tmem, a0 + i * 8, b0 + (uint64_t)((i >> 2) * 512 + (i & 3) * 2), idesc);
          if (MODE >= 1) ptx::umma_commit(ptx::smem_u32(&sbar[stage]));
          if (++stage == 5) stage = 0;
        }
        if (MODE == 2) ptx::umma_commit(ptx::smem_u32(&tfull[acc]));
        if (++acc == 2) { acc = 0; acc_phase ^= 1; }
      }
      ptx::umma_commit(ptx::smem_u32(&bar));
    }
    __syncwarp();
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    long long t1 = clock64();
    if (lane == 0) out[blockIdx.x] = t1 - t0;
  } else if (warp >= 2 && MODE == 2) {
    const int quarter = warp & 3;
    int acc = 0; uint32_t acc_phase = 0;
    float sink = 0.f;
    for (int t = 0; t < tiles; ++t) {
      ptx::mbar_wait(ptx::smem_u32(&tfull[acc]), acc_phase);
      ptx::tc_fence_after();
      uint32_t v0[32], v1[32];
      const uint32_t taddr = tb + ((uint32_t)(quarter * 32) << 16) + 384 + acc * 64;
      ptx::tmem_ld_32x32b_x32(taddr, v0);
      ptx::tmem_ld_32x32b_x32(taddr + 32, v1);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(ptx::smem_u32(&tempty[acc]));
#pragma unroll
      for (int j = 0; j < 32; ++j) sink += __uint_as_float(v0[j]) + __uint_as_float(v1[j]);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (sink == 123.456f) out[0] = 0;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tb); }
}

template <int MODE> void run(const char* name) {
  long long* d; cudaMalloc(&d, 148 * 8);
  const int tiles = 2000; const size_t smem = 1024 + 160 * 1024;
  cudaFuncSetAttribute(tile_kernel<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int w = 0; w < 2; ++w) tile_kernel<MODE><<<148, 192, smem>>>(tiles, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < 148; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("%-60s cycles/tile = %7.1f (ideal 1536) %s\n", name, (double)mx / tiles, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}
int main() {
  run<0>("free-running issue, single commit");
  run<1>("commit per stage");
  run<2>("commit per stage + tfull/tempty handshake with 4 drain warps");
  return 0;
}
