// umma_rate.cu -- microbenchmark: tensor-pipe cycles per tcgen05.mma (kind::f16, M=128, K=16) as a
// function of N and of where A comes from (TMEM vs shared memory).  One CTA per SM, one issuing
// thread, R back-to-back MMAs into one accumulator, one commit, wait.  Prints cycles/MMA.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o umma_rate umma_rate.cu && ./umma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../../sisap23-laion-challenge-learned-index_b200/csrc/ptx_sm100.cuh"
using namespace lmi;

template <int N, bool TS, int CTAS>
__global__ void __launch_bounds__(128, 1) rate_kernel(int reps, long long* out) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = raw + ((1024u - (ptx::smem_u32(raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_ptr;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;  // 1.0h
  if (threadIdx.x == 0) { ptx::mbar_init(ptx::smem_u32(&bar), 1); ptx::fence_barrier_init(); }
  ptx::fence_proxy_async();
  if (warp == 0) ptx::tmem_alloc<512>(ptx::smem_u32(&tmem_ptr));
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tb = tmem_ptr;
  if (warp == 0) {
    long long t0 = 0, t1 = 0;
    constexpr uint32_t idesc = ptx::umma_idesc_f16(128, N);
    const uint64_t a_desc = ptx::umma_desc_k_sw128(ptx::smem_u32(smem));             // 128 x 64 halves
    const uint64_t b_desc = ptx::umma_desc_k_sw128(ptx::smem_u32(smem + 16384));     // up to 256 x 64 halves
    if (ptx::elect_one()) {
      t0 = clock64();
      for (int r = 0; r < reps; r += 4) {
#pragma unroll
        for (int k4 = 0; k4 < 4; ++k4) {
          if (TS) ptx::umma_f16_ts_c<true>(tb + 256, tb + (uint32_t)(k4 * 8), b_desc + (uint64_t)(k4 * 2), idesc);
          else    ptx::umma_f16_ss(tb + 256, a_desc + (uint64_t)(k4 * 2), b_desc + (uint64_t)(k4 * 2), idesc, 1u);
        }
      }
      ptx::umma_commit(ptx::smem_u32(&bar));
    }
    __syncwarp();
    ptx::mbar_wait(ptx::smem_u32(&bar), 0);
    t1 = clock64();
    long long dt = t1 - __shfl_sync(0xffffffffu, t0, __ffs(__ballot_sync(0xffffffffu, t0 != 0)) - 1);
    if (lane == 0) out[blockIdx.x] = dt;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tb); }
}

template <int N, bool TS>
void run(const char* name, int grid) {
  long long* d; cudaMalloc(&d, grid * sizeof(long long));
  const int reps = 8192;
  const size_t smem = 1024 + 64 * 1024;
  cudaFuncSetAttribute(rate_kernel<N, TS, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int w = 0; w < 2; ++w) rate_kernel<N, TS, 1><<<grid, 128, smem>>>(reps, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  long long mx = 0; for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
  printf("%-28s grid=%3d  cycles/MMA = %7.1f   (ideal %5.1f)  %s\n", name, grid, (double)mx / reps, 128.0 * N / 256.0,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  for (int grid : {1, 148}) {
    run<64, true>("TS  M128 N64  K16", grid);
    run<128, true>("TS  M128 N128 K16", grid);
    run<256, true>("TS  M128 N256 K16", grid);
    run<64, false>("SS  M128 N64  K16", grid);
    run<128, false>("SS  M128 N128 K16", grid);
    run<256, false>("SS  M128 N256 K16", grid);
  }
  return 0;
}
