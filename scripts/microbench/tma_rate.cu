// tma_rate.cu -- microbenchmark: how fast can every SM ingest fp16 tiles through TMA?
// One CTA per SM (clusters of 2), one producer thread keeps a ring of 32 KB stages full, one
// consumer thread frees them.  Modes: unicast 64-row boxes, or 2-CTA multicast of 32-row boxes.
// Source: [rows, 768] fp16, either L2-sized or HBM-sized.  Prints aggregate TB/s into shared memory.
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "../../sisap23-laion-challenge-learned-index_b200/csrc/ptx_sm100.cuh"
using namespace lmi;

constexpr int STAGE_BYTES = 32768, ATOM = 8192;
#ifndef STAGES
#define STAGES 6
#endif

template <int MC>   // 0 unicast, 1 multicast half-boxes (32 rows), 2 multicast alternating whole atoms (64 rows)
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64, 1)
tma_kernel(const __grid_constant__ CUtensorMap t64, const __grid_constant__ CUtensorMap t32, int tiles_per_cta,
           int rows_total, int same_chunk, long long* cycles) {
  extern __shared__ unsigned char raw[];
  unsigned char* smem = raw + ((1024u - (ptx::smem_u32(raw) & 1023u)) & 1023u);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + STAGES * STAGE_BYTES);
  uint64_t* empty = full + STAGES;
  const uint32_t crank = ptx::cluster_ctarank();
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { ptx::mbar_init(ptx::smem_u32(&full[s]), 1); ptx::mbar_init(ptx::smem_u32(&empty[s]), MC ? 2 : 1); }
    ptx::fence_barrier_init();
  }
  __syncthreads();
  ptx::cluster_sync();
  const int pair = blockIdx.x >> 1;
  // MC: both CTAs of a pair stream the same rows; unicast: every CTA its own rows (or the pair's, same_chunk)
  // same_chunk = g >= 1: groups of g consecutive CTAs stream the same rows (g = 0: private rows)
  const long long row_base = (long long)(MC ? pair : (same_chunk ? (int)(blockIdx.x / same_chunk) : (int)blockIdx.x)) * tiles_per_cta * 64;
  long long t0 = clock64();
  if (threadIdx.x == 0) {
    int stage = 0; uint32_t phase = 0;
    for (int t = 0; t < tiles_per_cta; ++t) {
      const int obj0 = (int)((row_base + (long long)t * 64) % (rows_total - 64));
      for (int ks = 0; ks < 3; ++ks) {
        ptx::mbar_wait(ptx::smem_u32(&empty[stage]), phase ^ 1);
        const uint32_t fb = ptx::smem_u32(&full[stage]);
        ptx::mbar_arrive_expect_tx(fb, STAGE_BYTES);
        const uint32_t sb = ptx::smem_u32(smem + stage * STAGE_BYTES);
        for (int a = 0; a < 4; ++a) {
          if (MC == 1) ptx::tma_load_2d_mc(sb + a * ATOM + crank * (ATOM / 2), &t32, fb, (ks * 4 + a) * 64, obj0 + (int)crank * 32, (uint16_t)3);
          else if (MC == 2) { if ((a & 1) == (int)crank) ptx::tma_load_2d_mc(sb + a * ATOM, &t64, fb, (ks * 4 + a) * 64, obj0, (uint16_t)3); }
          else    ptx::tma_load_2d(sb + a * ATOM, &t64, fb, (ks * 4 + a) * 64, obj0);
        }
        if (++stage == STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else if (threadIdx.x == 32) {
    int stage = 0; uint32_t phase = 0;
    for (int t = 0; t < tiles_per_cta * 3; ++t) {
      ptx::mbar_wait(ptx::smem_u32(&full[stage]), phase);
      if (MC) { ptx::mbar_arrive(ptx::smem_u32(&empty[stage])); ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&empty[stage]), crank ^ 1)); }
      else ptx::mbar_arrive(ptx::smem_u32(&empty[stage]));
      if (++stage == STAGES) { stage = 0; phase ^= 1; }
    }
  }
  __syncthreads();
  long long t1 = clock64();
  ptx::cluster_sync();
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static void make_map(CUtensorMap* m, void* base, uint64_t rows, int box_rows) {
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
  cuuint64_t gdim[2] = {768, rows}; cuuint64_t gstr[1] = {1536}; cuuint32_t box[2] = {64, (cuuint32_t)box_rows}; cuuint32_t es[2] = {1, 1};
  CUresult r = ((EncodeTiledFn)p)(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, base, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) printf("encode failed %d\n", (int)r);
}

template <int MC>
void run(const char* name, void* buf, int rows_total, int tiles, int same_chunk) {
  CUtensorMap t64, t32; make_map(&t64, buf, rows_total, 64); make_map(&t32, buf, rows_total, 32);
  long long* d; cudaMalloc(&d, 148 * 8);
  const size_t smem = 1024 + STAGES * STAGE_BYTES + 256;
  cudaFuncSetAttribute(tma_kernel<MC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  tma_kernel<MC><<<148, 64, smem>>>(t64, t32, tiles, rows_total, same_chunk, d);
  cudaEventRecord(e0);
  tma_kernel<MC><<<148, 64, smem>>>(t64, t32, tiles, rows_total, same_chunk, d);
  cudaEventRecord(e1);
  cudaError_t e = cudaDeviceSynchronize();
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double bytes = 148.0 * tiles * 64 * 1536;
  printf("%-44s rows=%9d  %7.2f TB/s into smem  (%.3f ms)  %.1f B/clk/SM @1.965GHz %s\n", name, rows_total, bytes / ms / 1e9, ms,
         bytes / 148 / (ms * 1e-3 * 1.965e9), e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  size_t big = (size_t)6000000 * 1536;
  void* buf; cudaMalloc(&buf, big); cudaMemset(buf, 0, big);
  const int tiles = 2048;
  for (int rows : {32768, 6000000}) {   // 50 MB (L2-resident) / 9.2 GB (HBM)
    run<0>("unicast 64-row boxes, private rows", buf, rows, tiles, 0);
    run<0>("unicast 64-row boxes, 2 CTAs share rows", buf, rows, tiles, 2);
    run<0>("unicast 64-row boxes, 3 CTAs share rows", buf, rows, tiles, 3);
    run<0>("unicast 64-row boxes, 4 CTAs share rows", buf, rows, tiles, 4);
    run<1>("multicast 2 x 32-row boxes (pair shares)", buf, rows, tiles, 1);
  }
  return 0;
}
