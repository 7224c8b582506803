// Microbenchmark: issue rate of tcgen05.mma.cta_group::1.kind::f16 (bf16, M=128) from shared-memory operands,
// one CTA per SM, operands resident (no loads).  Prints cycles per MMA and the implied TFLOP/s for N = 64/128/256.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -I transformertts_b200/csrc tools/mma_peak.cu -o gpurun_out/mma_peak
#include <cstdio>
#include <cuda_runtime.h>
#include "ptx.cuh"
using namespace tts;

__global__ void __launch_bounds__(64, 1) mma_rate(int N, int iters, long long* cycles) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = align_smem_1024(raw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  fence_proxy_async();
  if (threadIdx.x < 32) { tmem_alloc(slot, 512); tmem_relinquish(); }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = *slot;
  if (threadIdx.x == 0) {
    const uint64_t da = umma_desc_sw128(smem_u32(smem)), db = umma_desc_sw128(smem_u32(smem + 16384));
    const uint32_t idesc = umma_idesc_bf16(128, N);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k = 0; k < 4; ++k) tc_mma_bf16(tm + (i & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1u);
    }
    tc_commit(bar);
    mbar_wait(bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) *cycles = t1 - t0;
  }
  __syncthreads();
  if (threadIdx.x < 32) tmem_dealloc(tm, 512);
}

// Same for CTA pairs: tcgen05.mma.cta_group::2, M = 256 over two SMs (each holds 128 rows of A and N/2 rows of B).
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(64, 1) mma_rate_pair(int N, int iters, long long* cycles) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = align_smem_1024(raw);
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem + 16384 + 32768);
  uint32_t* slot = reinterpret_cast<uint32_t*>(bar + 1);
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { mbar_init(bar, 1); fence_mbar_init(); }
  fence_proxy_async();
  if (threadIdx.x < 32) { tmem_alloc_cg2(slot, 512); tmem_relinquish_cg2(); }
  tc_fence_before();
  cluster_sync_all();
  tc_fence_after();
  const uint32_t tm = *slot;
  if (threadIdx.x == 0 && cluster_cta_rank() == 0) {
    const uint64_t da = umma_desc_sw128(smem_u32(smem)), db = umma_desc_sw128(smem_u32(smem + 16384));
    const uint32_t idesc = umma_idesc_bf16(256, N);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
#pragma unroll
      for (int k = 0; k < 4; ++k) tc_mma_bf16_cg2(tm + (i & 1) * 256, da + 2 * k, db + 2 * k, idesc, 1u);
    }
    tc_commit_cg2(bar, 1);
    mbar_wait(bar, 0);
    long long t1 = clock64();
    if (blockIdx.x == 0) *cycles = t1 - t0;
  }
  tc_fence_before();
  cluster_sync_all();
  if (threadIdx.x < 32) tmem_dealloc_cg2(tm, 512);
}

int main() {
  long long* d; cudaMalloc(&d, 8);
  int sms = 0; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  int khz = 0; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
  cudaFuncSetAttribute(mma_rate, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
  const int iters = 4096;
  for (int N : {64, 128, 256}) {
    for (int rep = 0; rep < 2; ++rep) {
      mma_rate<<<sms, 64, 60000>>>(N, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    }
    long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / (iters * 4);
    const double flops = 2.0 * 128 * N * 16 / per * (khz * 1e3) * sms;
    printf("N=%3d: %.1f cycles per MMA (floor 128*N/256 = %d) -> %.0f TFLOP/s at %d SMs x %.3f GHz (clock rate attr)\n",
           N, per, 128 * N / 256, flops / 1e12, sms, khz / 1e6);
  }
  cudaFuncSetAttribute(mma_rate_pair, cudaFuncAttributeMaxDynamicSharedMemorySize, 60000);
  for (int N : {64, 128, 256}) {
    for (int rep = 0; rep < 2; ++rep) {
      mma_rate_pair<<<sms / 2 * 2, 64, 60000>>>(N, iters, d);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
    }
    long long c; cudaMemcpy(&c, d, 8, cudaMemcpyDeviceToHost);
    const double per = (double)c / (iters * 4);
    const double flops = 2.0 * 256 * N * 16 / per * (khz * 1e3) * (sms / 2);
    printf("pair M=256 N=%3d: %.1f cycles per MMA -> %.0f TFLOP/s over %d pairs\n", N, per, flops / 1e12, sms / 2);
  }
  return 0;
}
