// Micro-benchmark of the producer/consumer handshake used by bt_tc2.cu: how many cycles one
// ring slot costs with (a) plain mbarrier arrives, (b) tcgen05.commit arrives, (c) + 4 UMMAs.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../bayesian-torch_b200/csrc/bt_tc.cuh"
using namespace btc;

template <int MODE, int NA>
__global__ void __launch_bounds__(128, 1) pipe_kernel(int iters, long long* out) {
  __shared__ __align__(1024) uint8_t tiles[2][16384];
  __shared__ uint64_t full[NA], empty[NA];
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int i = 0; i < NA; ++i) { mbar_init(smem_u32(&full[i]), 1); mbar_init(smem_u32(&empty[i]), 1); }
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 2 * 16384 / 4; i += 128) reinterpret_cast<uint32_t*>(tiles)[i] = 0;
  if (warp == 1) tmem_alloc(smem_u32(&tmem_slot), 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const long long t0 = clock64();
  if (warp == 0) {
    if (lane == 0) {
      int slot = 0; uint32_t ph = 0;
      for (int i = 0; i < iters; ++i) {
        mbar_wait(smem_u32(&empty[slot]), ph ^ 1u);
        mbar_arrive(smem_u32(&full[slot]));
        if (++slot == NA) { slot = 0; ph ^= 1u; }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      int slot = 0; uint32_t ph = 0;
      constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
      const uint64_t ad = umma_desc_sw128(smem_u32(tiles[0])), bd = umma_desc_sw128(smem_u32(tiles[1]));
      for (int i = 0; i < iters; ++i) {
        mbar_wait(smem_u32(&full[slot]), ph);
        if (MODE == 0) {
          mbar_arrive(smem_u32(&empty[slot]));
        } else {
          tc_fence_after();
          if (MODE == 2) {
#pragma unroll
            for (int k = 0; k < 4; ++k) umma_bf16(tmem, ad + 2 * k, bd + 2 * k, idesc, 1u);
          }
          umma_commit(smem_u32(&empty[slot]));
        }
        if (++slot == NA) { slot = 0; ph ^= 1u; }
      }
      // drain: wait for the last commit
      if (MODE != 0) {
        const int lslot = (iters - 1) % NA; const uint32_t lph = ((iters - 1) / NA) & 1;
        mbar_wait(smem_u32(&empty[lslot]), lph);
      }
    }
    __syncwarp();
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
}

template <int MODE, int NA>
void run(const char* name, int grid) {
  long long* d; cudaMalloc(&d, grid * sizeof(long long));
  const int iters = 20000;
  pipe_kernel<MODE, NA><<<grid, 128>>>(iters, d);
  pipe_kernel<MODE, NA><<<grid, 128>>>(iters, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[148]; cudaMemcpy(h, d, grid * sizeof(long long), cudaMemcpyDeviceToHost);
  printf("%-28s NA=%d grid=%3d: %7.1f cycles/slot  (%s)\n", name, NA, grid, (double)h[0] / iters, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<0, 8>("mbarrier.arrive ping-pong", 1);
  run<1, 8>("tcgen05.commit ping-pong", 1);
  run<2, 8>("4 UMMA(128x128x16) + commit", 1);
  run<2, 4>("4 UMMA(128x128x16) + commit", 1);
  run<2, 2>("4 UMMA(128x128x16) + commit", 1);
  run<0, 8>("mbarrier.arrive ping-pong", 148);
  run<1, 8>("tcgen05.commit ping-pong", 148);
  run<2, 8>("4 UMMA(128x128x16) + commit", 148);
  return 0;
}
