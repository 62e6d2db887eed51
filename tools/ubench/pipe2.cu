// Micro-benchmarks of tcgen05.mma issue behaviour on one SM (see pipe.cu for the handshake).
#include <cstdio>
#include <cuda_runtime.h>
#include "../../bayesian-torch_b200/csrc/bt_tc.cuh"
using namespace btc;

// MODE 0: NM MMAs per slot, wait+commit per slot (baseline)   MODE 1: same, try_wait of the NEXT slot issued before the MMAs
// MODE 2: no handshake at all, just MMAs + one final commit    MODE 3: two consumer warps, each with its own producer+ring+accumulator
template <int MODE, int N, int NM>
__global__ void __launch_bounds__(192, 1) k(int iters, long long* out) {
  constexpr int NA = 8;
  extern __shared__ __align__(1024) uint8_t dyn[];
  uint8_t (*tiles)[49152] = reinterpret_cast<uint8_t(*)[49152]>(dyn);
  __shared__ uint64_t full[2][NA], empty[2][NA], done;
  __shared__ uint32_t tmem_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    for (int r = 0; r < 2; ++r) for (int i = 0; i < NA; ++i) { mbar_init(smem_u32(&full[r][i]), 1); mbar_init(smem_u32(&empty[r][i]), 1); }
    mbar_init(smem_u32(&done), MODE == 3 ? 2 : 1);
    mbar_fence_init();
  }
  for (int i = threadIdx.x; i < 2 * 49152 / 4; i += 192) reinterpret_cast<uint32_t*>(dyn)[i] = 0;
  if (warp == 1) tmem_alloc(smem_u32(&tmem_slot), 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = tmem_slot;
  const long long t0 = clock64();
  const int nrings = MODE == 3 ? 2 : 1;
  if (warp == 0 || (warp == 2 && MODE == 3)) {       // producers
    const int r = warp == 0 ? 0 : 1;
    if (lane == 0 && MODE != 2) {
      int slot = 0; uint32_t ph = 0;
      for (int i = 0; i < iters; ++i) {
        mbar_wait(smem_u32(&empty[r][slot]), ph ^ 1u);
        mbar_arrive(smem_u32(&full[r][slot]));
        if (++slot == NA) { slot = 0; ph ^= 1u; }
      }
    }
    __syncwarp();
  } else if ((warp == 1 && MODE != 4) || (warp == 3 && MODE == 3)) {  // consumers
    const int r = warp == 1 ? 0 : 1;
    if (lane == 0) {
      int slot = 0; uint32_t ph = 0;
      constexpr uint32_t idesc = umma_idesc_bf16(128, N);
      const uint64_t ad = umma_desc_sw128(smem_u32(tiles[0])), bd = umma_desc_sw128(smem_u32(tiles[1]));
      const uint32_t d = tmem + r * 256;
      bool ok = MODE == 1 ? mbar_try_wait(smem_u32(&full[r][0]), 0) : false;
      for (int i = 0; i < iters; ++i) {
        if (MODE == 0 || MODE == 3) mbar_wait(smem_u32(&full[r][slot]), ph);
        if (MODE == 1) { if (!ok) mbar_wait(smem_u32(&full[r][slot]), ph); }
        if (MODE != 2) tc_fence_after();
        int nslot = slot + 1; uint32_t nph = ph;
        if (nslot == NA) { nslot = 0; nph ^= 1u; }
        if (MODE == 1) ok = mbar_try_wait(smem_u32(&full[r][nslot]), nph);
#pragma unroll
        for (int kk = 0; kk < NM; ++kk) umma_bf16(d, ad + 2 * (kk & 3) + 1024 * (kk >> 2), bd + 2 * (kk & 3) + 1024 * (kk >> 2), idesc, 1u);
        if (MODE != 2) umma_commit(smem_u32(&empty[r][slot]));
        slot = nslot; ph = nph;
      }
      umma_commit(smem_u32(&done));
      mbar_wait(smem_u32(&done), 0);
    }
    __syncwarp();
  }
  if (MODE == 4 && warp == 4) {   // warp-uniform issue loop: every lane waits, one elected lane issues
    int slot = 0; uint32_t ph = 0;
    constexpr uint32_t idesc = umma_idesc_bf16(128, N);
    const uint64_t ad = umma_desc_sw128(smem_u32(tiles[0])), bd = umma_desc_sw128(smem_u32(tiles[1]));
    for (int i = 0; i < iters; ++i) {
      mbar_wait(smem_u32(&full[0][slot]), ph);
      tc_fence_after();
      uint32_t is_leader;
      asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.b32 %0, 1, 0, p;\n\t}" : "=r"(is_leader));
      if (is_leader) {
#pragma unroll
        for (int kk = 0; kk < NM; ++kk) umma_bf16(tmem, ad + 2 * (kk & 3) + 1024 * (kk >> 2), bd + 2 * (kk & 3) + 1024 * (kk >> 2), idesc, 1u);
        umma_commit(smem_u32(&empty[0][slot]));
      }
      __syncwarp();
      if (++slot == NA) { slot = 0; ph ^= 1u; }
    }
    if (lane == 0) { umma_commit(smem_u32(&done)); mbar_wait(smem_u32(&done), 0); }
    __syncwarp();
  }
  __syncthreads();
  if (threadIdx.x == 0) out[blockIdx.x] = clock64() - t0;
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem, 512);
  (void)nrings;
}

template <int MODE, int N, int NM>
void run(const char* name) {
  long long* d; cudaMalloc(&d, sizeof(long long));
  const int iters = 20000;
  cudaFuncSetAttribute(k<MODE, N, NM>, cudaFuncAttributeMaxDynamicSharedMemorySize, 98304);
  k<MODE, N, NM><<<1, 192, 98304>>>(iters, d);
  k<MODE, N, NM><<<1, 192, 98304>>>(iters, d);
  cudaError_t e = cudaDeviceSynchronize();
  long long h; cudaMemcpy(&h, d, sizeof(long long), cudaMemcpyDeviceToHost);
  const double per_slot = (double)h / iters;
  const double ideal = NM * 128.0 * N / 256.0 * (MODE == 3 ? 2 : 1);
  printf("%-34s N=%3d MMAs/slot=%2d: %7.1f cycles/slot, ideal MMA %6.1f -> %4.1f%% (%s)\n", name, N, NM, per_slot, ideal, 100 * ideal / per_slot, cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run<2, 64, 4>("no handshake");
  run<2, 128, 4>("no handshake");
  run<2, 256, 4>("no handshake");
  run<0, 64, 4>("wait+commit per slot");
  run<0, 128, 4>("wait+commit per slot");
  run<0, 256, 4>("wait+commit per slot");
  run<0, 128, 8>("wait+commit per slot");
  run<0, 256, 8>("wait+commit per slot");
  run<1, 64, 4>("early try_wait of next slot");
  run<1, 128, 4>("early try_wait of next slot");
  run<1, 256, 4>("early try_wait of next slot");
  run<1, 128, 8>("early try_wait of next slot");
  run<4, 64, 4>("warp-uniform issue loop");
  run<4, 128, 4>("warp-uniform issue loop");
  run<4, 128, 8>("warp-uniform issue loop");
  run<4, 256, 4>("warp-uniform issue loop");
  run<3, 64, 4>("two issuing warps");
  run<3, 128, 4>("two issuing warps");
  run<3, 256, 4>("two issuing warps");
  return 0;
}
