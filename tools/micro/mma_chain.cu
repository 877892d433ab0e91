// Micro-benchmark: cycles per tcgen05.mma (M 128, K 16, bf16) as a function of N, of how many independent TMEM
// accumulators the instruction stream rotates over, and of where the A operand lives (shared memory / tensor memory).
// One CTA, one issuing thread; 384 instructions per measurement, timed with clock64 around issue + commit + wait.
//   nvcc -gencode arch=compute_100a,code=sm_100a -I fangyan_tts_b200/csrc -o tools/micro/mma_chain tools/micro/mma_chain.cu
#include <cstdio>
#include <cstdint>
#include <cuda.h>
#include <cuda_runtime.h>
#include "fy_ptx.cuh"
using namespace fy;

__host__ __device__ constexpr uint32_t idesc(uint32_t m, uint32_t n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

template <int N, int NACC, bool TS>
__global__ void __launch_bounds__(128, 1) chain_kernel(long long* out, int slot) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t holder;
  for (int i = threadIdx.x; i < (16384 + 32768) / 4; i += 128) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
  if (threadIdx.x < 32) ptx::tmem_alloc<512>(&holder);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = holder;
  if (threadIdx.x == 0) {
    const uint64_t da = ptx::make_desc_sw128(ptx::smem_u32(smem));
    const uint64_t db = ptx::make_desc_sw128(ptx::smem_u32(smem + 16384));
    constexpr uint32_t id = idesc(128, N);
    constexpr int kIters = 384;
    // accumulators at columns a * N (NACC * N <= 448), TS-mode A operand at columns [448, 480)
    for (int rep = 0; rep < 2; ++rep) {
      const long long t0 = clock64();
#pragma unroll 8
      for (int i = 0; i < kIters; ++i) {
        const uint32_t d = tmem + (i % NACC) * N;
        const int k = i & 3;
        if (TS) ptx::mma_bf16_ts(d, tmem + 448 + 8 * k, db + 2 * k, id, i >= NACC);
        else ptx::mma_bf16_ss(d, da + 2 * k, db + 2 * k, id, i >= NACC);
      }
      ptx::mma_commit(&bar);
      ptx::mbar_wait(&bar, rep & 1);
      const long long t1 = clock64();
      if (rep == 1) out[slot] = (t1 - t0) / kIters;
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem); }
}

// Same instruction stream issued two ways, optionally with 16 warps of dependent FP32 work sharing the schedulers:
//   CONV = false: "if (threadIdx.x == issuer)" -- the compiler wraps every uniform-datapath UTCHMMA of a divergent
//                 thread in an ELECT / BRA.U.ANY loop;
//   CONV = true : the whole warp runs the loop, the instruction itself is predicated on elect.sync.
__device__ __forceinline__ void mma_ts_elect(uint32_t d, uint32_t a, uint64_t db, uint32_t id, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}\n" ::"r"(d),
      "r"(a), "l"(db), "r"(id), "r"(acc)
      : "memory");
}
// BMN = 1: B is MN-major (the P V product of the attention kernels: V tiles as [key][d] rows), descriptor step 2048 B.
// BUSY = 2: the other warps run tcgen05.ld / tcgen05.st round trips on their TMEM lane quarters instead of FP32 work.
template <bool CONV, int BUSY, int BMN = 0>
__global__ void __launch_bounds__(32 + 32 * 16, 1) issue_kernel(long long* out, int slot, float* sink) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);
  __shared__ uint64_t bar;
  __shared__ uint32_t holder;
  __shared__ volatile int done;
  for (int i = threadIdx.x; i < 32768 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); done = 0; }
  if (threadIdx.x < 32) ptx::tmem_alloc<512>(&holder);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = holder;
  constexpr uint32_t id = idesc(128, 64) | (BMN ? (1u << 16) : 0u);
  constexpr int kIters = 384;
  constexpr int kStep = BMN ? 128 : 2;
  if (threadIdx.x < 32) {
    const uint64_t db = ptx::make_desc_sw128(ptx::smem_u32(smem));
    if (CONV) {
      const long long t0 = clock64();
#pragma unroll 12
      for (int i = 0; i < kIters; ++i) mma_ts_elect(tmem + (i / 12 % 3) * 64, tmem + 448 + 8 * (i & 3), db + kStep * (i & 3), id, i % 12 != 0);
      if (ptx::elect_one()) ptx::mma_commit(&bar);
      ptx::mbar_wait(&bar, 0);
      if (threadIdx.x == 0) out[slot] = (clock64() - t0) / kIters;
    } else if (threadIdx.x == 0) {
      const long long t0 = clock64();
#pragma unroll 12
      for (int i = 0; i < kIters; ++i) ptx::mma_bf16_ts(tmem + (i / 12 % 3) * 64, tmem + 448 + 8 * (i & 3), db + 2 * (i & 3), id, i % 12 != 0);
      ptx::mma_commit(&bar);
      ptx::mbar_wait(&bar, 0);
      out[slot] = (clock64() - t0) / kIters;
    }
    __syncwarp();
    if (threadIdx.x == 0) done = 1;
  } else if (BUSY == 2) {
    const uint32_t t_row = tmem + ((uint32_t)(((threadIdx.x >> 5) & 3) * 32) << 16) + 256 + 16 * ((threadIdx.x >> 7) & 3);
    uint32_t v[16];
    while (!done) {
      ptx::tmem_ld_32x16(t_row, v);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] += 1;
      ptx::tmem_st_32x16(t_row, v);
      ptx::tmem_st_wait();
    }
    if (v[0] == 0x12345678u) *sink = 1.f;
  } else if (BUSY) {
    float a = threadIdx.x, b = 1.0001f, c = 0.5f, d = 0.25f;
    while (!done) {
#pragma unroll
      for (int i = 0; i < 64; ++i) { a = fmaf(a, b, c); d = fmaf(d, b, a); c = fmaf(c, b, d); }
    }
    if (a + c + d == 12345.f) *sink = a;
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc<512>(tmem); }
}
template <bool CONV, int BUSY, int BMN = 0>
void run_issue(long long* d_out, int slot, float* sink) {
  cudaFuncSetAttribute(issue_kernel<CONV, BUSY, BMN>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  issue_kernel<CONV, BUSY, BMN><<<1, 32 + 32 * 16, 64 * 1024>>>(d_out, slot, sink);
}

template <int N, int NACC, bool TS>
void run(long long* d_out, int slot) {
  cudaFuncSetAttribute(chain_kernel<N, NACC, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
  chain_kernel<N, NACC, TS><<<1, 128, 64 * 1024>>>(d_out, slot);
}

int main() {
  long long* d_out;
  cudaMalloc(&d_out, 64 * sizeof(long long));
  cudaMemset(d_out, 0, 64 * sizeof(long long));
  int s = 0;
  run<64, 1, false>(d_out, s++); run<64, 2, false>(d_out, s++); run<64, 4, false>(d_out, s++);
  run<64, 1, true>(d_out, s++);  run<64, 2, true>(d_out, s++);  run<64, 4, true>(d_out, s++);
  run<128, 1, false>(d_out, s++); run<128, 2, false>(d_out, s++); run<128, 1, true>(d_out, s++); run<128, 2, true>(d_out, s++);
  run<256, 1, false>(d_out, s++); run<256, 1, true>(d_out, s++);
  run<32, 1, true>(d_out, s++); run<32, 4, true>(d_out, s++);
  const int s_issue = s;
  float* sink;
  cudaMalloc(&sink, 4);
  run_issue<false, 0>(d_out, s++, sink); run_issue<false, 1>(d_out, s++, sink);
  run_issue<true, 0>(d_out, s++, sink); run_issue<true, 1>(d_out, s++, sink);
  run_issue<true, 0, 1>(d_out, s++, sink); run_issue<true, 1, 1>(d_out, s++, sink);
  run_issue<true, 2, 0>(d_out, s++, sink); run_issue<true, 2, 1>(d_out, s++, sink);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("error %s\n", cudaGetErrorString(e)); return 1; }
  long long h[64];
  cudaMemcpy(h, d_out, sizeof(h), cudaMemcpyDeviceToHost);
  const char* names[] = {"N64 acc1 SS", "N64 acc2 SS", "N64 acc4 SS", "N64 acc1 TS", "N64 acc2 TS", "N64 acc4 TS",
                         "N128 acc1 SS", "N128 acc2 SS", "N128 acc1 TS", "N128 acc2 TS", "N256 acc1 SS", "N256 acc1 TS",
                         "N32 acc1 TS", "N32 acc4 TS"};
  const char* inames[] = {"lane-0 branch, idle SM", "lane-0 branch, 16 busy warps", "converged + elect, idle SM",
                          "converged + elect, 16 busy warps", "converged, B MN-major, idle SM",
                          "converged, B MN-major, 16 busy warps", "converged, 16 warps of TMEM ld/st",
                          "converged, B MN-major, TMEM ld/st"};
  for (int i = s_issue; i < s; ++i) printf("N64 TS, 12-long chains, %-34s %lld clk / mma\n", inames[i - s_issue], h[i]);
  s = s_issue;
  for (int i = 0; i < s; ++i) printf("%-14s %lld clk / mma (floor %d)\n", names[i], h[i], 128 * atoi(names[i] + 1) / 256);
  return 0;
}
