// Row-wise kernels of the S3 tokenizer encoder (rows A9/A10): layout change, LayerNorm, FSMN memory, FSQ (RoPE and
// the q/k/v operand split live in the QKV GEMM's epilogue, fy_gemm.cu).  All HBM-bound; activations are fp32 in the residual stream and bf16 (hi | lo) as tensor-core operands.
#pragma once
#include "fy_common.cuh"

namespace fy {
namespace s3k {

__device__ __forceinline__ void split_bf16(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// mel (reference layout [128][T_u] per utterance, fp32) -> time-major bf16 rows at row 4*P_u + 3 + t.
// rows: [Rmel, 128 * W] with hi at [0,128) and (W == 2) lo at [128,256).  Rows outside the utterances are cleared by
// zero_gap_rows_kernel.
// grid: (ceil(Tmax / 32), n_utts), block 256.
template <bool SPLIT>
__global__ void __launch_bounds__(256) mel_to_rows_kernel(const float* __restrict__ mel,
                                                          const int* __restrict__ frame_off,
                                                          const int* __restrict__ utt_P, __nv_bfloat16* __restrict__ rows) {
  __shared__ float tile[128][33];
  const int u = blockIdx.y;
  const int t0 = blockIdx.x * 32;
  const int f0 = frame_off[u];
  const int T = frame_off[u + 1] - f0;
  if (t0 >= T) return;
  const float* src = mel + (size_t)f0 * 128;
  for (int i = threadIdx.x; i < 128 * 32; i += 256) {
    const int b = i >> 5, t = i & 31;
    tile[b][t] = (t0 + t < T) ? src[(size_t)b * T + t0 + t] : 0.f;
  }
  __syncthreads();
  constexpr int LD = SPLIT ? 256 : 128;
  const size_t row0 = (size_t)4 * utt_P[u] + 3 + t0;
  for (int i = threadIdx.x; i < 32 * 128; i += 256) {
    const int t = i >> 7, b = i & 127;
    if (t0 + t < T) {
      __nv_bfloat16 hi, lo;
      split_bf16(tile[b][t], hi, lo);
      rows[(row0 + t) * LD + b] = hi;
      if (SPLIT) rows[(row0 + t) * LD + 128 + b] = lo;
    }
  }
}

// Zero every row of the mel operand buffer that is not an utterance's data: block b clears the rows between the data
// of utterance b - 1 (or the start of the buffer) and the data of utterance b (or the end).  Those rows are the zero
// padding of the stride-2 convolutions; clearing only them replaces a memset of the whole buffer.
// grid: n_utts + 1, block 128; row_bytes is a multiple of 16.
__global__ void __launch_bounds__(128) zero_gap_rows_kernel(__nv_bfloat16* __restrict__ rows, int row_bytes,
                                                            const int* __restrict__ frame_off,
                                                            const int* __restrict__ utt_P, int n_utts, int total_rows) {
  const int b = blockIdx.x;
  const int lo = b == 0 ? 0 : 4 * utt_P[b - 1] + 3 + (frame_off[b] - frame_off[b - 1]);
  const int hi = b == n_utts ? total_rows : 4 * utt_P[b] + 3;
  if (hi <= lo) return;
  uint4* p = reinterpret_cast<uint4*>(reinterpret_cast<char*>(rows) + (size_t)lo * row_bytes);
  const size_t n = (size_t)(hi - lo) * row_bytes / 16;
  for (size_t i = threadIdx.x; i < n; i += 128) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

// LayerNorm over D (multiple of 128) per row: one warp per row.  out bf16 [rows, D * W] (hi | lo).
template <bool SPLIT, int D>
__global__ void __launch_bounds__(256) layernorm_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                                                        const float* __restrict__ beta, float eps, int rows,
                                                        __nv_bfloat16* __restrict__ out) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  constexpr int NV = D / 128;  // float4 per lane
  const float4* xr = reinterpret_cast<const float4*>(x + (size_t)row * D);
  float4 v[NV];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    v[i] = xr[lane + 32 * i];
    s += v[i].x + v[i].y + v[i].z + v[i].w;
  }
  const float mean = warp_sum(s) * (1.0f / D);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
    q += a * a + b * b + c * c + d * d;
  }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / D) + eps);
  constexpr int LD = SPLIT ? 2 * D : D;
  __nv_bfloat16* o = out + (size_t)row * LD;
#pragma unroll
  for (int i = 0; i < NV; ++i) {
    const int c = (lane + 32 * i) * 4;
    const float4 g = __ldg(reinterpret_cast<const float4*>(gamma + c));
    const float4 b = __ldg(reinterpret_cast<const float4*>(beta + c));
    float y[4] = {(v[i].x - mean) * rstd * g.x + b.x, (v[i].y - mean) * rstd * g.y + b.y,
                  (v[i].z - mean) * rstd * g.z + b.z, (v[i].w - mean) * rstd * g.w + b.w};
    __nv_bfloat16 hi[4], lo[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) split_bf16(y[j], hi[j], lo[j]);
    *reinterpret_cast<uint2*>(o + c) = *reinterpret_cast<uint2*>(hi);
    if (SPLIT) *reinterpret_cast<uint2*>(o + D + c) = *reinterpret_cast<uint2*>(lo);
  }
}

// FSMN memory: x[row, c] += v[row, c] + sum_j w[j][c] * v[row + j - pad, c], taps confined to the row's utterance.
// v is the value projection as written by the fused QKV epilogue: bf16 [rows, ldv] with hi at column c and (SPLIT)
// lo at lo_off + c; hi + lo restores it to 2^-17.  One block = one 64-row tile of ONE utterance (the attention tile
// list is reused) x 64 channels: the tile plus its 15-row halos, clipped to the utterance and zero-filled outside, is
// staged in shared memory once, so the tap loop has no utterance tests.  A thread owns 8 consecutive rows x 4
// channels and slides an 8-row register window over the 31 taps (one LDS.128 of v and one of w per 32 FMAs); its
// residual-stream values are prefetched into L2 before the tap loop so the final read-modify-write is short.
// grid: (n_tiles, D / 64), block 128.
constexpr int kFsmnRows = 64, kFsmnCols = 64, kFsmnMaxK = 31, kFsmnPerThread = 8;
template <int KS, bool SPLIT>
__global__ void __launch_bounds__(128, 4) fsmn_kernel(const __nv_bfloat16* __restrict__ v, int ldv, int lo_off,
                                                      const int2* __restrict__ tiles, const int* __restrict__ utt_P,
                                                      const int* __restrict__ utt_L,
                                                      const float* __restrict__ w /* [KS][D] tap-major */, int D,
                                                      float* __restrict__ x) {
  constexpr int kSpan = kFsmnRows + KS - 1;
  constexpr int kPad = (KS - 1) / 2;
  __shared__ float4 sv[kSpan][kFsmnCols / 4];
  __shared__ float4 sw[KS][kFsmnCols / 4];
  const int2 tl = tiles[blockIdx.x];
  const int P = utt_P[tl.x], L = utt_L[tl.x], t0 = tl.y;
  const int c0 = blockIdx.y * kFsmnCols;
  const int c4 = threadIdx.x & 15;
  const int r0 = (threadIdx.x >> 4) * kFsmnPerThread;   // first output row of this thread inside the tile
  // pull this thread's residual-stream values into L2 now; they are read after the tap loop
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) {
    const int t = t0 + r0 + i;
    if (t < L) asm volatile("prefetch.global.L2 [%0];\n" ::"l"(x + (size_t)(P + t) * D + c0 + c4 * 4));
  }
  // stage v: all global loads of this thread are issued before any is consumed (one round trip, not twelve)
  constexpr int kIters = (kSpan * 16 + 127) / 128;
  uint2 gh[kIters], gl[kIters];
#pragma unroll
  for (int it = 0; it < kIters; ++it) {
    const int i = threadIdx.x + it * 128;
    const int rr = i >> 4, cc = i & 15;
    const int t = t0 - kPad + rr;
    gh[it] = make_uint2(0u, 0u);
    gl[it] = make_uint2(0u, 0u);
    if (i < kSpan * 16 && t >= 0 && t < L) {
      const __nv_bfloat16* src = v + (size_t)(P + t) * ldv + c0 + cc * 4;
      gh[it] = __ldg(reinterpret_cast<const uint2*>(src));
      if (SPLIT) gl[it] = __ldg(reinterpret_cast<const uint2*>(src + lo_off));
    }
  }
#pragma unroll
  for (int it = 0; it < kIters; ++it) {
    const int i = threadIdx.x + it * 128;
    if (i < kSpan * 16) {
      float4 val;
      val.x = __uint_as_float(gh[it].x << 16) + __uint_as_float(gl[it].x << 16);
      val.y = __uint_as_float(gh[it].x & 0xffff0000u) + __uint_as_float(gl[it].x & 0xffff0000u);
      val.z = __uint_as_float(gh[it].y << 16) + __uint_as_float(gl[it].y << 16);
      val.w = __uint_as_float(gh[it].y & 0xffff0000u) + __uint_as_float(gl[it].y & 0xffff0000u);
      sv[i >> 4][i & 15] = val;
    }
  }
  for (int i = threadIdx.x; i < KS * 16; i += 128)
    sw[i >> 4][i & 15] = __ldg(reinterpret_cast<const float4*>(w + (size_t)(i >> 4) * D + c0 + (i & 15) * 4));
  __syncthreads();
  float4 acc[kFsmnPerThread];
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) acc[i] = sv[r0 + i + kPad][c4];   // identity term
  float4 win[kFsmnPerThread];                                                // staged rows r0 + j .. r0 + j + 7
#pragma unroll
  for (int i = 0; i < kFsmnPerThread - 1; ++i) win[i] = sv[r0 + i][c4];
#pragma unroll
  for (int j = 0; j < KS; ++j) {
    win[(j + kFsmnPerThread - 1) % kFsmnPerThread] = sv[r0 + j + kFsmnPerThread - 1][c4];
    const float4 ww = sw[j][c4];
#pragma unroll
    for (int i = 0; i < kFsmnPerThread; ++i) {
      const float4 vv = win[(i + j) % kFsmnPerThread];
      acc[i].x = fmaf(ww.x, vv.x, acc[i].x); acc[i].y = fmaf(ww.y, vv.y, acc[i].y);
      acc[i].z = fmaf(ww.z, vv.z, acc[i].z); acc[i].w = fmaf(ww.w, vv.w, acc[i].w);
    }
  }
  float4 xv[kFsmnPerThread];
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) {
    const int t = t0 + r0 + i;
    xv[i] = t < L ? *reinterpret_cast<const float4*>(x + (size_t)(P + t) * D + c0 + c4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) {
    const int t = t0 + r0 + i;
    if (t < L) {
      float4 o = xv[i];
      o.x += acc[i].x; o.y += acc[i].y; o.z += acc[i].z; o.w += acc[i].w;
      *reinterpret_cast<float4*>(x + (size_t)(P + t) * D + c0 + c4 * 4) = o;
    }
  }
}

// FSQ: z = W[8, D] x + b -> tanh -> * scale -> pre; id = sum (rint(pre) + 1) * 3^i.  One warp per token row;
// output index = tok_off[u] + pos.
__global__ void __launch_bounds__(256) fsq_kernel(const float* __restrict__ x, const int* __restrict__ row_utt,
                                                  const int* __restrict__ row_pos, const int* __restrict__ tok_off,
                                                  const float* __restrict__ w, const float* __restrict__ b, int rows,
                                                  int D, float fsq_scale, int* __restrict__ tokens,
                                                  float* __restrict__ pre_out) {
  const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int u = __ldg(row_utt + row);
  if (u < 0) return;
  float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  const float* xr = x + (size_t)row * D;
  for (int c = lane * 4; c < D; c += 128) {
    const float4 xv = *reinterpret_cast<const float4*>(xr + c);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float4 wv = __ldg(reinterpret_cast<const float4*>(w + (size_t)i * D + c));
      acc[i] += xv.x * wv.x + xv.y * wv.y + xv.z * wv.z + xv.w * wv.w;
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) acc[i] = warp_sum(acc[i]);
  if (lane == 0) {
    const size_t o = (size_t)__ldg(tok_off + u) + __ldg(row_pos + row);
    int id = 0, p3 = 1;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float pre = tanhf(acc[i] + __ldg(b + i)) * fsq_scale;
      if (pre_out) pre_out[o * 8 + i] = pre;
      id += ((int)rintf(pre) + 1) * p3;
      p3 *= 3;
    }
    tokens[o] = id;
  }
}

}  // namespace s3k
}  // namespace fy
