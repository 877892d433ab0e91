// Row-wise kernels of the S3 tokenizer encoder (rows A9/A10): layout change, LayerNorm, RoPE + split, FSMN memory,
// FSQ.  All HBM-bound; activations are fp32 in the residual stream and bf16 (hi | lo) as tensor-core operands.
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
// rows: [Rmel, 128 * W] with hi at [0,128) and (W == 2) lo at [128,256).  Buffer is zeroed beforehand.
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

// qkv fp32 [rows, 3D] -> q, k (RoPE, rotate-half, then * scale) and v as bf16 [rows, D * W] (hi | lo).
// One thread per (row, head, d < DH/2) handles the pair (d, d + DH/2) of q and k and two v values.
template <bool SPLIT>
__global__ void __launch_bounds__(256) rope_split_kernel(const float* __restrict__ qkv, const int* __restrict__ row_pos,
                                                         const float* __restrict__ cosT, const float* __restrict__ sinT,
                                                         int rows, int D, int DH, float scale,
                                                         __nv_bfloat16* __restrict__ q, __nv_bfloat16* __restrict__ k,
                                                         __nv_bfloat16* __restrict__ v) {
  const int half = DH / 2;
  const int per_row = D / 2;
  const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (size_t)rows * per_row) return;
  const int row = (int)(idx / per_row);
  const int r = (int)(idx % per_row);
  const int h = r / half, d = r % half;
  const int c0 = h * DH + d, c1 = c0 + half;
  int pos = __ldg(row_pos + row);
  if (pos < 0) pos = 0;
  const float cs0 = __ldg(cosT + (size_t)pos * DH + d), sn0 = __ldg(sinT + (size_t)pos * DH + d);
  const float cs1 = __ldg(cosT + (size_t)pos * DH + d + half), sn1 = __ldg(sinT + (size_t)pos * DH + d + half);
  const float* src = qkv + (size_t)row * 3 * D;
  const int LD = SPLIT ? 2 * D : D;
  const size_t o = (size_t)row * LD;
  __nv_bfloat16 hi, lo;
  {
    const float a = src[c0], b = src[c1];
    split_bf16((a * cs0 - b * sn0) * scale, hi, lo);
    q[o + c0] = hi; if (SPLIT) q[o + D + c0] = lo;
    split_bf16((b * cs1 + a * sn1) * scale, hi, lo);
    q[o + c1] = hi; if (SPLIT) q[o + D + c1] = lo;
  }
  {
    const float a = src[D + c0], b = src[D + c1];
    split_bf16((a * cs0 - b * sn0) * scale, hi, lo);
    k[o + c0] = hi; if (SPLIT) k[o + D + c0] = lo;
    split_bf16((b * cs1 + a * sn1) * scale, hi, lo);
    k[o + c1] = hi; if (SPLIT) k[o + D + c1] = lo;
  }
  {
    split_bf16(src[2 * D + c0], hi, lo);
    v[o + c0] = hi; if (SPLIT) v[o + D + c0] = lo;
    split_bf16(src[2 * D + c1], hi, lo);
    v[o + c1] = hi; if (SPLIT) v[o + D + c1] = lo;
  }
}

// FSMN memory: x[row, c] += v[row, c] + sum_j w[j][c] * v[row + j - pad, c] over rows of the same utterance.
// v is the fp32 value projection inside qkv (column offset 2D).  A block stages a (64 + KS - 1) x 64 tile of v in
// shared memory (each value leaves L2 once); a thread owns 8 consecutive rows x 4 channels and slides an
// 8-row register window over the taps, so shared memory is read ~3 times per tap instead of 2 per tap per output.
// grid: (ceil(rows / 64), D / 64), block 128.
constexpr int kFsmnRows = 64, kFsmnCols = 64, kFsmnMaxK = 31, kFsmnPerThread = 8;
template <int KS>
__global__ void __launch_bounds__(128) fsmn_kernel(const float* __restrict__ qkv, const int* __restrict__ row_utt,
                                                   const float* __restrict__ w /* [KS][D] tap-major */, int rows, int D,
                                                   float* __restrict__ x) {
  constexpr int kSpan = kFsmnRows + KS - 1;
  constexpr int kPad = (KS - 1) / 2;
  __shared__ float4 sv[kSpan][kFsmnCols / 4];
  __shared__ float4 sw[KS][kFsmnCols / 4];
  __shared__ int sutt[kSpan];
  const int r0 = blockIdx.x * kFsmnRows, c0 = blockIdx.y * kFsmnCols;
  const size_t ldv = (size_t)3 * D;
  for (int i = threadIdx.x; i < kSpan * 16; i += 128) {
    const int rr = i >> 4, c4 = i & 15;
    const int row = r0 - kPad + rr;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (row >= 0 && row < rows) v = *reinterpret_cast<const float4*>(qkv + (size_t)row * ldv + 2 * D + c0 + c4 * 4);
    sv[rr][c4] = v;
    if (c4 == 0) sutt[rr] = (row >= 0 && row < rows) ? __ldg(row_utt + row) : -1;
  }
  for (int i = threadIdx.x; i < KS * 16; i += 128)
    sw[i >> 4][i & 15] = __ldg(reinterpret_cast<const float4*>(w + (size_t)(i >> 4) * D + c0 + (i & 15) * 4));
  __syncthreads();
  const int c4 = threadIdx.x & 15;
  const int t0 = (threadIdx.x >> 4) * kFsmnPerThread;   // first output row of this thread inside the tile
  float4 acc[kFsmnPerThread];
  int uo[kFsmnPerThread];
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) {
    acc[i] = sv[t0 + i + kPad][c4];                     // identity term
    uo[i] = sutt[t0 + i + kPad];
  }
  float4 win[kFsmnPerThread];                            // rows t0 + j .. t0 + j + 7 of the staged tile
  int wu[kFsmnPerThread];
#pragma unroll
  for (int i = 0; i < kFsmnPerThread - 1; ++i) { win[i] = sv[t0 + i][c4]; wu[i] = sutt[t0 + i]; }
#pragma unroll
  for (int j = 0; j < KS; ++j) {
    win[(j + kFsmnPerThread - 1) % kFsmnPerThread] = sv[t0 + j + kFsmnPerThread - 1][c4];
    wu[(j + kFsmnPerThread - 1) % kFsmnPerThread] = sutt[t0 + j + kFsmnPerThread - 1];
    const float4 ww = sw[j][c4];
#pragma unroll
    for (int i = 0; i < kFsmnPerThread; ++i) {
      const float4 vv = win[(i + j) % kFsmnPerThread];
      if (wu[(i + j) % kFsmnPerThread] == uo[i]) {
        acc[i].x += ww.x * vv.x; acc[i].y += ww.y * vv.y; acc[i].z += ww.z * vv.z; acc[i].w += ww.w * vv.w;
      }
    }
  }
#pragma unroll
  for (int i = 0; i < kFsmnPerThread; ++i) {
    const int row = r0 + t0 + i;
    if (row < rows && uo[i] >= 0) {
      float4* xp = reinterpret_cast<float4*>(x + (size_t)row * D + c0 + c4 * 4);
      float4 xv = *xp;
      xv.x += acc[i].x; xv.y += acc[i].y; xv.z += acc[i].z; xv.w += acc[i].w;
      *xp = xv;
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
