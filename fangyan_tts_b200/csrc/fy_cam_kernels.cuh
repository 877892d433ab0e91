// CUDA-core kernels of the CAMPPlus speaker network (row A6): the 1->32 stem convolution, BN+ReLU operand
// preparation, the context-aware mask (global + 100-frame segment means -> 2-layer MLP -> sigmoid),
// statistics pooling and the final dense layer.  Convolutions with >= 32 input channels run on the tcgen05 GEMM.
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace camk {

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Stem: Conv2d(1, 32, 3x3, pad 1) + BN + ReLU over (freq, time).  feat is packed [frames, 80] fp32.
// out: [R100, 82, 32] fp32 (rounded to tf32: it is the A operand of the first FCM GEMM) channels-last with zero
// columns at p = 0 and p = 81 (buffer pre-zeroed).
// rowsrc[r] = feat row of time row r or -1 for guard rows.  w: [32][3 (freq)][3 (time)].
// grid = R100, block = 320 (80 freq x 4 groups of 8 channels).
__global__ void __launch_bounds__(320) stem_kernel(const float* __restrict__ feat, const int* __restrict__ rowsrc,
                                                   int R100, const float* __restrict__ w,
                                                   const float* __restrict__ scale, const float* __restrict__ shift,
                                                   float* __restrict__ out) {
  const int r = blockIdx.x;
  const int src = rowsrc[r];
  if (src < 0) return;
  __shared__ float in[3][82];
  __shared__ float sw[32 * 9 + 64];
  for (int i = threadIdx.x; i < 32 * 9; i += 320) sw[i] = w[i];
  for (int i = threadIdx.x; i < 64; i += 320) sw[288 + i] = i < 32 ? scale[i] : shift[i - 32];
  for (int i = threadIdx.x; i < 3 * 82; i += 320) {
    const int j = i / 82, p = i % 82;
    const int rr = r + j - 1;
    const int s = (rr >= 0 && rr < R100) ? rowsrc[rr] : -1;
    in[j][p] = (s >= 0 && p >= 1 && p <= 80) ? feat[(size_t)s * 80 + p - 1] : 0.f;
  }
  __syncthreads();
  const int f = threadIdx.x % 80, cg = threadIdx.x / 80;
  float x[3][3];  // [freq i][time j]
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j < 3; ++j) x[i][j] = in[j][f + i];
  float y[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int c = cg * 8 + k;
    float a = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) a += sw[c * 9 + i * 3 + j] * x[i][j];
    y[k] = ptx::round_tf32(fmaxf(a * sw[288 + c] + sw[320 + c], 0.f));
  }
  float4* dst = reinterpret_cast<float4*>(out + ((size_t)r * 82 + f + 1) * 32 + cg * 8);
  dst[0] = make_float4(y[0], y[1], y[2], y[3]);
  dst[1] = make_float4(y[4], y[5], y[6], y[7]);
}

// mask[g] = -2 (skip) when slot g % Q is outside [lo, hi]; else -1 (write zero) on guard time rows; else 0
__global__ void pos_mask_kernel(const int* __restrict__ rowsrc, int R, int Q, int lo, int hi, int* __restrict__ mask) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * Q) return;
  const int p = g % Q;
  mask[g] = (p < lo || p > hi) ? -2 : (rowsrc[g / Q] >= 0 ? 0 : -1);
}

// y = tf32(relu(x * scale[c] + shift[c])) for c < C; x fp32 [R, ldx], y fp32 [R, ldy].  4 channels per thread.
__global__ void __launch_bounds__(256) bnrelu_kernel(const float* __restrict__ x, int ldx, int R, int C,
                                                     const float* __restrict__ scale, const float* __restrict__ shift,
                                                     float* __restrict__ y, int ldy) {
  const int per_row = C / 4;
  const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (size_t)R * per_row) return;
  const int r = (int)(idx / per_row), c = (int)(idx % per_row) * 4;
  const float4 v = *reinterpret_cast<const float4*>(x + (size_t)r * ldx + c);
  const float4 s = __ldg(reinterpret_cast<const float4*>(scale + c));
  const float4 b = __ldg(reinterpret_cast<const float4*>(shift + c));
  float4 o;
  o.x = ptx::round_tf32(fmaxf(v.x * s.x + b.x, 0.f));
  o.y = ptx::round_tf32(fmaxf(v.y * s.y + b.y, 0.f));
  o.z = ptx::round_tf32(fmaxf(v.z * s.z + b.z, 0.f));
  o.w = ptx::round_tf32(fmaxf(v.w * s.w + b.w, 0.f));
  *reinterpret_cast<float4*>(y + (size_t)r * ldy + c) = o;
}

// CAM gate, step 1: per-segment column sums of h (fp32 [R50, 128], post BN+ReLU bottleneck).
// One block of 128 threads per global segment g; seg_utt[g] = utterance, seg_base[u] = first segment of u.
__global__ void __launch_bounds__(128) cam_segsum_kernel(const float* __restrict__ h,
                                                         const int* __restrict__ utt_B, const int* __restrict__ utt_T,
                                                         const int* __restrict__ seg_base, const int* __restrict__ seg_utt,
                                                         int seg_len, float* __restrict__ seg_sum) {
  const int g = blockIdx.x, c = threadIdx.x;
  const int u = seg_utt[g];
  const int s = g - seg_base[u];
  const int B = utt_B[u], T = utt_T[u];
  const int t0 = s * seg_len, t1 = min(T, t0 + seg_len);
  float a = 0.f;
  for (int t = t0; t < t1; ++t) a += h[(size_t)(B + t) * 128 + c];
  seg_sum[(size_t)g * 128 + c] = a;
}

// CAM gate, step 2: ctx = mean over the whole utterance + mean over the segment (last segment partial, like
// avg_pool1d(ceil_mode=True)); gate[g] = sigmoid(W2 relu(W1 ctx + b1) + b2).  One block per global segment.
__global__ void __launch_bounds__(128) cam_gate_kernel(const float* __restrict__ seg_sum,
                                                       const int* __restrict__ utt_T, const int* __restrict__ seg_base,
                                                       const int* __restrict__ seg_utt, int seg_len,
                                                       const float* __restrict__ w1 /* [128][64] (in, out) */,
                                                       const float* __restrict__ b1,
                                                       const float* __restrict__ w2 /* [64][32] (in, out) */,
                                                       const float* __restrict__ b2, float* __restrict__ gate) {
  __shared__ float ctx[128];
  __shared__ float hid[64];
  const int g = blockIdx.x, c = threadIdx.x;
  const int u = seg_utt[g];
  const int sb = seg_base[u];
  const int s = g - sb;
  const int T = utt_T[u];
  const int nseg = (T + seg_len - 1) / seg_len;
  float total = 0.f;
  for (int k = 0; k < nseg; ++k) total += seg_sum[(size_t)(sb + k) * 128 + c];
  const int n = min(T, (s + 1) * seg_len) - s * seg_len;
  ctx[c] = total / (float)T + seg_sum[(size_t)g * 128 + c] / (float)n;
  __syncthreads();
  {  // 64 outputs x 128 inputs: two threads per output, combined by shuffle
    const int o = c >> 1, half = c & 1;
    float a = 0.f;
    for (int k = half * 64; k < half * 64 + 64; ++k) a += w1[k * 64 + o] * ctx[k];
    a += __shfl_xor_sync(0xffffffffu, a, 1);
    if (half == 0) hid[o] = fmaxf(a + b1[o], 0.f);
  }
  __syncthreads();
  if (c < 32) {
    float a = b2[c];
    for (int k = 0; k < 64; ++k) a += w2[k * 32 + c] * hid[k];
    gate[(size_t)g * 32 + c] = 1.0f / (1.0f + expf(-a));
  }
}

// Statistics pooling after BN+ReLU: stats[u] = [mean(1..C), std_unbiased(1..C)] over the utterance's rows.
// x fp32 [R50, ldx]; one block per utterance, one thread per channel (C <= 512).
__global__ void __launch_bounds__(512) stats_kernel(const float* __restrict__ x, int ldx, int C,
                                                    const int* __restrict__ utt_B, const int* __restrict__ utt_T,
                                                    const float* __restrict__ scale, const float* __restrict__ shift,
                                                    float* __restrict__ stats) {
  const int u = blockIdx.x, c = threadIdx.x;
  if (c >= C) return;
  const int B = utt_B[u], T = utt_T[u];
  const float s = scale[c], b = shift[c];
  float sum = 0.f;
  for (int t = 0; t < T; ++t) sum += fmaxf(x[(size_t)(B + t) * ldx + c] * s + b, 0.f);
  const float mean = sum / (float)T;
  float ss = 0.f;
  for (int t = 0; t < T; ++t) {
    const float d = fmaxf(x[(size_t)(B + t) * ldx + c] * s + b, 0.f) - mean;
    ss += d * d;
  }
  stats[(size_t)u * 2 * C + c] = mean;
  stats[(size_t)u * 2 * C + C + c] = sqrtf(ss / (float)(T - 1));
}

// emb[u][o] = (sum_k W[o][k] stats[u][k]) * scale[o] + shift[o]; one block per utterance, one warp per 8 outputs.
__global__ void __launch_bounds__(256) dense_kernel(const float* __restrict__ stats, int K, int O,
                                                    const float* __restrict__ w, const float* __restrict__ scale,
                                                    const float* __restrict__ shift, float* __restrict__ emb) {
  extern __shared__ float sx[];
  const int u = blockIdx.x;
  for (int i = threadIdx.x; i < K; i += blockDim.x) sx[i] = stats[(size_t)u * K + i];
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int o = warp; o < O; o += 8) {
    float a = 0.f;
    for (int k = lane; k < K; k += 32) a += w[(size_t)o * K + k] * sx[k];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) a += __shfl_xor_sync(0xffffffffu, a, d);
    if (lane == 0) emb[(size_t)u * O + o] = a * scale[o] + shift[o];
  }
}

}  // namespace camk
}  // namespace fy
