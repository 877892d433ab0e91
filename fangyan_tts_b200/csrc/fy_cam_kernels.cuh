// CUDA-core kernels of the CAMPPlus speaker network (row A6): the 1->32 stem convolution, BN+ReLU operand
// preparation, the context-aware mask (global + 100-frame segment means -> 2-layer MLP -> sigmoid),
// statistics pooling and the final dense layer.  Convolutions with >= 32 input channels run on the tcgen05 GEMM.
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace camk {

// Activations of this network are stored as IEEE fp16: the same 11-bit significand as tf32 at half the bytes.  Every
// stored value is post-BatchNorm / post-gate (O(1..100)), far inside fp16's range.
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ float2 unpack2(uint32_t u) { return __half22float2(*reinterpret_cast<const __half2*>(&u)); }

// Stem: Conv2d(1, 32, 3x3, pad 1) + BN + ReLU over (freq, time).  feat is packed [frames, 80] fp32.
// out: [R100, 82, 32] fp16 channels-last with zero columns at p = 0 and p = 81 (zero_pads_kernel) and zero guard rows.
// rowsrc[r] = feat row of time row r or -1 for guard rows.  w: [32][3 (freq)][3 (time)].
// One block = kStemRows consecutive time rows, 320 threads (80 freq x 4 groups of 8 channels): the rows' sources, the
// kStemRows + 2 input rows and the 288 weights are fetched once per block (one block per row spent its life on three
// dependent round trips for 100 instructions of work, profiles/r02_ncu_small_kernels.md) and a thread keeps its
// 8 x 9 weights in registers across the rows.  grid = ceil(R100 / kStemRows).
constexpr int kStemRows = 8;
__global__ void __launch_bounds__(320, 2) stem_kernel(const float* __restrict__ feat, const int* __restrict__ rowsrc,
                                                      int R100, const float* __restrict__ w,
                                                      const float* __restrict__ scale, const float* __restrict__ shift,
                                                      __half* __restrict__ out) {
  __shared__ int srow[kStemRows + 2];
  __shared__ float in[kStemRows + 2][82];
  __shared__ float sw[32 * 9 + 64];
  const int r0 = blockIdx.x * kStemRows;
  if (threadIdx.x < kStemRows + 2) {
    const int rr = r0 + (int)threadIdx.x - 1;
    srow[threadIdx.x] = (rr >= 0 && rr < R100) ? rowsrc[rr] : -1;
  }
  for (int i = threadIdx.x; i < 32 * 9; i += 320) sw[i] = w[i];
  for (int i = threadIdx.x; i < 64; i += 320) sw[288 + i] = i < 32 ? scale[i] : shift[i - 32];
  __syncthreads();
  for (int i = threadIdx.x; i < (kStemRows + 2) * 82; i += 320) {
    const int j = i / 82, p = i % 82;
    const int s = srow[j];
    in[j][p] = (s >= 0 && p >= 1 && p <= 80) ? feat[(size_t)s * 80 + p - 1] : 0.f;
  }
  __syncthreads();
  const int f = threadIdx.x % 80, cg = threadIdx.x / 80;
  float wr[8][9];
#pragma unroll
  for (int k = 0; k < 8; ++k)
#pragma unroll
    for (int q = 0; q < 9; ++q) wr[k][q] = sw[(cg * 8 + k) * 9 + q];
#pragma unroll 1
  for (int t = 0; t < kStemRows; ++t) {
    const int r = r0 + t;
    if (r >= R100) break;
    __half* dst = out + ((size_t)r * 82 + f + 1) * 32 + cg * 8;
    if (srow[t + 1] < 0) {   // guard row: explicit zeros (the workspace is not cleared; the pad slots come from zero_pads_kernel)
      *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
      continue;
    }
    float x[3][3];  // [freq i][time j]
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
      for (int j = 0; j < 3; ++j) x[i][j] = in[t + j][f + i];
    float y[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      float a = 0.f;
#pragma unroll
      for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 3; ++j) a += wr[k][i * 3 + j] * x[i][j];
      y[k] = fmaxf(a * sw[288 + cg * 8 + k] + sw[320 + cg * 8 + k], 0.f);
    }
    uint4 o;
    o.x = pack2(y[0], y[1]); o.y = pack2(y[2], y[3]); o.z = pack2(y[4], y[5]); o.w = pack2(y[6], y[7]);
    *reinterpret_cast<uint4*>(dst) = o;
  }
}

// The FCM activation buffers [R100][fp slots][32 ch] fp16 keep one zero pad slot at each end of every time row (the
// frequency padding of the 3x3 convolutions) plus 64 bytes after the last row (second half of the last overlapping
// view row).  The convolution epilogues never write those, and the workspace is reused across calls and stages, so
// they are cleared here -- 128 bytes per row instead of a memset of every buffer (3.5 GB at 250 utterances).
struct PadBufs {
  __half* ptr[9];
  int fp[9];
  int n;
};
__global__ void __launch_bounds__(256) zero_pads_kernel(PadBufs b, int R100) {
  const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
  const size_t per_buf = (size_t)(R100 + 1) * 8;        // 8 x 16 bytes per row, one extra "row" for the tail
  if (idx >= per_buf * b.n) return;
  const int buf = (int)(idx / per_buf);
  const size_t rem = idx % per_buf;
  const int r = (int)(rem >> 3), k = (int)(rem & 7);
  char* base = reinterpret_cast<char*>(b.ptr[buf]);
  const size_t row_bytes = (size_t)b.fp[buf] * 64;
  char* dst;
  if (r == R100) {
    if (k >= 4) return;
    dst = base + (size_t)R100 * row_bytes + k * 16;                                  // the 64-byte tail
  } else {
    dst = base + (size_t)r * row_bytes + (k < 4 ? 0 : row_bytes - 64) + (k & 3) * 16;  // first / last slot of row r
  }
  *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
}

// mask[g] = -2 (skip) when slot g % Q is outside [lo, hi]; else -1 (write zero) on guard time rows; else 0
__global__ void pos_mask_kernel(const int* __restrict__ rowsrc, int R, int Q, int lo, int hi, int* __restrict__ mask) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= R * Q) return;
  const int p = g % Q;
  mask[g] = (p < lo || p > hi) ? -2 : (rowsrc[g / Q] >= 0 ? 0 : -1);
}

// y = fp16(relu(x * scale[c] + shift[c])) for c < C; x fp16 [R, ldx], y fp16 [R, ldy].  8 channels per thread.
__global__ void __launch_bounds__(256) bnrelu_kernel(const __half* __restrict__ x, int ldx, int R, int C,
                                                     const float* __restrict__ scale, const float* __restrict__ shift,
                                                     __half* __restrict__ y, int ldy) {
  const int per_row = C / 8;
  const size_t idx = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (idx >= (size_t)R * per_row) return;
  const int r = (int)(idx / per_row), c = (int)(idx % per_row) * 8;
  const uint4 v = *reinterpret_cast<const uint4*>(x + (size_t)r * ldx + c);
  const float4 s0 = __ldg(reinterpret_cast<const float4*>(scale + c)), s1 = __ldg(reinterpret_cast<const float4*>(scale + c + 4));
  const float4 b0 = __ldg(reinterpret_cast<const float4*>(shift + c)), b1 = __ldg(reinterpret_cast<const float4*>(shift + c + 4));
  const float2 a0 = unpack2(v.x), a1 = unpack2(v.y), a2 = unpack2(v.z), a3 = unpack2(v.w);
  uint4 o;
  o.x = pack2(fmaxf(a0.x * s0.x + b0.x, 0.f), fmaxf(a0.y * s0.y + b0.y, 0.f));
  o.y = pack2(fmaxf(a1.x * s0.z + b0.z, 0.f), fmaxf(a1.y * s0.w + b0.w, 0.f));
  o.z = pack2(fmaxf(a2.x * s1.x + b1.x, 0.f), fmaxf(a2.y * s1.y + b1.y, 0.f));
  o.w = pack2(fmaxf(a3.x * s1.z + b1.z, 0.f), fmaxf(a3.y * s1.w + b1.w, 0.f));
  *reinterpret_cast<uint4*>(y + (size_t)r * ldy + c) = o;
}

// CAM gate, step 1: per-segment column sums of h (fp16 [R50, 128], post BN+ReLU bottleneck).
// One block of 128 threads per global segment g; seg_utt[g] = utterance, seg_base[u] = first segment of u.
__global__ void __launch_bounds__(128) cam_segsum_kernel(const __half* __restrict__ h,
                                                         const int* __restrict__ utt_B, const int* __restrict__ utt_T,
                                                         const int* __restrict__ seg_base, const int* __restrict__ seg_utt,
                                                         int seg_len, float* __restrict__ seg_sum) {
  const int g = blockIdx.x, c = threadIdx.x;
  const int u = seg_utt[g];
  const int s = g - seg_base[u];
  const int B = utt_B[u], T = utt_T[u];
  const int t0 = s * seg_len, t1 = min(T, t0 + seg_len);
  float a = 0.f;
  for (int t = t0; t < t1; ++t) a += __half2float(h[(size_t)(B + t) * 128 + c]);
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

// CAM gate in ONE launch per dense layer (replaces cam_segsum_kernel + cam_gate_kernel): one block of 512 threads per
// utterance.  The block is pure latency (~1,300 instructions per warp, profiles/r02_ncu_small_kernels.md), so the design
// is about how many dependent round trips to memory it makes:
//   * W1 (32 KB) is requested first and parked in registers, then stored to shared memory after phase 1: its latency
//     hides under the segment sums and phase 2 never waits for global memory on it;
//   * phase 1, segment sums of h: a thread = one column pair x one of EIGHT row groups, the eight row groups of a column
//     pair are neighbouring lanes, so their partial sums meet by three shuffles -- no shared memory, no block barrier in
//     the segment loop, one batch of <= 16 loads per 100-frame segment;
//   * phase 2, for every segment ctx = utterance mean + segment mean, gate = sigmoid(W2 relu(W1 ctx + b1) + b2): four
//     segments per pass (128 threads each), i.e. one pass for clips up to 8 s.
// An utterance is handled by one block whatever the batch holds and every sum has a fixed order, so the gates do not
// depend on the batch composition.  kGateMaxSeg segments fit the shared buffer (a 32 s clip at seg_len 100); longer
// utterances take the two-kernel path.  __launch_bounds__(512, 2): without the second number ptxas squeezes the kernel
// into 40 registers and, to get there, issues every load right in front of its use (sixteen serial round trips per
// segment instead of one batch -- visible in the SASS); 64 registers keep two blocks per SM, one wave for <= 296 clips.
constexpr int kGateMaxSeg = 16;
constexpr int kGateThreads = 512;
constexpr int kGateBatch = 16;                    // rows a thread requests at a time in phase 1
__global__ void __launch_bounds__(kGateThreads, 2) cam_ctx_gate_kernel(const __half* __restrict__ h, const int* __restrict__ utt_B,
                                                                    const int* __restrict__ utt_T, const int* __restrict__ seg_base,
                                                                    int seg_len, const float* __restrict__ w1, const float* __restrict__ b1,
                                                                    const float* __restrict__ w2, const float* __restrict__ b2,
                                                                    float* __restrict__ gate) {
  __shared__ float ssum[kGateMaxSeg][128];
  __shared__ float total[128];
  __shared__ float ctx[4][128];
  __shared__ float hid[4][64];
  __shared__ float4 sw1[128 * 64 / 4];             // W1 [128 inputs][64 outputs]
  const int u = blockIdx.x, tid = threadIdx.x;
  float4 wreg[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) wreg[i] = __ldg(reinterpret_cast<const float4*>(w1) + tid + kGateThreads * i);
  const int B = utt_B[u], T = utt_T[u], sb = seg_base[u];
  const int nseg = (T + seg_len - 1) / seg_len;
  const int cp = tid >> 3, rg = tid & 7;           // column pair, row group (eight neighbouring lanes)
  const int grp = tid >> 7, c = tid & 127;         // phase 2: segment of the pass, channel
  const float b1v = __ldg(b1 + (c >> 1)), b2v = __ldg(b2 + (c & 31));
  const __half2* hp = reinterpret_cast<const __half2*>(h) + (size_t)B * 64 + cp;
  float tx = 0.f, ty = 0.f;
  for (int s = 0; s < nseg; ++s) {
    const int t0 = s * seg_len, t1 = min(T, t0 + seg_len);
    float ax = 0.f, ay = 0.f;
    // kGateBatch predicated loads at a time (a fixed trip count: a `#pragma unroll` of the plain row loop leaves short
    // trip counts to a serial remainder loop, one round trip per row); rows beyond the segment add exact zeros
    for (int base = t0 + rg; base < t1; base += 8 * kGateBatch) {
      __half2 v[kGateBatch];
#pragma unroll
      for (int i = 0; i < kGateBatch; ++i)        // unconditional loads (row index clamped into the segment) ...
        v[i] = __ldg(hp + (size_t)min(base + 8 * i, t1 - 1) * 64);
#pragma unroll
      for (int i = 0; i < kGateBatch; ++i) {      // ... masked when they are added
        const float2 f = __half22float2(v[i]);
        const bool in = base + 8 * i < t1;
        ax += in ? f.x : 0.f; ay += in ? f.y : 0.f;
      }
    }
#pragma unroll
    for (int m = 1; m < 8; m <<= 1) {
      ax += __shfl_xor_sync(0xffffffffu, ax, m);
      ay += __shfl_xor_sync(0xffffffffu, ay, m);
    }
    if (rg == 0) { ssum[s][2 * cp] = ax; ssum[s][2 * cp + 1] = ay; }
    tx += ax; ty += ay;
  }
  if (rg == 0) { total[2 * cp] = tx; total[2 * cp + 1] = ty; }
#pragma unroll
  for (int i = 0; i < 4; ++i) sw1[tid + kGateThreads * i] = wreg[i];
  __syncthreads();
  const float* w1s = reinterpret_cast<const float*>(sw1);
  // four segments per pass: threads [128 g, 128 g + 128) take segment s0 + g
  for (int s0 = 0; s0 < nseg; s0 += 4) {
    const int s = s0 + grp;
    const bool on = s < nseg;
    if (on) {
      const int n = min(T, (s + 1) * seg_len) - s * seg_len;
      ctx[grp][c] = total[c] / (float)T + ssum[s][c] / (float)n;
    }
    __syncthreads();
    if (on) {   // 64 outputs x 128 inputs: two threads per output, combined by shuffle
      const int o = c >> 1, hf = c & 1;
      float a = 0.f;
#pragma unroll 16
      for (int k = hf * 64; k < hf * 64 + 64; ++k) a += w1s[k * 64 + o] * ctx[grp][k];
      a += __shfl_xor_sync(0xffffffffu, a, 1);
      if (hf == 0) hid[grp][o] = fmaxf(a + b1v, 0.f);
    }
    __syncthreads();
    if (on && c < 32) {
      float a = b2v;
#pragma unroll
      for (int k0 = 0; k0 < 64; k0 += 16) {     // sixteen weight loads in flight, then their FMAs in order
        float wv[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) wv[i] = __ldg(w2 + (k0 + i) * 32 + c);
#pragma unroll
        for (int i = 0; i < 16; ++i) a += wv[i] * hid[grp][k0 + i];
      }
      gate[(size_t)(sb + s) * 32 + c] = 1.0f / (1.0f + expf(-a));
    }
    __syncthreads();
  }
}

// Statistics pooling after BN+ReLU: stats[u] = [mean(1..C), std_unbiased(1..C)] over the utterance's rows.
// x fp16 [R50, ldx]; one block per utterance, one thread per channel (C <= 512).
__global__ void __launch_bounds__(512) stats_kernel(const __half* __restrict__ x, int ldx, int C,
                                                    const int* __restrict__ utt_B, const int* __restrict__ utt_T,
                                                    const float* __restrict__ scale, const float* __restrict__ shift,
                                                    float* __restrict__ stats) {
  const int u = blockIdx.x, c = threadIdx.x;
  if (c >= C) return;
  const int B = utt_B[u], T = utt_T[u];
  const float s = scale[c], b = shift[c];
  float sum = 0.f;
#pragma unroll 8
  for (int t = 0; t < T; ++t) sum += fmaxf(__half2float(x[(size_t)(B + t) * ldx + c]) * s + b, 0.f);
  const float mean = sum / (float)T;
  float ss = 0.f;
#pragma unroll 8
  for (int t = 0; t < T; ++t) {
    const float d = fmaxf(__half2float(x[(size_t)(B + t) * ldx + c]) * s + b, 0.f) - mean;
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
