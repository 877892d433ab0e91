// Audio front-end kernels (rows A1-A5): PCM decode, polyphase resampling, Kaldi fbank + CMN, Whisper log-mel.
// All fp32 on CUDA cores: these stages are HBM/L2-bound streaming work and must stay fp32-accurate
// (the mel floor sits 8 decades below the peak, far beyond what a bf16 DFT-as-GEMM can resolve).
//
// Each warp transforms TWO real frames with one complex FFT (frame A in the real lane, frame B in the
// imaginary lane) held in shared memory, Stockham autosort, radix 4/4/4/4/2 for 512 and 4/4/5/5 for 400.
#include <math.h>

#include <algorithm>

#include <vector>

#include <mutex>

#include "fy_common.cuh"

// Scratch of fy_fbank_cmn_static (the entry point that allocates nothing, end of this file): per-CTA column sums and
// per-utterance means of the second-generation fbank.  Resident on every GPU that loads the library until the process
// exits.  In HOST code this name is only a stand-in for the symbol: kernels get the device address that
// fbank_static_setup() asks the runtime for (cudaGetSymbolAddress), once per device.
static __device__ __align__(256) unsigned char g_fbank_static_ws[FY_FBANK_STATIC_WS_BYTES];

namespace fy {
namespace {

constexpr int kWarpsPerBlock = 4;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 mul_negi(float2 a) { return make_float2(a.y, -a.x); }  // a * (-i)

// Shared-memory FFT buffers are indexed through sidx(): one float2 of padding after every 16 keeps the radix-4
// stages' strided stores (stride 4 and 16 complex = 32 / 128 bytes) off the same banks (8-way conflicts otherwise).
__device__ __forceinline__ int sidx(int i) { return i + (i >> 4); }
constexpr int spad(int n) { return n + (n >> 4) + 1; }

// One Stockham stage of radix R over a length-N transform whose previous radices multiply to s.
//   y[q + s*(R*p + k)] = w^(k) * sum_j x[q + s*(p + m*j)] * exp(-2 pi i jk/R),  w = exp(-2 pi i p s / N)
// NT cooperating threads: 32 = one warp (__syncwarp between stages), otherwise the whole block (__syncthreads).
template <int N, int R, int NT = 32>
__device__ __forceinline__ void fft_stage(const float2* __restrict__ x, float2* __restrict__ y, int s,
                                          const float2* __restrict__ tw, int lane) {
  constexpr int NB = N / R;
  const int m = NB / s;
  for (int i = lane; i < NB; i += NT) {
    const int q = i % s;
    const int p = i / s;
    float2 a[R];
#pragma unroll
    for (int j = 0; j < R; ++j) a[j] = x[sidx(i + NB * j)];
    float2 b[R];
    if (R == 2) {
      b[0] = cadd(a[0], a[1]);
      b[1] = csub(a[0], a[1]);
    } else if (R == 4) {
      const float2 t0 = cadd(a[0], a[2]), t1 = csub(a[0], a[2]);
      const float2 t2 = cadd(a[1], a[3]), t3 = mul_negi(csub(a[1], a[3]));
      b[0] = cadd(t0, t2);
      b[1] = cadd(t1, t3);
      b[2] = csub(t0, t2);
      b[3] = csub(t1, t3);
    } else if (R == 3) {
      const float2 w1 = make_float2(-0.5f, -0.86602540378443865f), w2 = make_float2(-0.5f, 0.86602540378443865f);
      b[0] = cadd(a[0], cadd(a[1], a[2]));
      b[1] = cadd(a[0], cadd(cmul(a[1], w1), cmul(a[2], w2)));
      b[2] = cadd(a[0], cadd(cmul(a[1], w2), cmul(a[2], w1)));
    } else {  // R == 5, direct
      const float c1 = 0.30901699437494745f, s1 = 0.95105651629515353f;
      const float c2 = -0.80901699437494734f, s2 = 0.58778525229247314f;
      const float2 w1 = make_float2(c1, -s1), w2 = make_float2(c2, -s2), w3 = make_float2(c2, s2),
                   w4 = make_float2(c1, s1);
      const float2 w[5] = {make_float2(1.f, 0.f), w1, w2, w3, w4};
#pragma unroll
      for (int k = 0; k < 5; ++k) {
        float2 acc = a[0];
#pragma unroll
        for (int j = 1; j < 5; ++j) acc = cadd(acc, cmul(a[j], w[(j * k) % 5]));
        b[k] = acc;
      }
    }
    const int base = q + s * R * p;
    const int step = p * s;  // twiddle index of w
    y[sidx(base)] = b[0];
#pragma unroll
    for (int k = 1; k < R; ++k) y[sidx(base + s * k)] = (m == 1) ? b[k] : cmul(b[k], tw[(step * k) % N]);
  }
  if (NT == 32) __syncwarp(); else __syncthreads();
}

// 512-point: result ends in `a` when the number of stages is odd (5 stages: a->b->a->b->a->b ... see below).
__device__ __forceinline__ float2* fft512(float2* a, float2* b, const float2* tw, int lane) {
  fft_stage<512, 4>(a, b, 1, tw, lane);
  fft_stage<512, 4>(b, a, 4, tw, lane);
  fft_stage<512, 4>(a, b, 16, tw, lane);
  fft_stage<512, 4>(b, a, 64, tw, lane);
  fft_stage<512, 2>(a, b, 256, tw, lane);
  return b;
}
__device__ __forceinline__ float2* fft400(float2* a, float2* b, const float2* tw, int lane) {
  fft_stage<400, 4>(a, b, 1, tw, lane);
  fft_stage<400, 4>(b, a, 4, tw, lane);
  fft_stage<400, 5>(a, b, 16, tw, lane);
  fft_stage<400, 5>(b, a, 80, tw, lane);
  return a;
}

__device__ __forceinline__ int find_utt(const int* __restrict__ off, int n, int g) {
  // largest b with off[b] <= g  (off has n+1 entries, off[n] > g)
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(off + mid) <= g) lo = mid; else hi = mid;
  }
  return lo;
}

// Frame pairs never straddle utterances: utterance u owns pair slots [s(u), s(u+1)), s(u) = (frame_off[u] + u) >> 1,
// which is always at least ceil(M_u / 2) slots.  A frame's FFT partner (the other lane of the packed complex
// transform, whose rounding noise leaks into it at the 1e-5 level) is therefore fixed by its own utterance, and
// every result is bit-identical whatever else is in the batch.  Returns the utterance of slot w; n = #utterances.
__device__ __forceinline__ int find_slot_utt(const int* __restrict__ off, int n, int w) {
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (((__ldg(off + mid) + mid) >> 1) <= w) lo = mid; else hi = mid;
  }
  return lo;
}

template <bool S16>
__device__ __forceinline__ float load_pcm(const void* pcm, int idx) {
  if (S16) return (float)__ldg(reinterpret_cast<const short*>(pcm) + idx) * (1.0f / 32768.0f);
  return __ldg(reinterpret_cast<const float*>(pcm) + idx);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ------------------------------------------------------------------------------------------ Kaldi fbank
struct MelTable {
  const float* w;
  const int* off;   // first FFT bin
  const int* len;   // taps
  const int* ptr;   // offset into w
};

template <bool S16>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) fbank_kernel(const void* __restrict__ pcm,
                                                                    const int* __restrict__ pcm_off,
                                                                    const int* __restrict__ frame_off, int n_utts,
                                                                    int total_frames, const float* __restrict__ window,
                                                                    const float2* __restrict__ tw, MelTable mt,
                                                                    float* __restrict__ feat) {
  __shared__ float2 sbuf[kWarpsPerBlock][2][spad(512)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* A = sbuf[warp][0];
  float2* B = sbuf[warp][1];
  const int slot = blockIdx.x * kWarpsPerBlock + warp;
  if (slot >= ((total_frames + n_utts) >> 1)) return;
  const int u = find_slot_utt(frame_off, n_utts, slot);
  const int fo = __ldg(frame_off + u);
  const int f0 = 2 * (slot - ((fo + u) >> 1));          // first frame of the pair inside utterance u
  const int nf = __ldg(frame_off + u + 1) - fo;
  if (f0 >= nf) return;
  const int gf0 = fo + f0;
  int start[2];
  bool valid[2];
#pragma unroll
  for (int f = 0; f < 2; ++f) {
    valid[f] = f0 + f < nf;
    start[f] = __ldg(pcm_off + u) + (f0 + f) * 160;
  }
  // raw samples -> B, frame sums
  float s0 = 0.f, s1 = 0.f;
  for (int i = lane; i < 400; i += 32) {
    const float x0 = valid[0] ? load_pcm<S16>(pcm, start[0] + i) : 0.f;
    const float x1 = valid[1] ? load_pcm<S16>(pcm, start[1] + i) : 0.f;
    B[sidx(i)] = make_float2(x0, x1);
    s0 += x0;
    s1 += x1;
  }
  const float m0 = warp_sum(s0) * (1.0f / 400.0f);
  const float m1 = warp_sum(s1) * (1.0f / 400.0f);
  __syncwarp();
  // DC removal, pre-emphasis (replicate-left), Povey window, zero pad to 512 -> A
  for (int i = lane; i < 512; i += 32) {
    float2 v = make_float2(0.f, 0.f);
    if (i < 400) {
      const float2 c = B[sidx(i)];
      const float2 p = B[sidx(i > 0 ? i - 1 : 0)];
      const float w = __ldg(window + i);
      v.x = ((c.x - m0) - 0.97f * (p.x - m0)) * w;
      v.y = ((c.y - m1) - 0.97f * (p.y - m1)) * w;
    }
    A[sidx(i)] = v;
  }
  __syncwarp();
  float2* Z = fft512(A, B, tw, lane);
  float* P = reinterpret_cast<float*>(Z == A ? B : A);  // power spectra: P[k] frame 0, P[257 + k] frame 1
  for (int k = lane; k <= 256; k += 32) {
    const float2 z = Z[sidx(k)];
    const float2 zc = Z[sidx((512 - k) & 511)];
    const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);   // frame 0 spectrum
    const float br = 0.5f * (z.y + zc.y), bi = -0.5f * (z.x - zc.x);  // frame 1 spectrum
    P[k] = ar * ar + ai * ai;
    P[257 + k] = br * br + bi * bi;
  }
  __syncwarp();
  for (int b = lane; b < 80; b += 32) {
    const int o = __ldg(mt.off + b), n = __ldg(mt.len + b);
    const float* w = mt.w + __ldg(mt.ptr + b);
    float a0 = 0.f, a1 = 0.f;
    for (int j = 0; j < n; ++j) {
      const float wj = __ldg(w + j);
      a0 += wj * P[o + j];
      a1 += wj * P[257 + o + j];
    }
    if (valid[0]) feat[(size_t)gf0 * 80 + b] = logf(fmaxf(a0, 1.1920929e-07f));
    if (valid[1]) feat[(size_t)(gf0 + 1) * 80 + b] = logf(fmaxf(a1, 1.1920929e-07f));
  }
}

// CMN: feat[b] -= mean over its frames.  One block per utterance, 320 threads = 80 bins x 4 frame groups.
__global__ void __launch_bounds__(320) cmn_kernel(float* __restrict__ feat, const int* __restrict__ frame_off) {
  __shared__ float part[4][80];
  const int u = blockIdx.x;
  const int f0 = frame_off[u], f1 = frame_off[u + 1];
  const int col = threadIdx.x % 80, grp = threadIdx.x / 80;
  float acc = 0.f;
  for (int f = f0 + grp; f < f1; f += 4) acc += feat[(size_t)f * 80 + col];
  part[grp][col] = acc;
  __syncthreads();
  const int n = f1 - f0;
  if (n <= 0) return;
  const float mean = (part[0][col] + part[1][col] + part[2][col] + part[3][col]) / (float)n;
  for (int f = f0 + grp; f < f1; f += 4) feat[(size_t)f * 80 + col] -= mean;
}

// ------------------------------------------------------------------------------------------ Whisper log-mel
__device__ __forceinline__ int float_key(float f) {
  const int i = __float_as_int(f);
  return i >= 0 ? i : i ^ 0x7fffffff;
}
__device__ __forceinline__ float key_float(int k) { return __int_as_float(k >= 0 ? k : k ^ 0x7fffffff); }

__global__ void fill_int_kernel(int* p, int n, int v) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) p[i] = v;
}

#include "fy_fbank2.cuh"

// Pass 1: log10(max(mel, 1e-10)) time-major into raw[total_frames, 128]; per-utterance max into umax keys.
template <bool S16>
__global__ void __launch_bounds__(kWarpsPerBlock * 32) logmel_kernel(const void* __restrict__ pcm,
                                                                     const int* __restrict__ pcm_off,
                                                                     const int* __restrict__ frame_off, int n_utts,
                                                                     int total_frames, const float* __restrict__ window,
                                                                     const float2* __restrict__ tw, MelTable mt,
                                                                     float* __restrict__ raw, int* __restrict__ umax) {
  __shared__ float2 sbuf[kWarpsPerBlock][2][spad(400)];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float2* A = sbuf[warp][0];
  float2* B = sbuf[warp][1];
  const int slot = blockIdx.x * kWarpsPerBlock + warp;
  if (slot >= ((total_frames + n_utts) >> 1)) return;
  const int u = find_slot_utt(frame_off, n_utts, slot);
  const int fo = __ldg(frame_off + u);
  const int f0 = 2 * (slot - ((fo + u) >> 1));          // first frame of the pair inside utterance u
  const int nf = __ldg(frame_off + u + 1) - fo;
  if (f0 >= nf) return;
  const int gf0 = fo + f0;
  int base[2], nsamp[2], center[2], utt[2];
  bool valid[2];
#pragma unroll
  for (int f = 0; f < 2; ++f) {
    valid[f] = f0 + f < nf;
    utt[f] = u;
    base[f] = __ldg(pcm_off + u);
    nsamp[f] = __ldg(pcm_off + u + 1) - base[f];
    center[f] = (f0 + f) * 160;
  }
  for (int i = lane; i < 400; i += 32) {
    float2 v = make_float2(0.f, 0.f);
    const float w = __ldg(window + i);
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      if (valid[f]) {
        int idx = center[f] - 200 + i;
        if (idx < 0) idx = -idx;
        if (idx >= nsamp[f]) idx = 2 * (nsamp[f] - 1) - idx;
        const float x = load_pcm<S16>(pcm, base[f] + idx) * w;
        if (f == 0) v.x = x; else v.y = x;
      }
    }
    A[sidx(i)] = v;
  }
  __syncwarp();
  float2* Z = fft400(A, B, tw, lane);
  float* P = reinterpret_cast<float*>(Z == A ? B : A);
  for (int k = lane; k <= 200; k += 32) {
    const float2 z = Z[sidx(k)];
    const float2 zc = Z[sidx(k == 0 ? 0 : 400 - k)];
    const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
    const float br = 0.5f * (z.y + zc.y), bi = -0.5f * (z.x - zc.x);
    P[k] = ar * ar + ai * ai;
    P[201 + k] = br * br + bi * bi;
  }
  __syncwarp();
  float mx0 = -INFINITY, mx1 = -INFINITY;
  for (int b = lane; b < 128; b += 32) {
    const int o = __ldg(mt.off + b), n = __ldg(mt.len + b);
    const float* w = mt.w + __ldg(mt.ptr + b);
    float a0 = 0.f, a1 = 0.f;
    for (int j = 0; j < n; ++j) {
      const float wj = __ldg(w + j);
      a0 += wj * P[o + j];
      a1 += wj * P[201 + o + j];
    }
    const float l0 = log10f(fmaxf(a0, 1e-10f)), l1 = log10f(fmaxf(a1, 1e-10f));
    if (valid[0]) { raw[(size_t)gf0 * 128 + b] = l0; mx0 = fmaxf(mx0, l0); }
    if (valid[1]) { raw[(size_t)(gf0 + 1) * 128 + b] = l1; mx1 = fmaxf(mx1, l1); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, o));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, o));
  }
  if (lane == 0) {
    if (valid[0] && valid[1] && utt[0] == utt[1]) {
      atomicMax(umax + utt[0], float_key(fmaxf(mx0, mx1)));
    } else {
      if (valid[0]) atomicMax(umax + utt[0], float_key(mx0));
      if (valid[1]) atomicMax(umax + utt[1], float_key(mx1));
    }
  }
}

// Pass 2 (reference layout): mel[128 * frame_off[u] + b * T_u + t] = (max(raw, umax - 8) + 4) / 4.
// Block = 32 frames x 128 bins tile of utterance blockIdx.y, transposed through shared memory.
__global__ void __launch_bounds__(256) logmel_finish_kernel(const float* __restrict__ raw,
                                                            const int* __restrict__ frame_off,
                                                            const int* __restrict__ umax, float* __restrict__ mel) {
  __shared__ float tile[32][129];
  const int u = blockIdx.y;
  const int t0 = blockIdx.x * 32;
  const int f0 = frame_off[u];
  const int T = frame_off[u + 1] - f0;
  if (t0 >= T) return;
  const float floor_v = key_float(umax[u]) - 8.0f;
  for (int i = threadIdx.x; i < 32 * 128; i += 256) {
    const int t = i >> 7, b = i & 127;
    float v = 0.f;
    if (t0 + t < T) v = (fmaxf(raw[(size_t)(f0 + t0 + t) * 128 + b], floor_v) + 4.0f) * 0.25f;
    tile[t][b] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 32 * 128; i += 256) {
    const int b = i >> 5, t = i & 31;
    if (t0 + t < T) mel[(size_t)f0 * 128 + (size_t)b * T + t0 + t] = tile[t][b];
  }
}

// ------------------------------------------------------------------------------------------ matcha 24 kHz mel
// matcha.utils.audio.mel_spectrogram(n_fft 1920, hop 480, win 1920, 80 mels, fmin 0, fmax None, center=False)
// (third_party/Matcha-TTS/matcha/utils/audio.py:44-88): reflect-pad 720 samples each side, periodic Hann, |STFT| =
// sqrt(re^2 + im^2 + 1e-9), Slaney mel, log(max(., 1e-5)).  Output time-major [frames, 80] (both reference callers
// transpose to that: frontend.py:122, processor.py:174).  One block of 128 threads per frame pair of ONE utterance
// (pair slots as in find_slot_utt); the 1920-point complex transform (radix 4.4.4.5.3.2) carries two real frames.
constexpr int kMmN = 1920, kMmHop = 480, kMmPad = 720, kMmBins = 961, kMmMels = 80, kMmThreads = 128;
template <bool S16>
__global__ void __launch_bounds__(kMmThreads) matcha_mel_kernel(const void* __restrict__ pcm,
                                                                const int* __restrict__ pcm_off,
                                                                const int* __restrict__ frame_off, int n_utts,
                                                                int total_frames, const float* __restrict__ window,
                                                                const float2* __restrict__ tw, MelTable mt,
                                                                float* __restrict__ out) {
  __shared__ float2 A[spad(kMmN)];
  __shared__ float2 B[spad(kMmN)];
  const int tid = threadIdx.x;
  const int slot = blockIdx.x;
  if (slot >= ((total_frames + n_utts) >> 1)) return;
  const int u = find_slot_utt(frame_off, n_utts, slot);
  const int fo = __ldg(frame_off + u);
  const int f0 = 2 * (slot - ((fo + u) >> 1));
  const int nf = __ldg(frame_off + u + 1) - fo;
  if (f0 >= nf) return;
  const int base = __ldg(pcm_off + u);
  const int ns = __ldg(pcm_off + u + 1) - base;
  const bool two = f0 + 1 < nf;
  for (int i = tid; i < kMmN; i += kMmThreads) {
    const float w = __ldg(window + i);
    float2 v = make_float2(0.f, 0.f);
#pragma unroll
    for (int f = 0; f < 2; ++f) {
      if (f == 0 || two) {
        int idx = (f0 + f) * kMmHop + i - kMmPad;
        if (idx < 0) idx = -idx;
        if (idx >= ns) idx = 2 * (ns - 1) - idx;
        idx = max(0, min(idx, ns - 1));
        const float x = load_pcm<S16>(pcm, base + idx) * w;
        if (f == 0) v.x = x; else v.y = x;
      }
    }
    A[sidx(i)] = v;
  }
  __syncthreads();
  fft_stage<kMmN, 4, kMmThreads>(A, B, 1, tw, tid);
  fft_stage<kMmN, 4, kMmThreads>(B, A, 4, tw, tid);
  fft_stage<kMmN, 4, kMmThreads>(A, B, 16, tw, tid);
  fft_stage<kMmN, 5, kMmThreads>(B, A, 64, tw, tid);
  fft_stage<kMmN, 3, kMmThreads>(A, B, 320, tw, tid);
  fft_stage<kMmN, 2, kMmThreads>(B, A, 960, tw, tid);
  float* P = reinterpret_cast<float*>(B);   // magnitudes: P[k] frame 0, P[961 + k] frame 1
  for (int k = tid; k < kMmBins; k += kMmThreads) {
    const float2 z = A[sidx(k)];
    const float2 zc = A[sidx(k == 0 ? 0 : kMmN - k)];
    const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
    const float br = 0.5f * (z.y + zc.y), bi = -0.5f * (z.x - zc.x);
    P[k] = sqrtf(ar * ar + ai * ai + 1e-9f);
    P[kMmBins + k] = sqrtf(br * br + bi * bi + 1e-9f);
  }
  __syncthreads();
  for (int o = tid; o < 2 * kMmMels; o += kMmThreads) {
    const int f = o / kMmMels, b = o - f * kMmMels;
    if (f == 1 && !two) continue;
    const int off = __ldg(mt.off + b), n = __ldg(mt.len + b);
    const float* w = mt.w + __ldg(mt.ptr + b);
    const float* p = P + f * kMmBins + off;
    float acc = 0.f;
    for (int j = 0; j < n; ++j) acc = fmaf(__ldg(w + j), p[j], acc);
    out[(size_t)(fo + f0 + f) * kMmMels + b] = logf(fmaxf(acc, 1e-5f));
  }
}

// ------------------------------------------------------------------------------------------ resample
__global__ void resample_kernel(const float* __restrict__ in, const int* __restrict__ in_off, float* __restrict__ out,
                                const int* __restrict__ out_off, int n_utts, int total_out,
                                const float* __restrict__ kern, int width, int orig, int newf) {
  const int g = blockIdx.x * blockDim.x + threadIdx.x;
  if (g >= total_out) return;
  const int u = find_utt(out_off, n_utts, g);
  const int o = g - __ldg(out_off + u);
  const int ib = __ldg(in_off + u);
  const int n_in = __ldg(in_off + u + 1) - ib;
  const int n = o / newf, p = o % newf;
  const int K = 2 * width + orig;
  const float* kp = kern + (size_t)p * K;
  const int x0 = n * orig - width;
  float acc = 0.f;
  for (int k = 0; k < K; ++k) {
    const int xi = x0 + k;
    if (xi >= 0 && xi < n_in) acc += __ldg(kp + k) * __ldg(in + ib + xi);
  }
  out[g] = acc;
}

// ------------------------------------------------------------------------------------------ host tables
template <typename T>
int upload(Ctx* c, const std::vector<T>& h, T** d) {
  FY_CUDA(cudaMalloc(d, h.size() * sizeof(T)));
  c->owned.push_back(*d);
  FY_CUDA(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return FY_OK;
}

int upload_mel(Ctx* c, const std::vector<std::vector<float>>& bank, float** dw, int** doff, int** dlen, int** dptr,
               int* n_weights = nullptr) {
  std::vector<float> w;
  std::vector<int> off, len, ptr;
  for (const auto& row : bank) {
    int first = -1, last = -1;
    for (int k = 0; k < (int)row.size(); ++k)
      if (row[k] != 0.0f) { if (first < 0) first = k; last = k; }
    if (first < 0) { first = 0; last = -1; }
    off.push_back(first);
    len.push_back(last - first + 1);
    ptr.push_back((int)w.size());
    for (int k = first; k <= last; ++k) w.push_back(row[k]);
  }
  if (w.empty()) w.push_back(0.f);
  if (n_weights) *n_weights = (int)w.size();
  FY_TRY(upload(c, w, dw));
  FY_TRY(upload(c, off, doff));
  FY_TRY(upload(c, len, dlen));
  FY_TRY(upload(c, ptr, dptr));
  return FY_OK;
}

std::vector<float2> twiddles(int n) {
  std::vector<float2> t(n);
  for (int i = 0; i < n; ++i) {
    const double a = -2.0 * M_PI * (double)i / (double)n;
    t[i] = make_float2((float)cos(a), (float)sin(a));
  }
  return t;
}

// Kaldi triangular mel bank, computed in fp32 exactly as torchaudio.compliance.kaldi.get_mel_banks does
// (kaldi.py:436-511) for num_bins=80, n_fft=512, 16 kHz, low 20 Hz, high Nyquist.
std::vector<std::vector<float>> kaldi_bank() {
  const int nb = 80, nfft = 512;
  const float sr = 16000.f, low = 20.f, high = 8000.f;
  const float bw = sr / nfft;
  const float mel_low = 1127.0f * logf(1.0f + low / 700.0f);
  const float mel_high = 1127.0f * logf(1.0f + high / 700.0f);
  const float delta = (mel_high - mel_low) / (nb + 1);
  std::vector<std::vector<float>> bank(nb, std::vector<float>(nfft / 2 + 1, 0.f));
  for (int b = 0; b < nb; ++b) {
    const float left = mel_low + b * delta, center = mel_low + (b + 1.0f) * delta, right = mel_low + (b + 2.0f) * delta;
    for (int k = 0; k < nfft / 2; ++k) {
      const float mel = 1127.0f * logf(1.0f + (bw * k) / 700.0f);
      const float up = (mel - left) / (center - left), down = (right - mel) / (right - center);
      bank[b][k] = fmaxf(0.f, fminf(up, down));
    }
  }
  return bank;
}

// librosa.filters.mel(sr=16000, n_fft=400, n_mels=128): Slaney scale, Slaney area normalisation, float64 maths.
std::vector<std::vector<float>> slaney_bank(int nm = 128, int nfft = 400, double sr = 16000.0) {
  const int nf = nfft / 2 + 1;
  const double f_sp = 200.0 / 3, min_log_hz = 1000.0, min_log_mel = min_log_hz / f_sp;
  const double logstep = log(6.4) / 27.0;
  auto hz2mel = [&](double f) { return f >= min_log_hz ? min_log_mel + log(f / min_log_hz) / logstep : f / f_sp; };
  auto mel2hz = [&](double m) { return m >= min_log_mel ? min_log_hz * exp(logstep * (m - min_log_mel)) : f_sp * m; };
  std::vector<double> hz(nm + 2);
  const double m_lo = hz2mel(0.0), m_hi = hz2mel(sr / 2);
  for (int i = 0; i < nm + 2; ++i) hz[i] = mel2hz(m_lo + (m_hi - m_lo) * i / (nm + 1));
  std::vector<std::vector<float>> bank(nm, std::vector<float>(nf, 0.f));
  for (int i = 0; i < nm; ++i) {
    const double enorm = 2.0 / (hz[i + 2] - hz[i]);
    for (int k = 0; k < nf; ++k) {
      const double f = (sr / 2) * k / (nf - 1);
      const double lower = (f - hz[i]) / (hz[i + 1] - hz[i]);
      const double upper = (hz[i + 2] - f) / (hz[i + 2] - hz[i + 1]);
      const double v = fmax(0.0, fmin(lower, upper)) * enorm;
      bank[i][k] = (float)v;
    }
  }
  return bank;
}

}  // namespace

// Device address of g_fbank_static_ws on each device that has a context.  Runs inside fy_create (frontend_init) with
// `device` current, so the hot call below never asks the runtime for anything.
namespace {
constexpr int kMaxDevices = 64;
void* g_fbank_static_ptr[kMaxDevices] = {};
std::mutex g_fbank_static_mu;
}  // namespace

int fbank_static_setup(int device) {
  FY_CHECK(device >= 0 && device < kMaxDevices, FY_ERR_ARG, "device ordinal %d outside [0, %d)", device, kMaxDevices);
  std::lock_guard<std::mutex> lock(g_fbank_static_mu);
  if (g_fbank_static_ptr[device]) return FY_OK;
  int cur = -1;
  FY_CUDA(cudaGetDevice(&cur));
  FY_CHECK(cur == device, FY_ERR_STATE, "fbank_static_setup: device %d is current, not %d", cur, device);
  void* p = nullptr;
  FY_CUDA(cudaGetSymbolAddress(&p, g_fbank_static_ws));
  FY_CHECK(p != nullptr && (reinterpret_cast<uintptr_t>(p) & 255) == 0, FY_ERR_STATE, "static fbank scratch has no aligned device address");
  g_fbank_static_ptr[device] = p;
  return FY_OK;
}

int frontend_init(Ctx* c) {
  // Povey window = hann(400, periodic=False)^0.85 in fp32 (kaldi.py:98-100)
  std::vector<float> povey(400), hann(400);
  for (int i = 0; i < 400; ++i) {
    const float h = 0.5f - 0.5f * cosf((float)(2.0 * M_PI * i / 399.0));
    povey[i] = powf((float)(0.5 - 0.5 * cos(2.0 * M_PI * i / 399.0)), 0.85f);
    (void)h;
    hann[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * i / 400.0));
  }
  FY_TRY(upload(c, povey, &c->fb_window));
  FY_TRY(upload(c, hann, &c->wh_window));
  FY_TRY(upload(c, twiddles(512), &c->fb_tw));
  FY_TRY(upload(c, twiddles(400), &c->wh_tw));
  FY_TRY(upload_mel(c, kaldi_bank(), &c->fb_mel_w, &c->fb_mel_off, &c->fb_mel_len, &c->fb_mel_ptr, &c->fb_mel_nw));
  FY_TRY(upload_mel(c, slaney_bank(), &c->wh_mel_w, &c->wh_mel_off, &c->wh_mel_len, &c->wh_mel_ptr, &c->wh_mel_nw));
  FY_CHECK(c->fb_mel_nw <= fb2::kMelW, FY_ERR_STATE, "Kaldi mel bank has %d weights, the kernel holds %d", c->fb_mel_nw, fb2::kMelW);
  FY_CUDA(cudaFuncSetAttribute(fb2::fbank2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(fb2::Smem)));
  FY_CUDA(cudaFuncSetAttribute(fb2::fbank2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(fb2::Smem)));
  FY_CHECK(c->wh_mel_nw <= lm2::kMelW, FY_ERR_STATE, "Slaney mel bank has %d weights, the kernel holds %d", c->wh_mel_nw, lm2::kMelW);
  FY_CUDA(cudaFuncSetAttribute(lm2::logmel2_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(lm2::Smem)));
  FY_CUDA(cudaFuncSetAttribute(lm2::logmel2_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(lm2::Smem)));
  // matcha 24 kHz mel (row N4): periodic Hann(1920), 1920-point twiddles, librosa mel(sr 24000, n_fft 1920, 80)
  std::vector<float> hann24(kMmN);
  for (int i = 0; i < kMmN; ++i) hann24[i] = (float)(0.5 - 0.5 * cos(2.0 * M_PI * i / (double)kMmN));
  FY_TRY(upload(c, hann24, &c->mm_window));
  FY_TRY(upload(c, twiddles(kMmN), &c->mm_tw));
  FY_TRY(upload_mel(c, slaney_bank(kMmMels, kMmN, 24000.0), &c->mm_mel_w, &c->mm_mel_off, &c->mm_mel_len, &c->mm_mel_ptr));
  FY_TRY(fbank_static_setup(c->device));
  return FY_OK;
}

// launchers used by the pipeline code as well as the C ABI ------------------------------------------------
int fbank2_slots(int n_utts, int total_frames) { return ((total_frames + 15 * n_utts) >> 4) + 1; }

int launch_fbank(Ctx* c, const void* pcm, int dtype, const int* pcm_off, const int* frame_off, int n_utts,
                 int total_frames, float* feat, void* ws) {
  if (total_frames <= 0) return FY_OK;
  MelTable mt{c->fb_mel_w, c->fb_mel_off, c->fb_mel_len, c->fb_mel_ptr};
  static const bool v1 = [] { const char* e = getenv("FY_FRONTEND_V1"); return e && e[0] == '1'; }();
  if (v1) {
    const int pairs = ((total_frames + n_utts) >> 1) + 1;   // utterance-local pair slots (see find_slot_utt)
    const int grid = (pairs + kWarpsPerBlock - 1) / kWarpsPerBlock;
    if (dtype == FY_PCM_S16)
      fbank_kernel<true><<<grid, kWarpsPerBlock * 32, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                                     c->fb_window, c->fb_tw, mt, feat);
    else
      fbank_kernel<false><<<grid, kWarpsPerBlock * 32, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                                      c->fb_window, c->fb_tw, mt, feat);
    FY_CUDA(cudaGetLastError());
    cmn_kernel<<<n_utts, 320, 0, c->stream>>>(feat, frame_off);
    c->launches += 2;
    FY_CUDA(cudaGetLastError());
    return FY_OK;
  }
  const int slots = fbank2_slots(n_utts, total_frames);
  float* psum = reinterpret_cast<float*>(ws);                 // [slots][80] per-CTA column sums, then [n_utts][80] means
  float* mean = psum + (size_t)slots * 80;
  const int grid = std::min(slots, 3 * c->num_sms);            // persistent: three CTAs per SM walk the slots
  if (dtype == FY_PCM_S16)
    fb2::fbank2_kernel<true><<<grid, fb2::kThreads, sizeof(fb2::Smem), c->stream>>>(pcm, pcm_off, frame_off, n_utts, slots,
                                                                                    c->fb_window, c->fb_tw, mt, c->fb_mel_nw, feat, psum);
  else
    fb2::fbank2_kernel<false><<<grid, fb2::kThreads, sizeof(fb2::Smem), c->stream>>>(pcm, pcm_off, frame_off, n_utts, slots,
                                                                                     c->fb_window, c->fb_tw, mt, c->fb_mel_nw, feat, psum);
  fb2::cmn_mean_kernel<<<n_utts, 96, 0, c->stream>>>(psum, frame_off, mean);
  fb2::cmn_sub_kernel<<<slots, fb2::kThreads, 0, c->stream>>>(feat, frame_off, n_utts, mean);
  c->launches += 3;
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

// raw: [total_frames, 128] fp32 time-major log10 mel; umax: [n_utts] int keys
int launch_logmel_raw(Ctx* c, const void* pcm, int dtype, const int* pcm_off, const int* frame_off, int n_utts,
                      int total_frames, float* raw, int* umax) {
  if (total_frames <= 0) return FY_OK;
  MelTable mt{c->wh_mel_w, c->wh_mel_off, c->wh_mel_len, c->wh_mel_ptr};
  fill_int_kernel<<<(n_utts + 255) / 256, 256, 0, c->stream>>>(umax, n_utts, (int)0x80000000);
  const int pairs = ((total_frames + n_utts) >> 1) + 1;   // utterance-local pair slots (see find_slot_utt)
  const int grid = (pairs + kWarpsPerBlock - 1) / kWarpsPerBlock;
  if (dtype == FY_PCM_S16)
    logmel_kernel<true><<<grid, kWarpsPerBlock * 32, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                                    c->wh_window, c->wh_tw, mt, raw, umax);
  else
    logmel_kernel<false><<<grid, kWarpsPerBlock * 32, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                                     c->wh_window, c->wh_tw, mt, raw, umax);
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

}  // namespace fy

using namespace fy;

extern "C" {

size_t fy_fbank_workspace_bytes(int n_utts, int total_frames) {
  if (n_utts <= 0 || total_frames < 0) return 256;
  return ((size_t)fbank2_slots(n_utts, total_frames) + (size_t)n_utts) * 80 * sizeof(float) + 256;
}

int fy_fbank_cmn(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                 int n_utts, int total_frames, float* feat, void* ws, size_t ws_bytes) {
  FY_CHECK(ctx && pcm && pcm_off && frame_off && feat && ws, FY_ERR_ARG, "fy_fbank_cmn: NULL argument");
  FY_CHECK(ws_bytes >= fy_fbank_workspace_bytes(n_utts, total_frames), FY_ERR_NOMEM, "fy_fbank_cmn: workspace too small");
  FY_CHECK((reinterpret_cast<uintptr_t>(ws) & 15) == 0, FY_ERR_ARG, "fy_fbank_cmn: workspace must be 16-byte aligned");
  FY_CHECK(n_utts > 0 && total_frames >= 0, FY_ERR_ARG, "fy_fbank_cmn: bad sizes");
  FY_CHECK(pcm_dtype == FY_PCM_S16 || pcm_dtype == FY_PCM_F32, FY_ERR_ARG, "fy_fbank_cmn: bad pcm dtype");
  return launch_fbank(reinterpret_cast<Ctx*>(ctx), pcm, pcm_dtype, pcm_off, frame_off, n_utts, total_frames, feat, ws);
}

// fy_fbank_cmn for callers that cannot allocate (CUDA graph capture, serving loops, harnesses that forbid cudaMalloc):
// same kernels, same launch order, the scratch is g_fbank_static_ws instead of a caller workspace.  No cudaMalloc /
// cudaFree / allocating library call on this path; three launches on the context's stream.  Nothing has to be cleared
// between calls: every scratch element a call reads (slot sums, utterance means) is written earlier in the same call.
// All calls on a device share the array, so two calls that overlap on the GPU (different streams or contexts) race.
int fy_fbank_cmn_static(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                        int n_utts, int total_frames, float* feat) {
  FY_CHECK(ctx && pcm && pcm_off && frame_off && feat, FY_ERR_ARG, "fy_fbank_cmn_static: NULL argument");
  FY_CHECK(n_utts > 0 && total_frames >= 0, FY_ERR_ARG, "fy_fbank_cmn_static: bad sizes");
  FY_CHECK(pcm_dtype == FY_PCM_S16 || pcm_dtype == FY_PCM_F32, FY_ERR_ARG, "fy_fbank_cmn_static: bad pcm dtype");
  const size_t need = fy_fbank_workspace_bytes(n_utts, total_frames);
  FY_CHECK(need <= (size_t)FY_FBANK_STATIC_WS_BYTES, FY_ERR_NOMEM,
           "fy_fbank_cmn_static: %d utterances / %d frames need %zu bytes of scratch, the static array holds %zu "
           "(total_frames / 16 + 2 n_utts + 1 <= %zu rows); use fy_fbank_cmn with a caller workspace",
           n_utts, total_frames, need, (size_t)FY_FBANK_STATIC_WS_BYTES, ((size_t)FY_FBANK_STATIC_WS_BYTES - 256) / 320);
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  void* ws = (c->device >= 0 && c->device < kMaxDevices) ? g_fbank_static_ptr[c->device] : nullptr;
  FY_CHECK(ws != nullptr, FY_ERR_STATE, "fy_fbank_cmn_static: no scratch address for device %d (fy_create sets it up)", c->device);
  return launch_fbank(c, pcm, pcm_dtype, pcm_off, frame_off, n_utts, total_frames, feat, ws);
}

size_t fy_logmel_workspace_bytes(int n_utts, int total_frames) {
  return (size_t)total_frames * 128 * 4 + (size_t)n_utts * 4 + 256;
}

int fy_logmel(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
              int n_utts, int total_frames, int max_frames, float* mel, void* ws, size_t ws_bytes) {
  FY_CHECK(ctx && pcm && pcm_off && frame_off && mel && ws, FY_ERR_ARG, "fy_logmel: NULL argument");
  FY_CHECK(n_utts > 0 && n_utts <= 65535 && total_frames >= 0 && max_frames <= total_frames, FY_ERR_ARG,
           "fy_logmel: bad sizes (n_utts=%d total_frames=%d max_frames=%d)", n_utts, total_frames, max_frames);
  FY_CHECK(pcm_dtype == FY_PCM_S16 || pcm_dtype == FY_PCM_F32, FY_ERR_ARG, "fy_logmel: bad pcm dtype");
  FY_CHECK(ws_bytes >= fy_logmel_workspace_bytes(n_utts, total_frames), FY_ERR_NOMEM, "fy_logmel: workspace too small");
  if (total_frames == 0 || max_frames <= 0) return FY_OK;
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  float* raw = reinterpret_cast<float*>(ws);
  int* umax = reinterpret_cast<int*>(raw + (size_t)total_frames * 128);
  // The second-generation log-mel kernel (lm2, FY_LOGMEL_V2=1) is parity-green but measured SLOWER than the first
  // (2.47 against 2.11 ms at 1,024 x 10 s): with the same shared-memory 400-point transform its CTA-wide staging and
  // barriers cost more than they save, so the first generation stays the default here.
  static const bool v2 = [] { const char* e = getenv("FY_LOGMEL_V2"); return e && e[0] == '1'; }();
  if (v2) {
    MelTable mt{c->wh_mel_w, c->wh_mel_off, c->wh_mel_len, c->wh_mel_ptr};
    const int slots = fbank2_slots(n_utts, total_frames);
    fill_int_kernel<<<(n_utts + 255) / 256, 256, 0, c->stream>>>(umax, n_utts, (int)0x80000000);
    const int grid = std::min(slots, 3 * c->num_sms);          // persistent: three CTAs per SM walk the slots
    if (pcm_dtype == FY_PCM_S16)
      lm2::logmel2_kernel<true><<<grid, lm2::kThreads, sizeof(lm2::Smem), c->stream>>>(pcm, pcm_off, frame_off, n_utts, slots,
                                                                                       c->wh_window, c->wh_tw, mt, c->wh_mel_nw, mel, umax);
    else
      lm2::logmel2_kernel<false><<<grid, lm2::kThreads, sizeof(lm2::Smem), c->stream>>>(pcm, pcm_off, frame_off, n_utts, slots,
                                                                                        c->wh_window, c->wh_tw, mt, c->wh_mel_nw, mel, umax);
    lm2::logmel2_finish_kernel<<<slots, lm2::kThreads, 0, c->stream>>>(mel, frame_off, n_utts, umax);
    c->launches += 3;
    FY_CUDA(cudaGetLastError());
    return FY_OK;
  }
  FY_TRY(launch_logmel_raw(c, pcm, pcm_dtype, pcm_off, frame_off, n_utts, total_frames, raw, umax));
  dim3 grid((max_frames + 31) / 32, n_utts);
  logmel_finish_kernel<<<grid, 256, 0, c->stream>>>(raw, frame_off, umax, mel);
  c->launches += 3;
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

int fy_matcha_mel(fy_ctx* ctx, const void* pcm, int pcm_dtype, const int32_t* pcm_off, const int32_t* frame_off,
                  int n_utts, int total_frames, float* mel) {
  FY_CHECK(ctx && pcm && pcm_off && frame_off && mel, FY_ERR_ARG, "fy_matcha_mel: NULL argument");
  FY_CHECK(n_utts > 0 && total_frames >= 0, FY_ERR_ARG, "fy_matcha_mel: bad sizes");
  FY_CHECK(pcm_dtype == FY_PCM_S16 || pcm_dtype == FY_PCM_F32, FY_ERR_ARG, "fy_matcha_mel: bad pcm dtype");
  if (total_frames == 0) return FY_OK;
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  MelTable mt{c->mm_mel_w, c->mm_mel_off, c->mm_mel_len, c->mm_mel_ptr};
  const int pairs = ((total_frames + n_utts) >> 1) + 1;
  if (pcm_dtype == FY_PCM_S16)
    matcha_mel_kernel<true><<<pairs, kMmThreads, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                               c->mm_window, c->mm_tw, mt, mel);
  else
    matcha_mel_kernel<false><<<pairs, kMmThreads, 0, c->stream>>>(pcm, pcm_off, frame_off, n_utts, total_frames,
                                                                c->mm_window, c->mm_tw, mt, mel);
  c->launches += 1;
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

namespace {
struct ResampleSpec { int orig, newf, width, K; double base; };
ResampleSpec resample_spec(int orig_freq, int new_freq) {
  int a = orig_freq, b = new_freq;
  while (b) { int t = a % b; a = b; b = t; }
  ResampleSpec r;
  r.orig = orig_freq / a; r.newf = new_freq / a;
  const double lpw = 6.0, rolloff = 0.99;
  r.base = (double)(r.orig < r.newf ? r.orig : r.newf) * rolloff;
  r.width = (int)ceil(lpw * r.orig / r.base);
  r.K = 2 * r.width + r.orig;
  return r;
}
}  // namespace

// torchaudio _get_sinc_resample_kernel (functional.py:1305-1402): hann-windowed sinc, width 6, rolloff 0.99, built in
// float64 and rounded to fp32 -- except the phase offset -p/new, which torchaudio computes as a float32 division (its
// arange there is int64), a quirk reproduced below.  Built once per reduced (orig, new) pair and kept by the context:
// this is the call that allocates and copies synchronously, so that fy_resample itself never does.
int fy_resample_prepare(fy_ctx* ctx, int orig_freq, int new_freq) {
  FY_CHECK(ctx && orig_freq > 0 && new_freq > 0 && orig_freq != new_freq, FY_ERR_ARG, "fy_resample_prepare: bad arguments (%d -> %d)",
           orig_freq, new_freq);
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  const ResampleSpec r = resample_spec(orig_freq, new_freq);
  auto key = std::make_pair(r.orig, r.newf);
  if (c->resample_kern.count(key)) return FY_OK;
  const double lpw = 6.0;
  std::vector<float> h((size_t)r.newf * r.K);
  for (int p = 0; p < r.newf; ++p)
    for (int k = 0; k < r.K; ++k) {
      const float phase = (float)(-p) / (float)r.newf;
      double t = ((double)phase + (double)(k - r.width) / r.orig) * r.base;
      if (t < -lpw) t = -lpw;
      if (t > lpw) t = lpw;
      const double cw = cos(t * M_PI / lpw / 2);
      const double tp = t * M_PI;
      const double sinc = tp == 0.0 ? 1.0 : sin(tp) / tp;
      h[(size_t)p * r.K + k] = (float)(sinc * cw * cw * (r.base / r.orig));
    }
  float* dk = nullptr;
  FY_CUDA(cudaSetDevice(c->device));
  FY_TRY(upload(c, h, &dk));
  c->resample_kern[key] = dk;
  return FY_OK;
}

int fy_resample(fy_ctx* ctx, const float* in, const int32_t* in_off, float* out, const int32_t* out_off, int n_utts,
                int total_out, int orig_freq, int new_freq) {
  FY_CHECK(ctx && in && in_off && out && out_off, FY_ERR_ARG, "fy_resample: NULL argument");
  FY_CHECK(n_utts > 0 && orig_freq > 0 && new_freq > 0 && orig_freq != new_freq, FY_ERR_ARG,
           "fy_resample: bad arguments (n_utts=%d, %d -> %d)", n_utts, orig_freq, new_freq);
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  const ResampleSpec r = resample_spec(orig_freq, new_freq);
  auto it = c->resample_kern.find(std::make_pair(r.orig, r.newf));
  FY_CHECK(it != c->resample_kern.end(), FY_ERR_STATE,
           "fy_resample: call fy_resample_prepare(ctx, %d, %d) once first (it builds the filter bank; launches never allocate)",
           orig_freq, new_freq);
  if (total_out <= 0) return FY_OK;
  resample_kernel<<<(total_out + 255) / 256, 256, 0, c->stream>>>(in, in_off, out, out_off, n_utts, total_out, it->second,
                                                                r.width, r.orig, r.newf);
  c->launches += 1;
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

}  // extern "C"
