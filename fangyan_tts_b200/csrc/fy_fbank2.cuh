// Kaldi fbank + CMN, second generation (rows A3 + A4).  Included by fy_frontend.cu inside its anonymous namespace
// (uses its cmul / mul_negi / warp_sum / MelTable helpers).
//
// What changed against fbank_kernel (kept as FY_FRONTEND_V1=1): that kernel fetched every PCM sample 2.5 times with
// 2-byte loads, walked five shared-memory Stockham stages per transform, read the mel weights from global memory and
// left CMN to a kernel that re-read the whole feature matrix twice.  Here
//   * one CTA owns 16 consecutive frames of ONE utterance: their 2,800 samples are staged into shared memory once
//     (coalesced) and every frame is cut from that tile;
//   * a warp transforms one frame pair (real frame in the real lane, its successor in the imaginary lane; pairs are
//     utterance-local, so results do not depend on the batch) as a 512-point FFT = 8 x 8 x 8 held in REGISTERS:
//     sixteen points per lane, three radix-8 passes, two shared-memory exchanges (+ one to meet X[k] with X[512 - k]);
//   * window, twiddles and the sparse mel bank sit in shared memory;
//   * the CTA leaves the per-bin sum of its frames, so CMN needs no pass that reads the features to average them:
//     cmn_mean_kernel adds up an utterance's few partial rows in a fixed order, cmn_sub_kernel subtracts in place.
#pragma once

namespace fb2 {

constexpr int kFrames = 16, kWarps = 8, kThreads = 32 * kWarps;
constexpr int kSpan = 400 + 160 * (kFrames - 1);       // samples one CTA touches
constexpr int kXchg = 64 * 9;                            // float2 per warp: 64 rows of 8 (+1 pad), also holds 512 spectrum values
constexpr int kPStride = 260;                            // floats per power spectrum (257 used)
constexpr int kMelW = 512;                               // packed mel weights (501 non-zeros for the Kaldi bank)

struct Smem {
  float pcm[kSpan + 8];
  float win[400];
  float2 tw[512];
  float melw[kMelW];
  int mel_off[80], mel_len[80], mel_ptr[80];
  float2 x[kWarps][kXchg];        // per warp: exchange buffer, then the spectrum, then the two power spectra
};
static_assert(kFrames * 80 <= kSpan, "the CTA's output tile reuses the PCM tile");
static_assert(2 * kPStride <= 2 * kXchg, "power spectra fit the exchange buffer");
static_assert(sizeof(Smem) <= 58 * 1024, "three CTAs per SM with room to spare");

// Utterance u owns 16-frame slots [s(u), s(u+1)), s(u) = (frame_off[u] + 15 u) >> 4: always at least ceil(M_u / 16)
// slots, and a closed form of the offsets, so a CTA finds its utterance by binary search without any extra table.
__device__ __forceinline__ int slot_of(const int* __restrict__ off, int u) { return (__ldg(off + u) + 15 * u) >> 4; }
__device__ __forceinline__ int find_slot16_utt(const int* __restrict__ off, int n, int w) {
  int lo = 0, hi = n;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (slot_of(off, mid) <= w) lo = mid; else hi = mid;
  }
  return lo;
}

// In-place 8-point DFT, forward sign: a[k] <- sum_n a[n] exp(-2 pi i n k / 8).
__device__ __forceinline__ void dft8(float2 (&a)[8]) {
  const float h = 0.70710678118654752f;
  const float2 b0 = cadd(a[0], a[4]), b1 = csub(a[0], a[4]), b2 = cadd(a[2], a[6]), b3 = mul_negi(csub(a[2], a[6]));
  const float2 b4 = cadd(a[1], a[5]), b5 = csub(a[1], a[5]), b6 = cadd(a[3], a[7]), b7 = mul_negi(csub(a[3], a[7]));
  const float2 e0 = cadd(b0, b2), e2 = csub(b0, b2), e1 = cadd(b1, b3), e3 = csub(b1, b3);
  const float2 o0 = cadd(b4, b6), o2 = mul_negi(csub(b4, b6)), t1 = cadd(b5, b7), t3 = csub(b5, b7);
  const float2 o1 = make_float2(h * (t1.x + t1.y), h * (t1.y - t1.x));      // t1 * (1 - i) / sqrt 2
  const float2 o3 = make_float2(h * (t3.y - t3.x), -h * (t3.x + t3.y));     // t3 * (-1 - i) / sqrt 2
  a[0] = cadd(e0, o0); a[4] = csub(e0, o0);
  a[1] = cadd(e1, o1); a[5] = csub(e1, o1);
  a[2] = cadd(e2, o2); a[6] = csub(e2, o2);
  a[3] = cadd(e3, o3); a[7] = csub(e3, o3);
}

// 512-point FFT of the 16 points z[j] = x[lane + 32 j] held by each lane of a warp; returns with the spectrum in
// natural order in xs[0 .. 511].  xs: the warp's exchange buffer (kXchg float2).
__device__ __forceinline__ void fft512_regs(float2 (&z)[16], float2* __restrict__ xs, const float2* __restrict__ tw, int lane) {
  // pass 1: p = n0 + 8 n1 in {lane, lane + 32}; transform over n2 (points p + 64 n2), twiddle W64^(n1 k2)
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    float2 a[8];
#pragma unroll
    for (int n2 = 0; n2 < 8; ++n2) a[n2] = z[2 * n2 + h];
    dft8(a);
    const int p = lane + 32 * h, n1 = p >> 3;
    xs[p * 9] = a[0];
#pragma unroll
    for (int k2 = 1; k2 < 8; ++k2) xs[p * 9 + k2] = cmul(a[k2], tw[8 * n1 * k2]);
  }
  __syncwarp();
  // pass 2: (n0, k2) = (c & 7, c >> 3) for c in {lane, lane + 32}; transform over n1, twiddle W512^(n0 (k2 + 8 k1))
  float2 b[2][8];
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int c = lane + 32 * h, n0 = c & 7, k2 = c >> 3;
#pragma unroll
    for (int n1 = 0; n1 < 8; ++n1) b[h][n1] = xs[(n0 + 8 * n1) * 9 + k2];
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int c = lane + 32 * h, n0 = c & 7, k2 = c >> 3;
    dft8(b[h]);
#pragma unroll
    for (int k1 = 0; k1 < 8; ++k1)
      xs[(k1 * 8 + k2) * 9 + n0] = (n0 == 0) ? b[h][k1] : cmul(b[h][k1], tw[n0 * (k2 + 8 * k1)]);
  }
  __syncwarp();
  // pass 3: q = 8 k1 + k2 in {lane, lane + 32}; transform over n0; X[64 k0 + q]
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int q = lane + 32 * h;
#pragma unroll
    for (int n0 = 0; n0 < 8; ++n0) b[h][n0] = xs[q * 9 + n0];
  }
  __syncwarp();
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int q = lane + 32 * h;
    dft8(b[h]);
#pragma unroll
    for (int k0 = 0; k0 < 8; ++k0) xs[64 * k0 + q] = b[h][k0];
  }
  __syncwarp();
}

// Persistent: gridDim.x CTAs walk the slots with stride gridDim.x; the tables are staged once per CTA and the next
// slot's samples are already in flight (registers) while the current slot is transformed.
constexpr int kPerThread = (kSpan + kThreads - 1) / kThreads;   // samples a thread stages per slot

template <bool S16>
__global__ void __launch_bounds__(kThreads, 3) fbank2_kernel(const void* __restrict__ pcm, const int* __restrict__ pcm_off,
                                                          const int* __restrict__ frame_off, int n_utts, int n_slots,
                                                          const float* __restrict__ window, const float2* __restrict__ tw,
                                                          MelTable mt, int mel_nw, float* __restrict__ feat,
                                                          float* __restrict__ psum) {
  extern __shared__ uint8_t fb2_raw[];
  Smem& S = *reinterpret_cast<Smem*>(fb2_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 512; i += kThreads) S.tw[i] = __ldg(tw + i);
  for (int i = tid; i < mel_nw; i += kThreads) S.melw[i] = __ldg(mt.w + i);
  if (tid < 80) { S.mel_off[tid] = __ldg(mt.off + tid); S.mel_len[tid] = __ldg(mt.len + tid); S.mel_ptr[tid] = __ldg(mt.ptr + tid); }
  for (int i = tid; i < 400; i += kThreads) S.win[i] = __ldg(window + i);

  struct Where { int u, fo, f0, nf, base; };
  auto locate = [&](int slot) {
    Where w;
    w.u = find_slot16_utt(frame_off, n_utts, slot);
    w.fo = __ldg(frame_off + w.u);
    w.f0 = kFrames * (slot - slot_of(frame_off, w.u));            // first frame of the slot inside utterance u
    w.nf = __ldg(frame_off + w.u + 1) - w.fo;
    w.base = __ldg(pcm_off + w.u) + w.f0 * 160;
    return w;
  };
  float nx[kPerThread];
  auto prefetch = [&](const Where& w) {
    const int have = w.f0 < w.nf ? 400 + 160 * (min(kFrames, w.nf - w.f0) - 1) : 0;   // all inside the utterance
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
      const int i = tid + k * kThreads;
      nx[k] = i < have ? load_pcm<S16>(pcm, w.base + i) : 0.f;
    }
  };
  int slot = blockIdx.x;
  Where cur{0, 0, 0, 0, 0};
  if (slot < n_slots) { cur = locate(slot); prefetch(cur); }
  for (; slot < n_slots; slot += gridDim.x) {
    float* my_psum = psum + (size_t)slot * 80;
    const int fo = cur.fo, f0 = cur.f0, nf = cur.nf;
    const bool padding = f0 >= nf;                               // padding slot: contributes nothing to the mean
    const int nfr = padding ? 0 : min(kFrames, nf - f0);         // frames of this slot
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
      const int i = tid + k * kThreads;
      if (i < kSpan) S.pcm[i] = nx[k];
    }
    __syncthreads();
    if (slot + (int)gridDim.x < n_slots) { cur = locate(slot + gridDim.x); prefetch(cur); }   // in flight under the transform
    const int fa = 2 * warp;                                     // this warp's frame pair inside the slot: fa, fa + 1
    const bool active = fa < nfr;
    const bool two = fa + 1 < nfr;
    float2 z[16];
    if (active) {
      const float* xa = S.pcm + fa * 160;
      const float* xb = xa + 160;
      // frame means (DC removal), kaldi.py:183-186
      float sa = 0.f, sb = 0.f;
#pragma unroll
      for (int j = 0; j < 13; ++j) {
        const int n = lane + 32 * j;
        if (n < 400) { sa += xa[n]; sb += xb[n]; }
      }
      const float ma = warp_sum(sa) * (1.0f / 400.0f), mb = warp_sum(sb) * (1.0f / 400.0f);
      // pre-emphasis 0.97 with replicate-left, Povey window, zero pad to 512 (kaldi.py:193-217): z[j] = x[lane + 32 j]
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int n = lane + 32 * j;
        float2 v = make_float2(0.f, 0.f);
        if (n < 400) {
          const int np = n > 0 ? n - 1 : 0;
          const float w = S.win[n];
          v.x = ((xa[n] - ma) - 0.97f * (xa[np] - ma)) * w;
          if (two) v.y = ((xb[n] - mb) - 0.97f * (xb[np] - mb)) * w;
        }
        z[j] = v;
      }
    }
    __syncthreads();                                             // every frame has been cut: the PCM tile becomes the output tile
    float (*sfeat)[80] = reinterpret_cast<float (*)[80]>(S.pcm);
    if (active) {
      float2* xs = S.x[warp];
      fft512_regs(z, xs, S.tw, lane);
      // power spectra of the two real frames from the packed transform, written over the spectrum they come from
      float pa[9], pb[9];
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        const int k = min(lane + 32 * i, 256);
        const float2 a = xs[k], c = xs[(512 - k) & 511];
        const float ar = 0.5f * (a.x + c.x), ai = 0.5f * (a.y - c.y);
        const float br = 0.5f * (a.y + c.y), bi = -0.5f * (a.x - c.x);
        pa[i] = ar * ar + ai * ai;
        pb[i] = br * br + bi * bi;
      }
      __syncwarp();
      float* P = reinterpret_cast<float*>(xs);
#pragma unroll
      for (int i = 0; i < 9; ++i) {
        const int k = lane + 32 * i;
        if (k <= 256) { P[k] = pa[i]; P[kPStride + k] = pb[i]; }
      }
      __syncwarp();
      // sparse mel bank + log (kaldi.py:621-633): 160 outputs per warp
#pragma unroll
      for (int i = 0; i < 5; ++i) {
        const int o = lane + 32 * i;
        const int f = o >= 80 ? 1 : 0, b = o - 80 * f;
        const int off = S.mel_off[b], n = S.mel_len[b];
        const float* w = S.melw + S.mel_ptr[b];
        const float* p = P + f * kPStride + off;
        float acc = 0.f;
        for (int j = 0; j < n; ++j) acc += w[j] * p[j];
        const float v = logf(fmaxf(acc, 1.1920929e-07f));
        const bool ok = f == 0 || two;
        sfeat[fa + f][b] = ok ? v : 0.f;
        if (ok) feat[(size_t)(fo + f0 + fa + f) * 80 + b] = v;
      }
      __syncwarp();
    } else {
      for (int i = lane; i < 160; i += 32) sfeat[fa + (i >= 80)][i % 80] = 0.f;
    }
    __syncthreads();
    if (tid < 80) {
      float a = 0.f;
#pragma unroll
      for (int f = 0; f < kFrames; ++f) a += sfeat[f][tid];
      my_psum[tid] = a;
    }
    __syncthreads();                                             // the tile is free for the next slot's samples
  }
}

// mean[u][b] = (sum of the utterance's partial rows, in slot order) / M_u.  One block of 96 threads per utterance.
__global__ void __launch_bounds__(96) cmn_mean_kernel(const float* __restrict__ psum, const int* __restrict__ frame_off,
                                                      float* __restrict__ mean) {
  const int u = blockIdx.x, b = threadIdx.x;
  if (b >= 80) return;
  const int s0 = slot_of(frame_off, u);
  const int nf = __ldg(frame_off + u + 1) - __ldg(frame_off + u);
  const int ns = (nf + kFrames - 1) / kFrames;
  float a = 0.f;
  for (int s = 0; s < ns; ++s) a += psum[(size_t)(s0 + s) * 80 + b];
  mean[(size_t)u * 80 + b] = nf > 0 ? a / (float)nf : 0.f;
}

// feat[f][b] -= mean[u][b] over the CTA's 16 frames (same slot -> utterance mapping as fbank2_kernel).
__global__ void __launch_bounds__(kThreads) cmn_sub_kernel(float* __restrict__ feat, const int* __restrict__ frame_off, int n_utts,
                                                           const float* __restrict__ mean) {
  const int slot = blockIdx.x;
  const int u = find_slot16_utt(frame_off, n_utts, slot);
  const int fo = __ldg(frame_off + u);
  const int f0 = kFrames * (slot - slot_of(frame_off, u));
  const int nf = __ldg(frame_off + u + 1) - fo;
  if (f0 >= nf) return;
  const int n = min(kFrames, nf - f0) * 80;
  float* p = feat + (size_t)(fo + f0) * 80;
  const float* m = mean + (size_t)u * 80;
  for (int i = threadIdx.x; i < n; i += kThreads) p[i] -= __ldg(m + i % 80);
}

}  // namespace fb2

// ------------------------------------------------------------------------------------------------------------------
// Whisper log-mel, second generation (row A5).  Same CTA shape as fbank2 (16 frames of one utterance, PCM staged once
// with the reflect padding of torch.stft(center=True) applied while staging, window / twiddles / Slaney bank in shared
// memory, one frame pair per warp).  The 400-point transform is the shared-memory Stockham of the first generation
// (400 = 4 * 4 * 5 * 5 does not split into equal register-resident passes the way 512 does).  log10(mel) goes straight
// into the reference's [128][T] layout through a shared-memory tile -- 16 consecutive frames of a bin are one 64-byte
// run -- and the CTA's maximum into the utterance's atomicMax (exact whatever the order); what is left for the second
// kernel is the in-place clamp max(x, utterance max - 8), (x + 4) / 4: the values move twice, not three times, and
// nothing is transposed on the way.
namespace lm2 {

constexpr int kFrames = 16, kWarps = 8, kThreads = 32 * kWarps;
constexpr int kSpan = 400 + 160 * (kFrames - 1);
constexpr int kMelW = 1024;
constexpr int kTile = kFrames + 1;                       // row stride of the [128][16] output tile (bank-conflict pad)

struct Smem {
  float pcm[kSpan + 8];                                  // reused as the [128][kTile] output tile (2,176 floats)
  float win[400];
  float2 tw[400];
  float melw[kMelW];
  int mel_off[128], mel_len[128], mel_ptr[128];
  float2 x[kWarps][2][spad(400)];
  float wmax[kWarps];
};
static_assert(128 * kTile <= kSpan + 8, "the output tile reuses the PCM tile");
static_assert(sizeof(Smem) <= 75 * 1024, "three CTAs per SM");

// Persistent like fbank2_kernel: tables staged once per CTA, the next slot's samples (reflect padding resolved while
// loading) in flight in registers under the current slot's transforms.
constexpr int kPerThread = (kSpan + kThreads - 1) / kThreads;

template <bool S16>
__global__ void __launch_bounds__(kThreads, 3) logmel2_kernel(const void* __restrict__ pcm, const int* __restrict__ pcm_off,
                                                              const int* __restrict__ frame_off, int n_utts, int n_slots,
                                                              const float* __restrict__ window, const float2* __restrict__ tw,
                                                              MelTable mt, int mel_nw, float* __restrict__ mel,
                                                              int* __restrict__ umax) {
  extern __shared__ uint8_t lm2_raw[];
  Smem& S = *reinterpret_cast<Smem*>(lm2_raw);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < 400; i += kThreads) { S.win[i] = __ldg(window + i); S.tw[i] = __ldg(tw + i); }
  for (int i = tid; i < mel_nw; i += kThreads) S.melw[i] = __ldg(mt.w + i);
  if (tid < 128) { S.mel_off[tid] = __ldg(mt.off + tid); S.mel_len[tid] = __ldg(mt.len + tid); S.mel_ptr[tid] = __ldg(mt.ptr + tid); }

  struct Where { int u, fo, f0, T, base, ns; };
  auto locate = [&](int slot) {
    Where w;
    w.u = fb2::find_slot16_utt(frame_off, n_utts, slot);
    w.fo = __ldg(frame_off + w.u);
    w.f0 = kFrames * (slot - fb2::slot_of(frame_off, w.u));
    w.T = __ldg(frame_off + w.u + 1) - w.fo;
    w.base = __ldg(pcm_off + w.u);
    w.ns = __ldg(pcm_off + w.u + 1) - w.base;
    return w;
  };
  float nx[kPerThread];
  auto prefetch = [&](const Where& w) {
    const int have = w.f0 < w.T ? 400 + 160 * (min(kFrames, w.T - w.f0) - 1) : 0;
    const int first = w.f0 * 160 - 200;                  // sample index of tile position 0 (may be negative: reflect)
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
      const int i = tid + k * kThreads;
      float v = 0.f;
      if (i < have) {
        int idx = first + i;
        if (idx < 0) idx = -idx;
        if (idx >= w.ns) idx = 2 * (w.ns - 1) - idx;
        idx = max(0, min(idx, w.ns - 1));
        v = load_pcm<S16>(pcm, w.base + idx);
      }
      nx[k] = v;
    }
  };
  int slot = blockIdx.x;
  Where cur{0, 0, 0, 0, 0, 0};
  if (slot < n_slots) { cur = locate(slot); prefetch(cur); }
  for (; slot < n_slots; slot += gridDim.x) {
    const int u = cur.u, fo = cur.fo, f0 = cur.f0, T = cur.T;
    const int nfr = f0 < T ? min(kFrames, T - f0) : 0;   // 0: padding slot
#pragma unroll
    for (int k = 0; k < kPerThread; ++k) {
      const int i = tid + k * kThreads;
      if (i < kSpan) S.pcm[i] = nx[k];
    }
    __syncthreads();
    if (slot + (int)gridDim.x < n_slots) { cur = locate(slot + gridDim.x); prefetch(cur); }
    const int fa = 2 * warp;
    const bool active = fa < nfr, two = fa + 1 < nfr;
    float2* A = S.x[warp][0];
    float2* B = S.x[warp][1];
    if (active) {
      const float* xa = S.pcm + fa * 160;
      const float* xb = xa + 160;
      for (int i = lane; i < 400; i += 32) {
        const float w = S.win[i];
        A[sidx(i)] = make_float2(xa[i] * w, two ? xb[i] * w : 0.f);
      }
      __syncwarp();
    }
    __syncthreads();                                     // every frame has been cut: the PCM tile becomes the output tile
    float* tile = S.pcm;
    float mx = -INFINITY;
    if (active) {
      float2* Z = fft400(A, B, S.tw, lane);
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
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int b = lane + 32 * i;
        const int o = S.mel_off[b], n = S.mel_len[b];
        const float* w = S.melw + S.mel_ptr[b];
        float a0 = 0.f, a1 = 0.f;
        for (int j = 0; j < n; ++j) {
          const float wj = w[j];
          a0 += wj * P[o + j];
          a1 += wj * P[201 + o + j];
        }
        const float l0 = log10f(fmaxf(a0, 1e-10f)), l1 = log10f(fmaxf(a1, 1e-10f));
        tile[b * kTile + fa] = l0;
        mx = fmaxf(mx, l0);
        if (two) { tile[b * kTile + fa + 1] = l1; mx = fmaxf(mx, l1); }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    }
    if (lane == 0) S.wmax[warp] = mx;
    __syncthreads();
    if (nfr > 0) {
      if (tid == 0) {
        float m = S.wmax[0];
#pragma unroll
        for (int w = 1; w < kWarps; ++w) m = fmaxf(m, S.wmax[w]);
        atomicMax(umax + u, float_key(m));
      }
      // tile -> mel[128 * fo + b * T + f0 + t]: 16 threads per bin row, 16 rows per pass
      float* out = mel + (size_t)fo * 128 + f0;
      for (int i = tid; i < 128 * kFrames; i += kThreads) {
        const int b = i >> 4, t = i & 15;
        if (t < nfr) out[(size_t)b * T + t] = tile[b * kTile + t];
      }
    }
    __syncthreads();                                     // the tile is free for the next slot's samples
  }
}

// In place: mel = (max(mel, utterance max - 8) + 4) / 4 over the CTA's 16 frames x 128 bins.
__global__ void __launch_bounds__(kThreads) logmel2_finish_kernel(float* __restrict__ mel, const int* __restrict__ frame_off,
                                                                  int n_utts, const int* __restrict__ umax) {
  const int slot = blockIdx.x;
  const int u = fb2::find_slot16_utt(frame_off, n_utts, slot);
  const int fo = __ldg(frame_off + u);
  const int f0 = kFrames * (slot - fb2::slot_of(frame_off, u));
  const int T = __ldg(frame_off + u + 1) - fo;
  if (f0 >= T) return;
  const int nfr = min(kFrames, T - f0);
  const float floor_v = key_float(__ldg(umax + u)) - 8.0f;
  float* out = mel + (size_t)fo * 128 + f0;
  for (int i = threadIdx.x; i < 128 * kFrames; i += kThreads) {
    const int b = i >> 4, t = i & 15;
    if (t < nfr) {
      float* p = out + (size_t)b * T + t;
      *p = (fmaxf(*p, floor_v) + 4.0f) * 0.25f;
    }
  }
}

}  // namespace lm2
