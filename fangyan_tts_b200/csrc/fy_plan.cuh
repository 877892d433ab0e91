// Per-call layout planning ON THE DEVICE.
//
// Both networks run over a packed batch whose row layout (where each utterance starts, which rows are guard rows, the
// attention / pooling tile lists) depends on the utterance lengths.  The host only needs the TOTALS (grid sizes,
// workspace check), which it derives from its own copy of the offsets; the per-row maps are built here from the device
// copy by two tiny kernels -- a one-block scan over the utterances and a binary-search fill over the rows -- so a launch
// call neither stages metadata through pinned memory nor waits for an earlier call's copy: nothing in fy_s3tok_encode_fsq
// / fy_campplus_forward allocates or synchronises, and the whole sequence can be captured in a CUDA graph.
#pragma once
#include <cuda_runtime.h>

namespace fy {
namespace plan {

constexpr int kScanThreads = 1024;

// Exclusive scan of three ints across a 1024-thread block (warp shuffles + one shared-memory hop).
__device__ __forceinline__ void block_excl_scan3(int& a, int& b, int& c, int* sm /* 96 ints */) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int ia = a, ib = b, ic = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int ta = __shfl_up_sync(0xffffffffu, ia, o), tb = __shfl_up_sync(0xffffffffu, ib, o),
              tc = __shfl_up_sync(0xffffffffu, ic, o);
    if (lane >= o) { ia += ta; ib += tb; ic += tc; }
  }
  if (lane == 31) { sm[warp] = ia; sm[32 + warp] = ib; sm[64 + warp] = ic; }
  __syncthreads();
  if (warp == 0) {
    const int oa = sm[lane], ob = sm[32 + lane], oc = sm[64 + lane];
    int wa = oa, wb = ob, wc = oc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int ta = __shfl_up_sync(0xffffffffu, wa, o), tb = __shfl_up_sync(0xffffffffu, wb, o),
                tc = __shfl_up_sync(0xffffffffu, wc, o);
      if (lane >= o) { wa += ta; wb += tb; wc += tc; }
    }
    sm[lane] = wa - oa; sm[32 + lane] = wb - ob; sm[64 + lane] = wc - oc;
  }
  __syncthreads();
  a = ia - a + sm[warp];
  b = ib - b + sm[32 + warp];
  c = ic - c + sm[64 + warp];
}

// Largest u in [0, n) with arr[u] <= key (arr non-decreasing, arr[0] <= key).
__device__ __forceinline__ int last_le(const int* __restrict__ arr, int n, int key) {
  int lo = 0, hi = n - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (__ldg(arr + mid) <= key) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// ---------------------------------------------------------------------------------------------- S3 tokenizer
__device__ __forceinline__ int s3_token_len(int T) { const int l1 = (T - 1) / 2 + 1; return (l1 - 1) / 2 + 1; }

// P[u] = first token row (one guard row after every utterance), L[u] = tokens, t64 / t128 = first 64- / 128-row
// query tile of the utterance.
static __global__ void __launch_bounds__(kScanThreads) s3_scan_kernel(const int* __restrict__ frame_off, int n, int* __restrict__ P,
                                                               int* __restrict__ L, int* __restrict__ t64,
                                                               int* __restrict__ t128) {
  __shared__ int sm[96];
  const int per = (n + kScanThreads - 1) / kScanThreads;
  const int u0 = min(n, (int)threadIdx.x * per), u1 = min(n, u0 + per);
  int rows = 0, a = 0, b = 0;
  for (int u = u0; u < u1; ++u) {
    const int l = s3_token_len(__ldg(frame_off + u + 1) - __ldg(frame_off + u));
    rows += l + 1; a += (l + 63) >> 6; b += (l + 127) >> 7;
  }
  block_excl_scan3(rows, a, b, sm);
  for (int u = u0; u < u1; ++u) {
    const int l = s3_token_len(__ldg(frame_off + u + 1) - __ldg(frame_off + u));
    P[u] = rows; L[u] = l; t64[u] = a; t128[u] = b;
    rows += l + 1; a += (l + 63) >> 6; b += (l + 127) >> 7;
  }
}

// row1[r] (conv1 output rows, utterance u at 2 P_u + 1 + t) = utterance or -1; rutt / rpos over token rows; the two
// attention tile lists (utterance, first query row).
static __global__ void s3_fill_kernel(const int* __restrict__ frame_off, int n, const int* __restrict__ P, const int* __restrict__ L,
                               const int* __restrict__ t64, const int* __restrict__ t128, int Rtok, int R1, int n64, int n128,
                               int* __restrict__ row1, int* __restrict__ rutt, int* __restrict__ rpos,
                               int2* __restrict__ tiles, int2* __restrict__ tiles128) {
  const int top = max(max(Rtok, R1), max(n64, n128));
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < top; i += gridDim.x * blockDim.x) {
    if (i < Rtok) {
      const int u = last_le(P, n, i), t = i - __ldg(P + u);
      const bool ok = t < __ldg(L + u);
      rutt[i] = ok ? u : -1;
      rpos[i] = ok ? t : -1;
    }
    if (i < R1) {
      int v = -1;
      if (i >= 1) {
        const int u = last_le(P, n, (i - 1) >> 1), t = i - (2 * __ldg(P + u) + 1);
        const int L1 = (__ldg(frame_off + u + 1) - __ldg(frame_off + u) - 1) / 2 + 1;
        if (t < L1) v = u;
      }
      row1[i] = v;
    }
    if (i < n64) { const int u = last_le(t64, n, i); tiles[i] = make_int2(u, (i - __ldg(t64 + u)) << 6); }
    if (i < n128) { const int u = last_le(t128, n, i); tiles128[i] = make_int2(u, (i - __ldg(t128 + u)) << 7); }
  }
}

// ---------------------------------------------------------------------------------------------- CAMPPlus
// B[u] = first 50 Hz row (two guard rows after every utterance), T2[u] = 50 Hz frames, SB[u] = first pooling segment.
static __global__ void __launch_bounds__(kScanThreads) cam_scan_kernel(const int* __restrict__ frame_off, int n, int seg_len,
                                                                int* __restrict__ B, int* __restrict__ T2,
                                                                int* __restrict__ SB) {
  __shared__ int sm[96];
  const int per = (n + kScanThreads - 1) / kScanThreads;
  const int u0 = min(n, (int)threadIdx.x * per), u1 = min(n, u0 + per);
  int rows = 0, segs = 0, unused = 0;
  for (int u = u0; u < u1; ++u) {
    const int t2 = (__ldg(frame_off + u + 1) - __ldg(frame_off + u) - 1) / 2 + 1;
    rows += t2 + 2; segs += (t2 + seg_len - 1) / seg_len;
  }
  block_excl_scan3(rows, segs, unused, sm);
  for (int u = u0; u < u1; ++u) {
    const int t2 = (__ldg(frame_off + u + 1) - __ldg(frame_off + u) - 1) / 2 + 1;
    B[u] = rows; T2[u] = t2; SB[u] = segs;
    rows += t2 + 2; segs += (t2 + seg_len - 1) / seg_len;
  }
}

// rowsrc[r] (100 Hz rows, utterance u at 2 B_u + 2 + t) = source fbank frame or -1; rowseg[r] (50 Hz rows) = pooling
// segment or -1; segutt[s] = utterance of segment s.
static __global__ void cam_fill_kernel(const int* __restrict__ frame_off, int n, int seg_len, const int* __restrict__ B,
                                const int* __restrict__ T2, const int* __restrict__ SB, int R100, int R50, int n_seg,
                                int* __restrict__ rowsrc, int* __restrict__ rowseg, int* __restrict__ segutt) {
  const int top = max(max(R100, R50), n_seg);
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < top; i += gridDim.x * blockDim.x) {
    if (i < R100) {
      int v = -1;
      if (i >= 2) {
        const int u = last_le(B, n, (i - 2) >> 1), t = i - 2 * __ldg(B + u) - 2;
        const int f0 = __ldg(frame_off + u);
        if (t < __ldg(frame_off + u + 1) - f0) v = f0 + t;
      }
      rowsrc[i] = v;
    }
    if (i < R50) {
      const int u = last_le(B, n, i), t = i - __ldg(B + u);
      rowseg[i] = t < __ldg(T2 + u) ? __ldg(SB + u) + t / seg_len : -1;
    }
    if (i < n_seg) segutt[i] = last_le(SB, n, i);
  }
}

}  // namespace plan
}  // namespace fy
