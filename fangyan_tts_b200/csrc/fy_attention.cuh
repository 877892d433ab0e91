// Variable-length, non-causal self-attention for head_dim 64 (S3 encoder, row A9).
// One CTA = 64 query rows of one (utterance, head); 4 warps x 16 rows; keys/values streamed in 64-row tiles
// through shared memory with cp.async double buffering; flash-style online softmax in fp32.
// Tensor-core work uses mma.sync m16n8k16 (attention is <10% of the encoder's FLOPs; the GEMMs are tcgen05).
// SPLIT = true evaluates every product as hi*hi + hi*lo + lo*hi on bf16 pairs (fp32-faithful mode).
#pragma once
#include "fy_common.cuh"

namespace fy {
namespace attn {

constexpr int kBQ = 64, kBK = 64, kDH = 64, kThreads = 128;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int sz = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;\n" ::"r"(dst), "l"(src), "r"(sz) : "memory");
}
__device__ __forceinline__ void cp_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_wait() { asm volatile("cp.async.wait_group %0;\n" ::"n"(N) : "memory"); }

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];\n"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};\n"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// A [64 rows][64 bf16] tile: 128-byte rows, 16-byte chunks XOR-swizzled by (row & 7).
__device__ __forceinline__ uint32_t tile_addr(uint32_t base, int row, int chunk) {
  return base + row * 128 + ((chunk ^ (row & 7)) << 4);
}

// Per-thread constants of load_tile's access pattern: thread t moves chunk (t & 7) of rows (t >> 3) + 16 i, i < 4.
// With them a tile load is four cp.async whose addresses are base + i * stride (no per-call index arithmetic).
struct TileLane {
  uint32_t soff;      // byte offset of (row r0, chunk ch) inside a swizzled tile
  int r0;             // first row of this thread
  size_t goff;        // element offset r0 * ld + ch * 8 inside the global tile
  size_t gstep;       // 16 * ld elements
};
__device__ __forceinline__ TileLane tile_lane(int tid, int ld) {
  TileLane t;
  t.r0 = tid >> 3;
  const int ch = tid & 7;
  t.soff = t.r0 * 128 + ((ch ^ (t.r0 & 7)) << 4);     // (r0 + 16 i) & 7 == r0 & 7
  t.goff = (size_t)t.r0 * ld + ch * 8;
  t.gstep = (size_t)16 * ld;
  return t;
}
__device__ __forceinline__ void load_tile(uint32_t sbase, const __nv_bfloat16* g, const TileLane& t, int nvalid) {
  const __nv_bfloat16* p = g + t.goff;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const bool ok = t.r0 + 16 * i < nvalid;
    cp_async16(sbase + t.soff + i * 2048, p, ok);   // !ok: src-size 0, nothing is read (rows past the utterance are
                                                    // still inside the caller's q/k/v workspace)
    p += t.gstep;
  }
}

// Load a [64][64] bf16 tile: global rows row0.. (stride ld elements), zero-filled for rows >= nvalid.
__device__ __forceinline__ void load_tile(uint32_t sbase, const __nv_bfloat16* g, size_t ld, int nvalid, int tid) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int c = tid + i * kThreads;  // 512 chunks
    const int row = c >> 3, ch = c & 7;
    const bool ok = row < nvalid;
    cp_async16(tile_addr(sbase, row, ch), g + (size_t)(ok ? row : 0) * ld + ch * 8, ok);
  }
}

__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// 2^x without the denormal fix-up sequence exp2f expands to (results below 2^-126 flush to zero, as they should here)
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// bf16 pair of the residuals (a - hi(a), b - hi(b)) given the already packed pair hi = (bf16(a), bf16(b))
__device__ __forceinline__ uint32_t pack_lo(uint32_t hi, float a, float b) {
  return pack2(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xffff0000u));
}
__device__ __forceinline__ float bf_round(float a) { return __bfloat162float(__float2bfloat16_rn(a)); }

// q, k, v, out: bf16 [rows, ld] with head h at columns h*64 (hi) and lo_off + h*64 (lo, SPLIT only).
// tiles: int2 (utt, q0) per blockIdx.x; utt_P / utt_L give the utterance's first row and length.
template <bool SPLIT>
__global__ void __launch_bounds__(kThreads, 3) attention_kernel(const __nv_bfloat16* __restrict__ q,
                                                             const __nv_bfloat16* __restrict__ k,
                                                             const __nv_bfloat16* __restrict__ v,
                                                             __nv_bfloat16* __restrict__ out, int ld, int lo_off,
                                                             const int2* __restrict__ tiles,
                                                             const int* __restrict__ utt_P,
                                                             const int* __restrict__ utt_L) {
  extern __shared__ __align__(1024) uint8_t smem[];
  constexpr int NP = SPLIT ? 2 : 1;            // parts (hi, lo)
  constexpr int TILE = kBK * kDH * 2;          // 8 KB
  // layout: stage0: K[NP] V[NP] | stage1: K[NP] V[NP].  Q[NP] is staged in stage 1 and moved to registers before
  // the first prefetch overwrites it (64 KB per CTA in SPLIT mode: three CTAs per SM).
  const uint32_t sKV = smem_u32(smem);
  const uint32_t sQ = sKV + 2 * NP * TILE;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int2 tl = tiles[blockIdx.x];
  const int u = tl.x, q0 = tl.y, h = blockIdx.y;
  const int P = utt_P[u], L = utt_L[u];
  const size_t col = (size_t)h * kDH;
  const int nq = min(kBQ, L - q0);
  const int nkt = (L + kBK - 1) / kBK;

  // prologue: Q and first K/V tile
  const TileLane tlane = tile_lane(tid, ld);
  const __nv_bfloat16* kp = k + (size_t)P * ld + col;    // this (utterance, head)'s key / value rows, advanced per tile
  const __nv_bfloat16* vp = v + (size_t)P * ld + col;
  const size_t tile_step = (size_t)kBK * ld;
#pragma unroll
  for (int p = 0; p < NP; ++p) load_tile(sQ + p * TILE, q + (size_t)(P + q0) * ld + col + p * lo_off, tlane, nq);
#pragma unroll
  for (int p = 0; p < NP; ++p) {
    load_tile(sKV + p * TILE, kp + p * lo_off, tlane, min(kBK, L));
    load_tile(sKV + (NP + p) * TILE, vp + p * lo_off, tlane, min(kBK, L));
  }
  cp_commit();

  uint32_t qf[NP][4][4];
  float o[8][4];
#pragma unroll
  for (int j = 0; j < 8; ++j) { o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f; }
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};
  const float kLog2e = 1.4426950408889634f;

  // Q fragments -> registers, then stage 1 is free for the first prefetch
  cp_wait<0>();
  __syncthreads();
#pragma unroll
  for (int p = 0; p < NP; ++p)
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const int row = warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
      const int ch = ks * 2 + (lane >> 4);
      ldsm_x4(tile_addr(sQ + p * TILE, row, ch), qf[p][ks][0], qf[p][ks][1], qf[p][ks][2], qf[p][ks][3]);
    }
  __syncthreads();

  for (int kt = 0; kt < nkt; ++kt) {
    const int st = kt & 1;
    if (kt + 1 < nkt) {
      const int k0n = (kt + 1) * kBK;
      const uint32_t sb = sKV + (st ^ 1) * 2 * NP * TILE;
      kp += tile_step;
      vp += tile_step;
#pragma unroll
      for (int p = 0; p < NP; ++p) {
        load_tile(sb + p * TILE, kp + p * lo_off, tlane, min(kBK, L - k0n));
        load_tile(sb + (NP + p) * TILE, vp + p * lo_off, tlane, min(kBK, L - k0n));
      }
      cp_commit();
    }
    if (kt > 0) {   // tile 0 arrived with Q
      if (kt + 1 < nkt) cp_wait<1>(); else cp_wait<0>();
      __syncthreads();
    }
    const uint32_t sK = sKV + st * 2 * NP * TILE;
    const uint32_t sV = sK + NP * TILE;

    // ---- S = Q K^T  (16 x 64 per warp)
    float s[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) { s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f; }
    // Four key blocks (independent accumulators) are interleaved so back-to-back mma.sync never wait on each other.
#pragma unroll
    for (int half = 0; half < 2; ++half) {
#pragma unroll
      for (int jg = 0; jg < 2; ++jg) {
        uint32_t b[4][NP][4];
#pragma unroll
        for (int jj = 0; jj < 4; ++jj) {
          const int row = (jg * 4 + jj) * 8 + (lane & 7);
          const int ch = half * 4 + (lane >> 3);
#pragma unroll
          for (int p = 0; p < NP; ++p)
            ldsm_x4(tile_addr(sK + p * TILE, row, ch), b[jj][p][0], b[jj][p][1], b[jj][p][2], b[jj][p][3]);
        }
#pragma unroll
        for (int kk = 0; kk < 2; ++kk) {
          const int ks = half * 2 + kk;
#pragma unroll
          for (int jj = 0; jj < 4; ++jj) mma16816(s[jg * 4 + jj], qf[0][ks], b[jj][0][2 * kk], b[jj][0][2 * kk + 1]);
          if (SPLIT) {
#pragma unroll
            for (int jj = 0; jj < 4; ++jj)
              mma16816(s[jg * 4 + jj], qf[0][ks], b[jj][NP - 1][2 * kk], b[jj][NP - 1][2 * kk + 1]);
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) mma16816(s[jg * 4 + jj], qf[NP - 1][ks], b[jj][0][2 * kk], b[jj][0][2 * kk + 1]);
          }
        }
      }
    }
    // ---- mask keys beyond L, online softmax
    if ((kt + 1) * kBK > L) {   // only the last tile of an utterance can hold keys beyond L
      const int kbase = kt * kBK + (lane & 3) * 2;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int kc = kbase + j * 8;
        if (kc >= L) { s[j][0] = -INFINITY; s[j][2] = -INFINITY; }
        if (kc + 1 >= L) { s[j][1] = -INFINITY; s[j][3] = -INFINITY; }
      }
    }
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      mx[0] = fmaxf(mx[0], fmaxf(s[j][0], s[j][1]));
      mx[1] = fmaxf(mx[1], fmaxf(s[j][2], s[j][3]));
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
    }
    // mrow holds the running maximum already scaled by log2(e): p = 2^(s log2e - mrow) with the product exact inside
    // the FFMA, and the rescale 2^(mrow_old - mrow_new) uses the same rounded values, so every tile of a row is
    // referenced consistently and the rounding of mrow cancels in the final division.
    float corr[2], moff[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      moff[r] = fmaxf(mrow[r], mx[r] * kLog2e);    // finite: every tile holds at least one valid key
      corr[r] = ex2(mrow[r] - moff[r]);
      mrow[r] = moff[r];
    }
    float ls[2] = {0.f, 0.f};
#pragma unroll
    for (int j = 0; j < 8; ++j) {   // one FFMA + one MUFU per score
      s[j][0] = ex2(fmaf(s[j][0], kLog2e, -moff[0]));
      s[j][1] = ex2(fmaf(s[j][1], kLog2e, -moff[0]));
      s[j][2] = ex2(fmaf(s[j][2], kLog2e, -moff[1]));
      s[j][3] = ex2(fmaf(s[j][3], kLog2e, -moff[1]));
      ls[0] += s[j][0] + s[j][1];
      ls[1] += s[j][2] + s[j][3];
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) lrow[r] = lrow[r] * corr[r] + ls[r];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      o[j][0] *= corr[0]; o[j][1] *= corr[0];
      o[j][2] *= corr[1]; o[j][3] *= corr[1];
    }
    // ---- O += P V
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {  // 16 keys per step
      uint32_t pa[NP][4];
      pa[0][0] = pack2(s[2 * ks][0], s[2 * ks][1]);
      pa[0][1] = pack2(s[2 * ks][2], s[2 * ks][3]);
      pa[0][2] = pack2(s[2 * ks + 1][0], s[2 * ks + 1][1]);
      pa[0][3] = pack2(s[2 * ks + 1][2], s[2 * ks + 1][3]);
      if (SPLIT) {   // lo = p - hi, with hi read back from the packed pair (one shift / one mask per element)
        pa[NP - 1][0] = pack_lo(pa[0][0], s[2 * ks][0], s[2 * ks][1]);
        pa[NP - 1][1] = pack_lo(pa[0][1], s[2 * ks][2], s[2 * ks][3]);
        pa[NP - 1][2] = pack_lo(pa[0][2], s[2 * ks + 1][0], s[2 * ks + 1][1]);
        pa[NP - 1][3] = pack_lo(pa[0][3], s[2 * ks + 1][2], s[2 * ks + 1][3]);
      }
#pragma unroll
      for (int jpg = 0; jpg < 2; ++jpg) {  // two pairs of 8-wide d tiles = four independent accumulators
        uint32_t b[2][NP][4];
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
          const int row = ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8;
          const int ch = (jpg * 2 + jj) * 2 + (lane >> 4);
#pragma unroll
          for (int p = 0; p < NP; ++p)
            ldsm_x4_t(tile_addr(sV + p * TILE, row, ch), b[jj][p][0], b[jj][p][1], b[jj][p][2], b[jj][p][3]);
        }
#pragma unroll
        for (int jj = 0; jj < 2; ++jj)
#pragma unroll
          for (int t = 0; t < 2; ++t) mma16816(o[(jpg * 2 + jj) * 2 + t], pa[0], b[jj][0][2 * t], b[jj][0][2 * t + 1]);
        if (SPLIT) {
#pragma unroll
          for (int jj = 0; jj < 2; ++jj)
#pragma unroll
            for (int t = 0; t < 2; ++t)
              mma16816(o[(jpg * 2 + jj) * 2 + t], pa[NP - 1], b[jj][0][2 * t], b[jj][0][2 * t + 1]);
#pragma unroll
          for (int jj = 0; jj < 2; ++jj)
#pragma unroll
            for (int t = 0; t < 2; ++t)
              mma16816(o[(jpg * 2 + jj) * 2 + t], pa[0], b[jj][NP - 1][2 * t], b[jj][NP - 1][2 * t + 1]);
        }
      }
    }
    __syncthreads();  // everyone is done with stage st before it is refilled
  }

  // ---- finalise: divide by the row sums, write bf16 (hi | lo)
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 1);
    lrow[r] += __shfl_xor_sync(0xffffffffu, lrow[r], 2);
  }
  const int r0 = warp * 16 + (lane >> 2);
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int qr = r0 + r * 8;
    if (qr < nq) {
      const float inv = 1.0f / lrow[r];
      __nv_bfloat16* dst = out + (size_t)(P + q0 + qr) * ld + col + (lane & 3) * 2;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float a = o[j][2 * r] * inv, b = o[j][2 * r + 1] * inv;
        *reinterpret_cast<uint32_t*>(dst + j * 8) = pack2(a, b);
        if (SPLIT) *reinterpret_cast<uint32_t*>(dst + lo_off + j * 8) = pack2(a - bf_round(a), b - bf_round(b));
      }
    }
  }
}

template <bool SPLIT>
constexpr int smem_bytes() { return (SPLIT ? 2 : 1) * kBK * kDH * 2 * 4; }

}  // namespace attn
}  // namespace fy
