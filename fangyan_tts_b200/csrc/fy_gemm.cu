// Persistent warp-specialised bf16 GEMM for sm_100a: TMA -> shared (128B swizzle) -> tcgen05.mma -> TMEM
// -> tcgen05.ld epilogue.   D[M,N] = sum_s A_s[M,K_s] * B_s[N,K_s]^T, fp32 accumulate.
//
// "Segments" let one launch accumulate several (row-shifted A view, B column range) products into the
// same TMEM accumulator.  That one mechanism expresses
//   * the 3-term split-bf16 product (hi*hi + hi*lo + lo*hi) used for fp32-faithful results,
//   * 1-D convolutions over time-major activations (one segment per tap, row offset = tap * dilation),
//   * strided convolutions through a "two rows per row" view of the input.
// Rows outside the tensor (negative or past the end) are zero-filled by TMA, which is the convolution's
// zero padding.
#include <math.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>

#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {

namespace {

constexpr int BM = 128;
constexpr int BK = 64;
// 4 control warps (TMA producer, MMA issuer, TMEM allocator, spare) + the epilogue warps.  BN = 256 tiles get two
// epilogue warpgroups, each draining one half of the accumulator's columns, so the erf-GELU / hi-lo-split epilogues
// of the S3 GEMMs stay hidden behind the next tile's main loop; narrow tiles keep one warpgroup (two CTAs per SM).
template <int BN> constexpr int epi_warps() { return BN == 256 ? 8 : 4; }
template <int BN> constexpr int cta_threads() { return 128 + 32 * epi_warps<BN>(); }

__device__ __forceinline__ uint32_t pack_f16(float a, float b) {
  __half2 v = __floats2half2_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// BN = 256 (the S3 encoder's big GEMMs) owns the SM; the narrow tiles of CAMPPlus (BN = 128 / 32) have short K loops
// and latency-bound epilogues, so two CTAs share an SM and overlap each other's epilogue with their main loops.
// NCTA = 2 is the CTA-pair form of the BN = 256 kernel: two SMs of one cluster compute a 256 x 256 tile with
// tcgen05.mma.cta_group::2, each holding 128 rows of A and HALF of the B tile, which cuts the B traffic from L2 and
// the shared-memory reads per MMA by a third and frees room for a 6-deep ring.
template <int BN, int NCTA = 1>
struct Cfg {
  static constexpr int kCtasPerSm = BN == 256 ? 1 : 2;
  static constexpr int kStages = NCTA == 2 ? 6 : (BN == 256 ? 4 : (BN == 128 ? 3 : 4));
  static constexpr int kABytes = BM * BK * 2;
  static constexpr int kBBytes = BN / NCTA * BK * 2;
  static constexpr int kStageBytes = kABytes + kBBytes;
  static constexpr int kAccStages = BN == 32 ? 4 : 2;   // TMEM accumulator ring
  static constexpr int kTmemCols = kAccStages * BN < 32 ? 32 : kAccStages * BN;
  // tiles + barriers + scale/bias staging + alignment slack
  static constexpr int kSmemBytes = kStages * kStageBytes + 256 + 2 * BN * 4 + 1024;
};

// Exact (erf) GELU, x * Phi(x), with erfc from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, below fp32 rounding of
// the product): Phi(x) = h for x < 0 and 1 - h otherwise, h = 0.5 * poly(t) * exp(-x^2 / 2), t = 1 / (1 + p |x| / sqrt 2).
// Evaluating h directly avoids the 1 + erf cancellation for negative x and costs two MUFU ops instead of erff's ~30
// instructions, which is what kept the fc1 epilogue (5120 columns of GELU + hi/lo split) longer than its main loop.
__device__ __forceinline__ float gelu_erf(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(t, 1.061405429f, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float h = 0.5f * p * t * __expf(-z * z);
  return x * (x < 0.0f ? h : 1.0f - h);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// OP = OP_BF16 / OP_F16: 2-byte operands, 64-element K blocks (kind::f16).  OP = OP_TF32: fp32 data consumed as tf32,
// 32-element K blocks (kind::tf32).  All are 128-byte rows, so tiles, swizzle and descriptors are identical.
template <int BN, int OP, int NCTA>
__global__ void __launch_bounds__(cta_threads<BN>(), Cfg<BN, NCTA>::kCtasPerSm) gemm_kernel(const __grid_constant__ GemmArgs args) {
  using C = Cfg<BN, NCTA>;
  constexpr bool PAIR = NCTA == 2;
  const uint32_t cta_rank = PAIR ? ptx::cluster_ctarank() : 0u;   // rank 0 = the pair's leader (issues the MMAs)
  const int worker = PAIR ? blockIdx.x >> 1 : blockIdx.x;         // persistent tile worker = CTA or CTA pair
  const int n_workers = PAIR ? gridDim.x >> 1 : gridDim.x;
  constexpr bool TF32 = OP == OP_TF32;
  constexpr int BKE = TF32 ? 32 : 64;  // K elements per block
  constexpr int kEpiThreads = 32 * epi_warps<BN>();
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tfull_bar = empty_bar + C::kStages;
  uint64_t* tempty_bar = tfull_bar + C::kAccStages;
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(tempty_bar + C::kAccStages);
  float* s_scale = reinterpret_cast<float*>(smem + C::kStages * C::kStageBytes + 256);
  float* s_bias = s_scale + BN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr int TM = BM * NCTA;                                  // rows per tile (per pair)
  const int num_m = (args.M + TM - 1) / TM;
  const int num_n = args.N / BN;
  const int num_tiles = num_m * num_n;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&args.tmap_a);
    ptx::prefetch_tmap(&args.tmap_b);
  }
  if (warp == 1 && lane == 0) {
    // pair: the leader's full barrier takes one arrival (its own producer, which posts the byte count of BOTH CTAs'
    // loads) plus both CTAs' TMA bytes; its tempty barrier collects one arrival per epilogue warp of both CTAs.
    // empty / tfull are signalled in both CTAs by multicast commits.
    for (int i = 0; i < C::kStages; ++i) {
      ptx::mbar_init(&full_bar[i], 1);
      ptx::mbar_init(&empty_bar[i], 1);
    }
    for (int i = 0; i < C::kAccStages; ++i) {
      ptx::mbar_init(&tfull_bar[i], 1);
      ptx::mbar_init(&tempty_bar[i], epi_warps<BN>() * NCTA);
    }
    ptx::fence_mbar_init();
  }
  if (PAIR) ptx::cluster_sync();   // both CTAs are resident before TMEM is allocated pair-wide
  if (warp == 2) {
    if (PAIR) ptx::tmem_alloc_pair<C::kTmemCols>(tmem_holder);
    else ptx::tmem_alloc<C::kTmemCols>(tmem_holder);
  }
  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync(); else __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = worker; tile < num_tiles; tile += n_workers) {
        const int m0 = (tile / num_n) * TM + (int)cta_rank * BM;
        const int n0 = (tile % num_n) * BN + (int)cta_rank * (BN / NCTA);   // pair: this CTA's half of the B tile
        for (int s = 0; s < args.nseg; ++s) {
          const GemmSeg sg = args.seg[s];
          for (int kb = 0; kb < sg.nkb; ++kb) {
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            uint8_t* a_dst = tiles + stage * C::kStageBytes;
            uint8_t* b_dst = a_dst + C::kABytes;
            if (PAIR) {
              // The peer's bytes may land on the leader's barrier before this expect_tx: harmless, the phase cannot
              // complete until this (the only) arrival, and the peer cannot run a ring revolution ahead because its
              // empty barrier is only released by the leader's MMA commit.
              if (cta_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], C::kStageBytes * 2);
              ptx::tma_load_2d_pair(a_dst, &args.tmap_a, &full_bar[stage], sg.a_col + kb * BKE, m0 + sg.a_row_off);
              ptx::tma_load_2d_pair(b_dst, &args.tmap_b, &full_bar[stage], sg.b_col + kb * BKE, n0);
            } else {
              ptx::mbar_expect_tx(&full_bar[stage], C::kStageBytes);
              ptx::tma_load_2d(a_dst, &args.tmap_a, &full_bar[stage], sg.a_col + kb * BKE, m0 + sg.a_row_off);
              ptx::tma_load_2d(b_dst, &args.tmap_b, &full_bar[stage], sg.b_col + kb * BKE, n0);
              if (BN == 256)   // tensor-map boxes are at most 128 rows (shared with the pair kernel): second half
                ptx::tma_load_2d(b_dst + 128 * 128, &args.tmap_b, &full_bar[stage], sg.b_col + kb * BKE, n0 + 128);
            }
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0 && cta_rank == 0) {
      constexpr uint32_t idesc = TF32 ? ptx::make_idesc_tf32(TM, BN)
                                      : (OP == OP_F16 ? ptx::make_idesc_f16(TM, BN) : ptx::make_idesc_bf16(TM, BN));
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      for (int tile = worker; tile < num_tiles; tile += n_workers) {
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        uint32_t accumulate = 0;
        for (int s = 0; s < args.nseg; ++s) {
          const int nkb = args.seg[s].nkb;
          for (int kb = 0; kb < nkb; ++kb) {
            ptx::mbar_wait(&full_bar[stage], phase);
            ptx::tc_fence_after();
            const uint32_t a_addr = ptx::smem_u32(tiles + stage * C::kStageBytes);
            const uint64_t da = ptx::make_desc_sw128(a_addr);
            const uint64_t db = ptx::make_desc_sw128(a_addr + C::kABytes);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              if (PAIR) ptx::mma_bf16_ss_pair(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              else if (TF32) ptx::mma_tf32_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              else ptx::mma_bf16_ss(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              accumulate = 1;
            }
            if (PAIR) ptx::mma_commit_pair(&empty_bar[stage]);
            else ptx::mma_commit(&empty_bar[stage]);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
        if (PAIR) ptx::mma_commit_pair(&tfull_bar[as]);
        else ptx::mma_commit(&tfull_bar[as]);
        if (++as == C::kAccStages) { as = 0; aphase ^= 1; }
      }
    }
  } else if (warp >= 4) {
    // ===================== epilogue (warp w drains TMEM lane quarter w % 4; warps 8..11 take the upper columns) ====
    const int q = warp & 3;
    const int et = threadIdx.x - 128;
    constexpr int kChunks = BN / 32 / (epi_warps<BN>() / 4);   // 32-column chunks per warp
    const int c_begin = ((warp - 4) >> 2) * kChunks;
    const GemmEpi& e = args.epi;
    int as = 0;
    uint32_t aphase = 0;
    int staged_n0 = -1;
    for (int tile = worker; tile < num_tiles; tile += n_workers) {
      const int m0 = (tile / num_n) * TM + (int)cta_rank * BM;
      const int n0 = (tile % num_n) * BN;
      if (n0 != staged_n0) {  // per-column scale / bias of this N block -> shared memory (uniform branch)
        asm volatile("bar.sync 1, %0;\n" ::"n"(kEpiThreads) : "memory");
        for (int i = et; i < BN; i += kEpiThreads) {
          s_scale[i] = e.scale ? __ldg(e.scale + n0 + i) : 1.0f;
          s_bias[i] = e.bias ? __ldg(e.bias + n0 + i) : 0.0f;
        }
        asm volatile("bar.sync 1, %0;\n" ::"n"(kEpiThreads) : "memory");
        staged_n0 = n0;
      }
      const int row = m0 + q * 32 + lane;
      bool row_ok = row < args.M;
      int ridx = 0;
      if (e.row_idx) ridx = row_ok ? __ldg(e.row_idx + row) : -1;
      if (ridx == -2) row_ok = false;  // -2: leave the output row untouched; -1: write zeros
      const bool keep = ridx >= 0;
      const size_t orow = e.remap_q ? (size_t)(row / e.remap_q) * e.remap_fp + row % e.remap_q + e.remap_add : (size_t)row;
      uint4 rpre[4];
      if (e.resid && row_ok) {  // first chunk of the residual: issue the loads before waiting for the accumulator
        const uint4* rp = reinterpret_cast<const uint4*>(e.resid + orow * e.ldo + (size_t)e.col_off + n0 + c_begin * 32);
#pragma unroll
        for (int j = 0; j < 4; ++j) rpre[j] = __ldg(rp + j);
      }
      ptx::mbar_wait(&tfull_bar[as], aphase);
      ptx::tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
      if (BN == 256 && e.mode == EPI_QKV_ROPE) {
        // 64-column heads: columns [h*64, +32) pair with [h*64 + 32, +32).  Drained 16 + 16 columns at a time.
        const int which = n0 / e.rope_D;
        const int cbase = n0 - which * e.rope_D;
        __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(which == 0 ? e.out : (which == 1 ? e.out2 : e.out3));
        __nv_bfloat16* o = ob + (size_t)row * e.ldo + cbase;
        int pos = row_ok ? __ldg(e.rope_pos + row) : 0;
        if (pos < 0) pos = 0;
        const float* ct = e.rope_cos + (size_t)pos * 64;
        const float* sn = e.rope_sin + (size_t)pos * 64;
#pragma unroll 1
        for (int hs = 0; hs < kChunks; ++hs) {           // kChunks / 2 heads x 2 sub-chunks of 16 pairs
          const int ca = c_begin * 32 + (hs >> 1) * 64 + (hs & 1) * 16;   // tile-local column of the first half
          uint32_t va[16], vb[16];
          ptx::tmem_ld_32x16(t_row + ca, va);
          ptx::tmem_ld_32x16(t_row + ca + 32, vb);
          ptx::tmem_ld_wait();
          if (row_ok) {
            float xa[16], xb[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              xa[j] = __uint_as_float(va[j]) + s_bias[ca + j];
              xb[j] = __uint_as_float(vb[j]) + s_bias[ca + 32 + j];
            }
            if (which < 2) {
              const int d0 = (hs & 1) * 16;
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 c4 = __ldg(reinterpret_cast<const float4*>(ct + d0) + j4);
                const float4 s4 = __ldg(reinterpret_cast<const float4*>(sn + d0) + j4);
                const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                  const float a = xa[4 * j4 + k], b = xb[4 * j4 + k];
                  xa[4 * j4 + k] = (a * cc[k] - b * ss[k]) * e.rope_scale;
                  xb[4 * j4 + k] = (b * cc[k] + a * ss[k]) * e.rope_scale;
                }
              }
            }
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              float* xs = half ? xb : xa;
              uint4* dhi = reinterpret_cast<uint4*>(o + ca + half * 32);
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                uint4 w;
                w.x = pack_bf16(xs[8 * j], xs[8 * j + 1]); w.y = pack_bf16(xs[8 * j + 2], xs[8 * j + 3]);
                w.z = pack_bf16(xs[8 * j + 4], xs[8 * j + 5]); w.w = pack_bf16(xs[8 * j + 6], xs[8 * j + 7]);
                dhi[j] = w;
              }
              if (e.lo_col_off >= 0) {
                uint4* dlo = reinterpret_cast<uint4*>(o + e.lo_col_off + ca + half * 32);
#pragma unroll
                for (int j = 0; j < 16; ++j) xs[j] -= __bfloat162float(__float2bfloat16_rn(xs[j]));
#pragma unroll
                for (int j = 0; j < 2; ++j) {
                  uint4 w;
                  w.x = pack_bf16(xs[8 * j], xs[8 * j + 1]); w.y = pack_bf16(xs[8 * j + 2], xs[8 * j + 3]);
                  w.z = pack_bf16(xs[8 * j + 4], xs[8 * j + 5]); w.w = pack_bf16(xs[8 * j + 6], xs[8 * j + 7]);
                  dlo[j] = w;
                }
              }
            }
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) {
          if (PAIR) ptx::mbar_arrive_cluster(&tempty_bar[as], 0); else ptx::mbar_arrive(&tempty_bar[as]);
        }
        if (++as == C::kAccStages) { as = 0; aphase ^= 1; }
        continue;
      }
#pragma unroll 1
      for (int c = c_begin; c < c_begin + kChunks; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(t_row + c * 32, v);
        ptx::tmem_ld_wait();
        if (row_ok) {
          float x[32];
          const size_t col = (size_t)e.col_off + n0 + c * 32;
          float rs[32];
          if (e.resid) {
            const uint4* rp = reinterpret_cast<const uint4*>(e.resid + orow * e.ldo + col);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint4 w = c == c_begin ? rpre[j] : __ldg(rp + j);
              const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                if (e.out_f16) {
                  const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&ww[k]));
                  rs[8 * j + 2 * k] = f.x;
                  rs[8 * j + 2 * k + 1] = f.y;
                } else {
                  rs[8 * j + 2 * k] = __uint_as_float(ww[k] << 16);
                  rs[8 * j + 2 * k + 1] = __uint_as_float(ww[k] & 0xffff0000u);
                }
              }
            }
          }
          if (e.resid_f32) {
            const float4* rp = reinterpret_cast<const float4*>(e.resid_f32 + orow * e.ldo + col);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 w = __ldg(rp + j);
              rs[4 * j] = w.x; rs[4 * j + 1] = w.y; rs[4 * j + 2] = w.z; rs[4 * j + 3] = w.w;
            }
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            float t = __uint_as_float(v[j]) * s_scale[c * 32 + j] + s_bias[c * 32 + j];
            if (e.resid || e.resid_f32) t += rs[j];
            if (e.act == ACT_GELU) t = gelu_erf(t);
            else if (e.act == ACT_RELU) t = fmaxf(t, 0.0f);
            x[j] = t;
          }
          if (e.row_gate && keep) {
            const float4* g = reinterpret_cast<const float4*>(e.row_gate + (size_t)ridx * args.N + n0 + c * 32);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float4 gv = __ldg(g + j);
              x[4 * j] *= gv.x; x[4 * j + 1] *= gv.y; x[4 * j + 2] *= gv.z; x[4 * j + 3] *= gv.w;
            }
          }
          if (!keep) {
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = 0.0f;
          }
          if (e.mode == EPI_BF16) {
            __nv_bfloat16* o = reinterpret_cast<__nv_bfloat16*>(e.out) + orow * e.ldo;
            uint4* dst = reinterpret_cast<uint4*>(o + col);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 w;
              if (e.out_f16) {
                w.x = pack_f16(x[8 * j], x[8 * j + 1]);
                w.y = pack_f16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_f16(x[8 * j + 4], x[8 * j + 5]);
                w.w = pack_f16(x[8 * j + 6], x[8 * j + 7]);
              } else {
                w.x = pack_bf16(x[8 * j], x[8 * j + 1]);
                w.y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_bf16(x[8 * j + 4], x[8 * j + 5]);
                w.w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
              }
              dst[j] = w;
            }
            if (e.lo_col_off >= 0) {
              uint4* dlo = reinterpret_cast<uint4*>(o + (size_t)e.lo_col_off + n0 + c * 32);
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] -= __bfloat162float(__float2bfloat16_rn(x[j]));
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                uint4 w;
                w.x = pack_bf16(x[8 * j], x[8 * j + 1]);
                w.y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_bf16(x[8 * j + 4], x[8 * j + 5]);
                w.w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
                dlo[j] = w;
              }
            }
          } else {
            if (e.round_tf32) {
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] = ptx::round_tf32(x[j]);
            }
            float4* dst = reinterpret_cast<float4*>(reinterpret_cast<float*>(e.out) + orow * e.ldo + col);
            if (e.mode == EPI_F32_ADD) {
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                float4 r = dst[j];
                r.x += x[4 * j]; r.y += x[4 * j + 1]; r.z += x[4 * j + 2]; r.w += x[4 * j + 3];
                dst[j] = r;
              }
            } else {
#pragma unroll
              for (int j = 0; j < 8; ++j) dst[j] = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
        if (lane == 0) {
          if (PAIR) ptx::mbar_arrive_cluster(&tempty_bar[as], 0); else ptx::mbar_arrive(&tempty_bar[as]);
        }
      if (++as == C::kAccStages) { as = 0; aphase ^= 1; }
    }
  }

  ptx::tc_fence_before();
  if (PAIR) ptx::cluster_sync(); else __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    if (PAIR) ptx::tmem_dealloc_pair<C::kTmemCols>(tmem_base);
    else ptx::tmem_dealloc<C::kTmemCols>(tmem_base);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn get_encode() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

template <int BN, int OP, int NCTA>
int launch_bn(const GemmArgs& a, int num_sms, cudaStream_t st) {
  using C = Cfg<BN, NCTA>;
  static bool configured = false;
  if (!configured) {
    FY_CUDA(cudaFuncSetAttribute(gemm_kernel<BN, OP, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    configured = true;
  }
  const int tiles = ((a.M + BM * NCTA - 1) / (BM * NCTA)) * (a.N / BN);
  const int slots = num_sms * C::kCtasPerSm / NCTA;
  const int workers = tiles < slots ? tiles : slots;
  if (NCTA == 1) {
    gemm_kernel<BN, OP, NCTA><<<workers, cta_threads<BN>(), C::kSmemBytes, st>>>(a);
  } else {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(workers * NCTA);
    cfg.blockDim = dim3(cta_threads<BN>());
    cfg.dynamicSmemBytes = C::kSmemBytes;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = NCTA;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    FY_CUDA(cudaLaunchKernelEx(&cfg, gemm_kernel<BN, OP, NCTA>, a));
  }
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

// CTA pairs pay off once there are enough 256-row tiles to keep all pairs busy; FY_GEMM_2CTA=0 turns them off.
bool use_pair(const GemmArgs& a) {
  static int enabled = -1;
  if (enabled < 0) {
    const char* e = getenv("FY_GEMM_2CTA");
    enabled = (e && e[0] == '0') ? 0 : 1;
  }
  return enabled && a.op == OP_BF16 && a.N % 256 == 0 && a.M >= 2048;
}

}  // namespace

int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  if (box_rows > 128) box_rows = 128;  // 256-wide B tiles are fetched as two 128-row boxes
  EncodeTiledFn enc = get_encode();
  FY_CHECK(enc != nullptr, FY_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  FY_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 2) % 16 == 0, FY_ERR_ARG,
           "tensor map base/stride must be 16-byte aligned (ld=%llu)", (unsigned long long)ld);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 2};
  cuuint32_t box[2] = {BK, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FY_CHECK(r == CUDA_SUCCESS, FY_ERR_CUDA, "cuTensorMapEncodeTiled failed with %d (rows=%llu cols=%llu ld=%llu)", (int)r,
           (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
  return FY_OK;
}

int make_tmap_f32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  if (box_rows > 128) box_rows = 128;  // 256-wide B tiles are fetched as two 128-row boxes
  EncodeTiledFn enc = get_encode();
  FY_CHECK(enc != nullptr, FY_ERR_CUDA, "cuTensorMapEncodeTiled entry point not available");
  FY_CHECK((reinterpret_cast<uintptr_t>(base) & 15) == 0 && (ld * 4) % 16 == 0, FY_ERR_ARG,
           "tensor map base/stride must be 16-byte aligned (ld=%llu)", (unsigned long long)ld);
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {ld * 4};
  cuuint32_t box[2] = {32, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FY_CHECK(r == CUDA_SUCCESS, FY_ERR_CUDA, "cuTensorMapEncodeTiled(f32) failed with %d (rows=%llu cols=%llu ld=%llu)",
           (int)r, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
  return FY_OK;
}

int gemm_block_n(int N) {
  if (N % 256 == 0) return 256;
  if (N % 128 == 0) return 128;
  if (N % 32 == 0) return 32;
  return 0;
}

int launch_gemm(const GemmArgs& a, int num_sms, cudaStream_t st) {
  FY_CHECK(a.M > 0 && a.nseg > 0 && a.nseg <= kMaxSeg, FY_ERR_ARG, "bad GEMM shape M=%d nseg=%d", a.M, a.nseg);
  FY_CHECK(a.epi.out != nullptr && a.epi.ldo > 0, FY_ERR_ARG, "GEMM epilogue has no output");
  switch (gemm_block_n(a.N)) {
#define FY_DISPATCH_OP(BN_)                                                          \
  switch (a.op) {                                                                    \
    case OP_BF16: return launch_bn<BN_, OP_BF16, 1>(a, num_sms, st);                 \
    case OP_TF32: return launch_bn<BN_, OP_TF32, 1>(a, num_sms, st);                 \
    case OP_F16: return launch_bn<BN_, OP_F16, 1>(a, num_sms, st);                   \
    default: FY_CHECK(false, FY_ERR_ARG, "bad GEMM operand type %d", a.op);          \
  }                                                                                  \
  break
    case 256:
      if (use_pair(a)) return launch_bn<256, OP_BF16, 2>(a, num_sms, st);
      FY_DISPATCH_OP(256);
    case 128: FY_DISPATCH_OP(128);
    case 32: FY_DISPATCH_OP(32);
#undef FY_DISPATCH_OP
    default: FY_CHECK(false, FY_ERR_ARG, "GEMM N=%d must be a multiple of 32", a.N);
  }
  return FY_OK;
}

}  // namespace fy
