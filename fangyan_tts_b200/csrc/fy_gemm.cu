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

// ex2 / rcp without the denormal fix-up sequences the default (non -ftz) intrinsics expand to
__device__ __forceinline__ float ex2_ftz(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float rcp_ftz(float x) {
  float y;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Exact (erf) GELU, x * Phi(x), with erfc from Abramowitz & Stegun 7.1.26 (|error| <= 1.5e-7, below fp32 rounding of
// the product): Phi(x) = h for x < 0 and 1 - h otherwise, h = 0.5 * poly(t) * exp(-x^2 / 2), t = 1 / (1 + p |x| / sqrt 2).
// Evaluating h directly avoids the 1 + erf cancellation for negative x and costs two MUFU ops instead of erff's ~30
// instructions, which is what kept the fc1 epilogue (5120 columns of GELU + hi/lo split) longer than its main loop.
__device__ __forceinline__ float gelu_erf(float x) {
  const float z = fabsf(x) * 0.70710678118654752f;
  const float t = rcp_ftz(fmaf(0.3275911f, z, 1.0f));
  float p = fmaf(t, 1.061405429f, -1.453152027f);
  p = fmaf(p, t, 1.421413741f);
  p = fmaf(p, t, -0.284496736f);
  p = fmaf(p, t, 0.254829592f);
  const float h = 0.5f * p * t * ex2_ftz(z * z * -1.4426950408889634f);
  return x * (x < 0.0f ? h : 1.0f - h);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// The epilogue holds one accumulator row per lane, so a plain 16-byte store touches 32 different rows (32 L1
// wavefronts, 32 half-written sectors).  Transposing the four 16-byte pieces across each lane quad turns it into
// 8 rows x 64 contiguous bytes: after the call, lane (Q + q) holds piece q of rows Q .. Q+3 in w[0..3].
__device__ __forceinline__ uint4 shfl_xor_u4(uint4 v, int m) {
  v.x = __shfl_xor_sync(0xffffffffu, v.x, m);
  v.y = __shfl_xor_sync(0xffffffffu, v.y, m);
  v.z = __shfl_xor_sync(0xffffffffu, v.z, m);
  v.w = __shfl_xor_sync(0xffffffffu, v.w, m);
  return v;
}
__device__ __forceinline__ void quad_transpose(uint4 (&w)[4], int lane) {
  const bool b0 = lane & 1, b1 = lane & 2;
#pragma unroll
  for (int a = 0; a < 4; a += 2) {   // 2x2 blocks across lane bit 0: pieces (a, a+1)
    const uint4 r = shfl_xor_u4(b0 ? w[a] : w[a + 1], 1);
    if (b0) w[a] = r; else w[a + 1] = r;
  }
#pragma unroll
  for (int a = 0; a < 2; ++a) {      // across lane bit 1: pieces (a, a+2)
    const uint4 r = shfl_xor_u4(b1 ? w[a] : w[a + 2], 2);
    if (b1) w[a] = r; else w[a + 2] = r;
  }
}

// OP = OP_BF16 / OP_F16: 2-byte operands, 64-element K blocks (kind::f16).  OP = OP_TF32: fp32 data consumed as tf32,
// 32-element K blocks (kind::tf32).  All are 128-byte rows, so tiles, swizzle and descriptors are identical.
#ifdef FY_STALLPROF
// Debug build only (tools/gemm_stalls.py): cycles each role spends blocked on its barrier, summed over CTAs.
// slot = GEMM kind (0 QKV, 1 fc1, 2 O, 3 fc2, 4 other) x 8 counters.
__device__ unsigned long long g_stall[5 * 8];
#define FY_SP(...) __VA_ARGS__
#else
#define FY_SP(...)
#endif

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
  uint8_t* smem = smem_raw + ((1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u);   // stays a shared-space pointer
  uint8_t* tiles = smem;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tfull_bar = empty_bar + C::kStages;
  uint64_t* tempty_bar = tfull_bar + C::kAccStages;
  uint64_t* xf_bar = tempty_bar + C::kAccStages;          // [kStages] A tile transformed (GemmArgs::a_scale), 2 arrivals
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(xf_bar + C::kStages);
  static_assert((3 * C::kStages + 2 * C::kAccStages) * 8 + 8 <= 256, "barrier block");
  const bool xform = !PAIR && OP == OP_F16 && args.a_scale != nullptr;
  float* s_scale = reinterpret_cast<float*>(smem + C::kStages * C::kStageBytes + 256);
  float* s_bias = s_scale + BN;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  constexpr int TM = BM * NCTA;                                  // rows per tile (per pair)
  const int num_m = (args.M + TM - 1) / TM;
  const int num_n = args.N / BN;
  const int num_tiles = num_m * num_n;
#ifdef FY_STALLPROF
  int sp_kb = 0;
  for (int s = 0; s < args.nseg; ++s) sp_kb += args.seg[s].nkb * (args.seg[s].a_lo ? 3 : 1);
  const int sp_slot = 8 * (args.N == 3840 ? 0 : args.N == 5120 ? 1 : (args.N == 1280 && sp_kb == 60) ? 2
                                                                      : (args.N == 1280 && sp_kb == 240) ? 3 : 4);
#endif

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
      ptx::mbar_init(&xf_bar[i], 2);
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
      FY_SP(long long sp_wait = 0; const long long sp_t0 = clock64();)
      for (int tile = worker; tile < num_tiles; tile += n_workers) {
        const int m0 = (tile / num_n) * TM + (int)cta_rank * BM;
        const int n0 = (tile % num_n) * BN + (int)cta_rank * (BN / NCTA);   // pair: this CTA's half of the B tile
        for (int s = 0; s < args.nseg; ++s) {
          const GemmSeg sg = args.seg[s];
          const int nparts = sg.a_lo ? 2 : 1;   // fused split segment: (A_hi, B_hi) then (A_lo, B_lo) per K block
          for (int kbp = 0; kbp < sg.nkb * nparts; ++kbp) {
            const int kb = sg.a_lo ? kbp >> 1 : kbp;
            const int a_c = sg.a_col + ((sg.a_lo && (kbp & 1)) ? sg.a_lo : 0) + kb * BKE;
            const int b_c = sg.b_col + ((sg.a_lo && (kbp & 1)) ? sg.b_lo : 0) + kb * BKE;
            FY_SP(const long long sp_a = clock64();)
            ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
            FY_SP(sp_wait += clock64() - sp_a;)
            uint8_t* a_dst = tiles + stage * C::kStageBytes;
            uint8_t* b_dst = a_dst + C::kABytes;
            if (PAIR) {
              // The peer's bytes may land on the leader's barrier before this expect_tx: harmless, the phase cannot
              // complete until this (the only) arrival, and the peer cannot run a ring revolution ahead because its
              // empty barrier is only released by the leader's MMA commit.
              if (cta_rank == 0) ptx::mbar_expect_tx(&full_bar[stage], C::kStageBytes * 2);
              ptx::tma_load_2d_pair(a_dst, &args.tmap_a, &full_bar[stage], a_c, m0 + sg.a_row_off);
              ptx::tma_load_2d_pair(b_dst, &args.tmap_b, &full_bar[stage], b_c, n0);
            } else {
              ptx::mbar_expect_tx(&full_bar[stage], C::kStageBytes);
              ptx::tma_load_2d(a_dst, &args.tmap_a, &full_bar[stage], a_c, m0 + sg.a_row_off);
              ptx::tma_load_2d(b_dst, &args.tmap_b, &full_bar[stage], b_c, n0);
              if (BN == 256)   // tensor-map boxes are at most 128 rows (shared with the pair kernel): second half
                ptx::tma_load_2d(b_dst + 128 * 128, &args.tmap_b, &full_bar[stage], b_c, n0 + 128);
            }
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
      }
      FY_SP(if (cta_rank == 0) { atomicAdd(&g_stall[sp_slot + 0], (unsigned long long)sp_wait);
                                 atomicAdd(&g_stall[sp_slot + 1], (unsigned long long)(clock64() - sp_t0)); })
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // The whole warp runs the loop (waits included); one elected lane issues each instruction (ptx::*_elect): a
    // tcgen05.mma under "if (lane == 0)" costs ~50 clk to issue, which the N = 32 / 64 tiles cannot hide.
    if (cta_rank == 0) {
      constexpr uint32_t idesc = TF32 ? ptx::make_idesc_tf32(TM, BN)
                                      : (OP == OP_F16 ? ptx::make_idesc_f16(TM, BN) : ptx::make_idesc_bf16(TM, BN));
      int stage = 0;
      uint32_t phase = 0;
      int as = 0;
      uint32_t aphase = 0;
      FY_SP(long long sp_wf = 0, sp_we = 0; const long long sp_t0 = clock64();)
      for (int tile = worker; tile < num_tiles; tile += n_workers) {
        FY_SP(const long long sp_a = clock64();)
        ptx::mbar_wait(&tempty_bar[as], aphase ^ 1);
        FY_SP(sp_we += clock64() - sp_a;)
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + as * BN;
        uint32_t accumulate = 0;
        for (int s = 0; s < args.nseg; ++s) {
          const int nkb = args.seg[s].nkb;
          if (args.seg[s].a_lo) {
            // fused split: slots (stage, stage + 1) hold (A_hi, B_hi) and (A_lo, B_lo); the ring depth is even and
            // every segment of the GEMM is fused, so stage is even here and the pair never wraps
            for (int kb = 0; kb < nkb; ++kb) {
              FY_SP(const long long sp_b = clock64();)
              ptx::mbar_wait(&full_bar[stage], phase);
              ptx::mbar_wait(&full_bar[stage + 1], phase);
              FY_SP(sp_wf += clock64() - sp_b;)
              ptx::tc_fence_after();
              const uint32_t s0 = ptx::smem_u32(tiles + stage * C::kStageBytes);
              const uint32_t s1 = s0 + C::kStageBytes;
              const uint64_t a_hi = ptx::make_desc_sw128(s0), b_hi = ptx::make_desc_sw128(s0 + C::kABytes);
              const uint64_t a_lo = ptx::make_desc_sw128(s1), b_lo = ptx::make_desc_sw128(s1 + C::kABytes);
#pragma unroll
              for (int pr = 0; pr < 3; ++pr) {
                const uint64_t da = pr == 2 ? a_lo : a_hi, db = pr == 1 ? b_lo : b_hi;
#pragma unroll
                for (int k = 0; k < BK / 16; ++k) {
                  if (PAIR) ptx::mma_bf16_ss_pair_elect(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
                  else ptx::mma_bf16_ss_elect(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
                  accumulate = 1;
                }
              }
              if (PAIR) { ptx::mma_commit_pair_elect(&empty_bar[stage]); ptx::mma_commit_pair_elect(&empty_bar[stage + 1]); }
              else { ptx::mma_commit_elect(&empty_bar[stage]); ptx::mma_commit_elect(&empty_bar[stage + 1]); }
              stage += 2;
              if (stage == C::kStages) { stage = 0; phase ^= 1; }
            }
            continue;
          }
          for (int kb = 0; kb < nkb; ++kb) {
            FY_SP(const long long sp_b = clock64();)
            ptx::mbar_wait(xform ? &xf_bar[stage] : &full_bar[stage], phase);   // transformed tiles imply landed tiles
            FY_SP(sp_wf += clock64() - sp_b;)
            ptx::tc_fence_after();
            const uint32_t a_addr = ptx::smem_u32(tiles + stage * C::kStageBytes);
            const uint64_t da = ptx::make_desc_sw128(a_addr);
            const uint64_t db = ptx::make_desc_sw128(a_addr + C::kABytes);
#pragma unroll
            for (int k = 0; k < BK / 16; ++k) {
              if (PAIR) ptx::mma_bf16_ss_pair_elect(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              else if (TF32) ptx::mma_tf32_ss_elect(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              else ptx::mma_bf16_ss_elect(d_tmem, da + 2 * k, db + 2 * k, idesc, accumulate);
              accumulate = 1;
            }
            if (PAIR) ptx::mma_commit_pair_elect(&empty_bar[stage]);
            else ptx::mma_commit_elect(&empty_bar[stage]);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
        if (PAIR) ptx::mma_commit_pair_elect(&tfull_bar[as]);
        else ptx::mma_commit_elect(&tfull_bar[as]);
        if (++as == C::kAccStages) { as = 0; aphase ^= 1; }
      }
      FY_SP(if (lane == 0) { atomicAdd(&g_stall[sp_slot + 2], (unsigned long long)sp_wf);
                             atomicAdd(&g_stall[sp_slot + 3], (unsigned long long)sp_we);
                             atomicAdd(&g_stall[sp_slot + 4], (unsigned long long)(clock64() - sp_t0)); })
    }
  } else if (warp < 4) {
    // ===================== A-tile transform (warps 2 and 3, only when GemmArgs::a_scale is set) =====================
    // The tile is [128 rows][128 bytes] with the 128-byte swizzle: the 16-byte chunk at position c' of row r holds the
    // logical chunk c' ^ (r & 7), i.e. A columns 8 (c' ^ (r & 7)) .. + 7 of the K block.  Thread t takes position t & 7
    // of the rows (t >> 3) + 8 i: its logical chunk -- and so its eight scale / shift pairs -- is the same for all
    // sixteen rows, loaded once per K block.
    if (xform) {
      const int t = threadIdx.x - 64;
      const int cpos = t & 7, r0 = t >> 3;
      const int clog = cpos ^ (r0 & 7);
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = worker; tile < num_tiles; tile += n_workers) {
        for (int s = 0; s < args.nseg; ++s) {
          const GemmSeg sg = args.seg[s];
          for (int kb = 0; kb < sg.nkb; ++kb) {
            const int k0 = sg.a_col + kb * BKE + clog * 8;
            const uint4 sc4 = __ldg(reinterpret_cast<const uint4*>(args.a_scale + k0));
            const uint4 sh4 = __ldg(reinterpret_cast<const uint4*>(args.a_shift + k0));
            const __half2* sc = reinterpret_cast<const __half2*>(&sc4);
            const __half2* sh = reinterpret_cast<const __half2*>(&sh4);
            const __half2 zero = __float2half2_rn(0.f);
            ptx::mbar_wait(&full_bar[stage], phase);
            uint8_t* a_tile = tiles + stage * C::kStageBytes + cpos * 16;
#pragma unroll 4
            for (int i = 0; i < 16; ++i) {
              uint4* p = reinterpret_cast<uint4*>(a_tile + (r0 + 8 * i) * 128);
              uint4 v = *p;
              __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
              for (int j = 0; j < 4; ++j) h[j] = __hmax2(__hfma2(h[j], sc[j], sh[j]), zero);
              *p = v;
            }
            asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");   // generic-proxy writes -> the MMA's async reads
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&xf_bar[stage]);
            if (++stage == C::kStages) { stage = 0; phase ^= 1; }
          }
        }
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
    FY_SP(long long sp_wt = 0; const long long sp_t0 = clock64();)
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
      FY_SP(const long long sp_a = clock64();)
      ptx::mbar_wait(&tfull_bar[as], aphase);
      FY_SP(sp_wt += clock64() - sp_a;)
      ptx::tc_fence_after();
      const uint32_t t_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + as * BN;
      if (BN == 256 && e.mode == EPI_QKV_ROPE) {
        // 64-column heads: columns [h*64, +32) pair with [h*64 + 32, +32).  Drained 16 + 16 columns at a time.
        const int which = n0 / e.rope_D;
        const int cbase = n0 - which * e.rope_D;
        __nv_bfloat16* ob = reinterpret_cast<__nv_bfloat16*>(which == 0 ? e.out : (which == 1 ? e.out2 : e.out3));
        const int qrow0 = m0 + q * 32 + (lane & ~3);   // first row of this lane's quad (see quad_transpose)
        const int qp = lane & 3;
        int pos = row_ok ? __ldg(e.rope_pos + row) : 0;
        if (pos < 0) pos = 0;
        const float* ct = e.rope_cos + (size_t)pos * 64;
        const float* sn = e.rope_sin + (size_t)pos * 64;
        // Sub-chunk sc covers pair indices [16 sc, 16 sc + 16); its 16 cos / 16 sin values are loaded once (issued
        // while the first head's TMEM load is in flight) and reused for every head this warp drains.
#pragma unroll 1
        for (int sc = 0; sc < 2; ++sc) {
          const int d0 = sc * 16;
          float cc[16], ss[16];
#pragma unroll 1
          for (int hh = 0; hh < kChunks / 2; ++hh) {
            const int ca = c_begin * 32 + hh * 64 + d0;     // tile-local column of the first half
            uint32_t va[16], vb[16];
            ptx::tmem_ld_32x16(t_row + ca, va);
            ptx::tmem_ld_32x16(t_row + ca + 32, vb);
            if (hh == 0 && which < 2) {
#pragma unroll
              for (int j4 = 0; j4 < 4; ++j4) {
                const float4 c4 = __ldg(reinterpret_cast<const float4*>(ct + d0) + j4);
                const float4 s4 = __ldg(reinterpret_cast<const float4*>(sn + d0) + j4);
                cc[4 * j4] = c4.x; cc[4 * j4 + 1] = c4.y; cc[4 * j4 + 2] = c4.z; cc[4 * j4 + 3] = c4.w;
                ss[4 * j4] = s4.x; ss[4 * j4 + 1] = s4.y; ss[4 * j4 + 2] = s4.z; ss[4 * j4 + 3] = s4.w;
              }
            }
            ptx::tmem_ld_wait();
            {   // rows past M compute on zeros and are not stored; the quad shuffles need every lane
              float xa[16], xb[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) {
                xa[j] = __uint_as_float(va[j]) + s_bias[ca + j];
                xb[j] = __uint_as_float(vb[j]) + s_bias[ca + 32 + j];
              }
              if (which < 2) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  const float a = xa[j], b = xb[j];
                  xa[j] = (a * cc[j] - b * ss[j]) * e.rope_scale;
                  xb[j] = (b * cc[j] + a * ss[j]) * e.rope_scale;
                }
              }
              // pieces [xa 0-7 | xa 8-15 | xb 0-7 | xb 8-15] of this row -> quad-transposed 2 x 32-byte runs per row
              const int pcol = ca + (qp & 1) * 8 + (qp >> 1) * 32;
              uint4 w[4];
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float* xs = (j >> 1) ? xb : xa;
                const int k = (j & 1) * 8;
                w[j].x = pack_bf16(xs[k], xs[k + 1]); w[j].y = pack_bf16(xs[k + 2], xs[k + 3]);
                w[j].z = pack_bf16(xs[k + 4], xs[k + 5]); w[j].w = pack_bf16(xs[k + 6], xs[k + 7]);
              }
              quad_transpose(w, lane);
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (qrow0 + j < args.M) *reinterpret_cast<uint4*>(ob + (size_t)(qrow0 + j) * e.ldo + cbase + pcol) = w[j];
              if (e.lo_col_off >= 0) {
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                  xa[j] -= __bfloat162float(__float2bfloat16_rn(xa[j]));
                  xb[j] -= __bfloat162float(__float2bfloat16_rn(xb[j]));
                }
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const float* xs = (j >> 1) ? xb : xa;
                  const int k = (j & 1) * 8;
                  w[j].x = pack_bf16(xs[k], xs[k + 1]); w[j].y = pack_bf16(xs[k + 2], xs[k + 3]);
                  w[j].z = pack_bf16(xs[k + 4], xs[k + 5]); w[j].w = pack_bf16(xs[k + 6], xs[k + 7]);
                }
                quad_transpose(w, lane);
#pragma unroll
                for (int j = 0; j < 4; ++j)
                  if (qrow0 + j < args.M)
                    *reinterpret_cast<uint4*>(ob + (size_t)(qrow0 + j) * e.ldo + e.lo_col_off + cbase + pcol) = w[j];
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
      // Affine rows (no row_idx, no remap: every S3 GEMM): global accesses are quad-transposed, see quad_transpose.
      const bool fast = !e.row_idx && !e.remap_q;
      const int qrow0 = m0 + q * 32 + (lane & ~3);   // first row of this lane's quad
      const int qp = lane & 3;                       // the 16-byte piece this lane moves for each quad row
#pragma unroll 1
      for (int c = c_begin; c < c_begin + kChunks; ++c) {
        uint32_t v[32];
        ptx::tmem_ld_32x32(t_row + c * 32, v);
        // Everything this chunk reads from global memory is requested while the TMEM load is in flight.
        const size_t col = (size_t)e.col_off + n0 + c * 32;
        float rs[32];
        if (fast && e.mode == EPI_F32_ADD) {   // transposed layout: rs[16 g + 4 j ..] = piece 4 g + qp of row qrow0 + j
          const float* ob = reinterpret_cast<const float*>(e.out) + col + qp * 4;
#pragma unroll
          for (int g = 0; g < 2; ++g)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              float4 w = make_float4(0.f, 0.f, 0.f, 0.f);
              if (qrow0 + j < args.M) w = *reinterpret_cast<const float4*>(ob + (size_t)(qrow0 + j) * e.ldo + g * 16);
              rs[16 * g + 4 * j] = w.x; rs[16 * g + 4 * j + 1] = w.y; rs[16 * g + 4 * j + 2] = w.z; rs[16 * g + 4 * j + 3] = w.w;
            }
        } else if (row_ok) {
          if (e.mode == EPI_F32_ADD) {
            const float4* rp = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(e.out) + orow * e.ldo + col);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 w = rp[j];
              rs[4 * j] = w.x; rs[4 * j + 1] = w.y; rs[4 * j + 2] = w.z; rs[4 * j + 3] = w.w;
            }
          } else if (e.resid) {
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
          } else if (e.resid_f32) {
            const float4* rp = reinterpret_cast<const float4*>(e.resid_f32 + orow * e.ldo + col);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float4 w = __ldg(rp + j);
              rs[4 * j] = w.x; rs[4 * j + 1] = w.y; rs[4 * j + 2] = w.z; rs[4 * j + 3] = w.w;
            }
          }
        }
        ptx::tmem_ld_wait();
        float x[32];
        // One uniform branch per stage, never per element: the 32 columns stay independent instruction streams.
        // Rows past M compute on zero accumulators (TMA zero-fills them) and are never stored.
#pragma unroll
        for (int j = 0; j < 32; ++j) x[j] = fmaf(__uint_as_float(v[j]), s_scale[c * 32 + j], s_bias[c * 32 + j]);
        if ((e.resid || e.resid_f32) && e.mode != EPI_F32_ADD && row_ok) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] += rs[j];
        }
        if (e.act == ACT_GELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = gelu_erf(x[j]);
        } else if (e.act == ACT_RELU) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.0f);
        }
        if (fast) {
          if (e.mode == EPI_BF16) {
            char* ob = reinterpret_cast<char*>(e.out) + (col * 2 + qp * 16);
            const size_t pitch = (size_t)e.ldo * 2;
            uint4 w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (e.out_f16) {
                w[j].x = pack_f16(x[8 * j], x[8 * j + 1]); w[j].y = pack_f16(x[8 * j + 2], x[8 * j + 3]);
                w[j].z = pack_f16(x[8 * j + 4], x[8 * j + 5]); w[j].w = pack_f16(x[8 * j + 6], x[8 * j + 7]);
              } else {
                w[j].x = pack_bf16(x[8 * j], x[8 * j + 1]); w[j].y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w[j].z = pack_bf16(x[8 * j + 4], x[8 * j + 5]); w[j].w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
              }
            }
            quad_transpose(w, lane);
#pragma unroll
            for (int j = 0; j < 4; ++j)
              if (qrow0 + j < args.M) *reinterpret_cast<uint4*>(ob + (size_t)(qrow0 + j) * pitch) = w[j];
            if (e.lo_col_off >= 0) {
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] -= __bfloat162float(__float2bfloat16_rn(x[j]));
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                w[j].x = pack_bf16(x[8 * j], x[8 * j + 1]); w[j].y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w[j].z = pack_bf16(x[8 * j + 4], x[8 * j + 5]); w[j].w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
              }
              quad_transpose(w, lane);
              char* ol = ob + ((size_t)e.lo_col_off - (size_t)e.col_off) * 2;
#pragma unroll
              for (int j = 0; j < 4; ++j)
                if (qrow0 + j < args.M) *reinterpret_cast<uint4*>(ol + (size_t)(qrow0 + j) * pitch) = w[j];
            }
          } else {
            if (e.round_tf32) {
#pragma unroll
              for (int j = 0; j < 32; ++j) x[j] = ptx::round_tf32(x[j]);
            }
            float* ob = reinterpret_cast<float*>(e.out) + col + qp * 4;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              uint4 w[4];
#pragma unroll
              for (int j = 0; j < 4; ++j)
                w[j] = make_uint4(__float_as_uint(x[16 * g + 4 * j]), __float_as_uint(x[16 * g + 4 * j + 1]),
                                  __float_as_uint(x[16 * g + 4 * j + 2]), __float_as_uint(x[16 * g + 4 * j + 3]));
              quad_transpose(w, lane);
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                float4 r = make_float4(__uint_as_float(w[j].x), __uint_as_float(w[j].y), __uint_as_float(w[j].z),
                                       __uint_as_float(w[j].w));
                if (e.mode == EPI_F32_ADD) {
                  r.x += rs[16 * g + 4 * j]; r.y += rs[16 * g + 4 * j + 1];
                  r.z += rs[16 * g + 4 * j + 2]; r.w += rs[16 * g + 4 * j + 3];
                }
                if (qrow0 + j < args.M) *reinterpret_cast<float4*>(ob + (size_t)(qrow0 + j) * e.ldo + g * 16) = r;
              }
            }
          }
        } else if (row_ok) {
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
                w.x = pack_f16(x[8 * j], x[8 * j + 1]); w.y = pack_f16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_f16(x[8 * j + 4], x[8 * j + 5]); w.w = pack_f16(x[8 * j + 6], x[8 * j + 7]);
              } else {
                w.x = pack_bf16(x[8 * j], x[8 * j + 1]); w.y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_bf16(x[8 * j + 4], x[8 * j + 5]); w.w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
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
                w.x = pack_bf16(x[8 * j], x[8 * j + 1]); w.y = pack_bf16(x[8 * j + 2], x[8 * j + 3]);
                w.z = pack_bf16(x[8 * j + 4], x[8 * j + 5]); w.w = pack_bf16(x[8 * j + 6], x[8 * j + 7]);
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
              for (int j = 0; j < 8; ++j)
                dst[j] = make_float4(rs[4 * j] + x[4 * j], rs[4 * j + 1] + x[4 * j + 1], rs[4 * j + 2] + x[4 * j + 2],
                                     rs[4 * j + 3] + x[4 * j + 3]);
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
    FY_SP(if (threadIdx.x == 128 && cta_rank == 0) {
      atomicAdd(&g_stall[sp_slot + 5], (unsigned long long)sp_wt);
      atomicAdd(&g_stall[sp_slot + 6], (unsigned long long)(clock64() - sp_t0));
      atomicAdd(&g_stall[sp_slot + 7], 1ull);
    })
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
  // function attributes belong to a device's context: configured once per device, not once per process
  static bool configured[64] = {false};
  int dev = 0;
  FY_CUDA(cudaGetDevice(&dev));
  if (!configured[dev & 63]) {
    FY_CUDA(cudaFuncSetAttribute(gemm_kernel<BN, OP, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    configured[dev & 63] = true;
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
  int n_fused = 0;
  for (int i = 0; i < a.nseg; ++i) n_fused += a.seg[i].a_lo != 0;
  FY_CHECK(n_fused == 0 || (n_fused == a.nseg && a.op == OP_BF16 && gemm_block_n(a.N) == 256), FY_ERR_ARG,
           "fused split segments need every segment fused, bf16 operands and a 256-wide tile (N=%d)", a.N);
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

#ifdef FY_STALLPROF
// Debug build only: copy the role-stall counters to the host (40 x u64) and optionally clear them.
extern "C" int fy_debug_gemm_stalls(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out, fy::g_stall, sizeof(unsigned long long) * 40) != cudaSuccess) return 1;
  if (reset) {
    unsigned long long z[40] = {0};
    if (cudaMemcpyToSymbol(fy::g_stall, z, sizeof(z)) != cudaSuccess) return 1;
  }
  return 0;
}
#endif
