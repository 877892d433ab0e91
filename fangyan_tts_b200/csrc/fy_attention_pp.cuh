// Variable-length self-attention on tcgen05, two-group ping-pong (S3 encoder, row A9).  Default kernel of both encoder
// precisions; FY_ATTN_TC=1 selects the previous single-group tcgen05 kernel (fy_attention_tc.cuh, split-bf16 only) and
// FY_ATTN_TC=0 the mma.sync kernel (fy_attention.cuh).
//
// Why: the single-group kernel walks one serial chain per key tile (S load -> row maximum exchanged between four
// threads -> ex2 -> pack -> P store -> signal) with all sixteen softmax warps in lock-step, ~2,400 clk per tile against
// 768 clk of tensor work (profiles/r01_ncu_attention_tc.md).  Here one CTA still owns 128 query rows of one utterance
// and four heads, but the heads are split between TWO independent softmax groups (A: heads h0, h0+1; B: h0+2, h0+3):
//   * ONE thread per query row (4 warps per group): a row's 64 scores live in one thread's registers, so the row maximum
//     and sum need no exchange, no shared memory and no named barrier, and the instruction stream has 64-way ILP;
//   * each group has its own S (double-buffered by tile), O and Q in tensor memory -- 256 of the 512 columns each -- and P
//     is stored over the S tile it was computed from (hi halves in its first 32 columns, lo halves in the last 32);
//   * each group has its own MMA-issuing warp (S_G(i+1) then PV_G(i)); the two groups' products interleave in the tensor
//     pipe, so one group's softmax runs under the other group's products;
//   * tcgen05.ld / st move a whole row at a time (two .x32 each way per tile).
// Everything else follows the single-group kernel: K / V tiles arrive by TMA into a ring (K and V tracked apart), both
// products take their A operand from tensor memory (Q, P) and B from shared memory (K K-major, V MN-major as it lands),
// split-bf16 = three products hi*hi + hi*lo + lo*hi into one fp32 accumulator, O accumulates in TMEM over a head with a
// lazily moved running maximum (2^slack), a head is normalised and stored under the next head's first products.
// Both halves of P are TRUNCATED bf16 (hi = upper half-word of P, lo = upper half-word of P - hi: PRMT, no
// conversion-unit instruction next to the ex2 stream); hi + lo carries P to 2^-16.  W = 1 (single-pass bf16 encoder mode): one product, rounded P.
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace attn_pp {

constexpr int kBQ = 128, kBK = 64, kDH = 64;
constexpr int kGroups = 2, kHeadsPerGroup = 2, kHeadsPerCta = kGroups * kHeadsPerGroup;
constexpr int kSoftmaxWarps = 4 * kGroups, kSoftmaxThreads = 32 * kSoftmaxWarps;
constexpr int kThreads = kSoftmaxThreads + 96;       // + one issuing warp per group + producer warp
constexpr int kKVBytes = kBK * 128;                  // one [64][64] bf16 tile
template <int W> constexpr int stage_bytes() { return 2 * W * kKVBytes; }   // K (hi, lo), V (hi, lo)
template <int W> constexpr int stages() { return W == 2 ? 5 : 8; }
template <int W> constexpr int q_bytes() { return W * kBQ * 128; }          // one group's Q tile: [128][64] bf16 hi (, lo)
constexpr int kCtrlBytes = 512;
template <int W> constexpr int smem_bytes() { return stages<W>() * stage_bytes<W>() + kGroups * q_bytes<W>() + kCtrlBytes + 1024; }
static_assert(smem_bytes<2>() <= 227 * 1024 && smem_bytes<1>() > 114 * 1024, "one CTA per SM (it owns all of TMEM)");
// Tensor-memory columns of group G (base 256 G): S/P buffer 0 [0,64), S/P buffer 1 [64,128), O [128,192), Q [192,256)
// (Q: 32 columns of bf16 pairs for the hi halves, 32 for the lo halves).
constexpr uint32_t kTmemCols = 512, kGroupCols = 256, kColO = 128, kColQ = 192;
constexpr float kRescaleSlack = 8.0f;

// kind::f16 instruction descriptor, bf16 A (K-major) x bf16 B, fp32 D; b_mn = 1: B is MN-major.
__host__ __device__ constexpr uint32_t idesc(uint32_t m, uint32_t n, uint32_t b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (b_mn << 16) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2_rn(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
// (a, b) -> bf16 pair by truncation: the upper half-words of b and a (a in the low half-word).
__device__ __forceinline__ uint32_t pack2_trunc(float a, float b) {
  return __byte_perm(__float_as_uint(a), __float_as_uint(b), 0x7632);
}

// Transpose the four 16-byte pieces held by each lane across its lane quad: afterwards lane (Q + q) holds piece q of the
// rows of lanes Q .. Q + 3 in w[0..3] (same exchange as the GEMM epilogue's).
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
  for (int a = 0; a < 4; a += 2) {
    const uint4 r = shfl_xor_u4(b0 ? w[a] : w[a + 1], 1);
    if (b0) w[a] = r; else w[a + 1] = r;
  }
#pragma unroll
  for (int a = 0; a < 2; ++a) {
    const uint4 r = shfl_xor_u4(b1 ? w[a] : w[a + 2], 2);
    if (b1) w[a] = r; else w[a + 2] = r;
  }
}

__device__ __forceinline__ void tmem_st_32x32(uint32_t taddr, const uint32_t* v) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x32.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, "
      "%17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31, %32};\n" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15]), "r"(v[16]), "r"(v[17]), "r"(v[18]),
      "r"(v[19]), "r"(v[20]), "r"(v[21]), "r"(v[22]), "r"(v[23]), "r"(v[24]), "r"(v[25]), "r"(v[26]), "r"(v[27]),
      "r"(v[28]), "r"(v[29]), "r"(v[30]), "r"(v[31])
      : "memory");
}
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

struct Params {
  CUtensorMap tm_k;    // [rows, W D] bf16, box 64 cols x 64 rows
  CUtensorMap tm_v;
  CUtensorMap tm_q;    // same tensor geometry over q, box 64 cols x 128 rows
  const __nv_bfloat16* q;
  __nv_bfloat16* out;
  int ld, lo_off, rows;
  float slack;
  const int2* tiles;   // (utterance, q0) per blockIdx.x, q0 a multiple of 128
  const int* utt_P;
  const int* utt_L;
};

#ifdef FY_STALLPROF
// Debug build only (tools/attn_stalls.py): cycles spent per role and phase, summed over CTAs.
//  issuer (one lane): 0 q_full 1 k_full 2 v_full 3 p_full 4 o_empty 5 life 6 CTAs
//  softmax thread 0 : 8 s_full 9 S load 10 max + ex2 11 pack + P store 12 o_full (head end) 13 head epilogue 14 life
//                     15 Q staging 19 prologue (launch -> first barrier passed)
//  producer         : 16 k_empty 17 v_empty 18 life
__device__ unsigned long long g_pp_stall[32];
#define PP_T(slot, ...) { const long long pp_a = clock64(); __VA_ARGS__; pw[slot] += clock64() - pp_a; }
#define PP_P(...) __VA_ARGS__
#else
#define PP_T(slot, ...) { __VA_ARGS__; }
#define PP_P(...)
#endif

template <int W>
__global__ void __maxnreg__(168) attention_pp_kernel(const __grid_constant__ Params p) {
  constexpr int kStages = stages<W>();
  constexpr int kStageBytes = stage_bytes<W>();
  extern __shared__ uint8_t smem_raw[];
  const uint32_t align_pad = (1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u;
  uint8_t* sKV = smem_raw + align_pad;               // stage s: K hi, (K lo,) V hi (, V lo)
  uint8_t* sQ = sKV + kStages * kStageBytes;         // group G: Q hi (, lo) of the head it stages next, 128-byte swizzle
  uint64_t* bars = reinterpret_cast<uint64_t*>(sQ + kGroups * q_bytes<W>());
  uint64_t* k_full = bars;                    // [kStages]
  uint64_t* k_empty = bars + kStages;         // [kStages]
  uint64_t* v_full = bars + 2 * kStages;      // [kStages]
  uint64_t* v_empty = bars + 3 * kStages;     // [kStages]
  uint64_t* gb = bars + 4 * kStages;          // per group G, 8 barriers at gb + 8 G:
  //   +0,+1 s_full[buffer]   +2,+3 p_full[buffer]   +4 q_full   +5 o_full   +6 o_empty   +7 all_done
  uint64_t* qs_full = gb + 16;                // [kGroups] a head's Q tile has landed in shared memory (TMA bytes)
  uint64_t* qs_empty = qs_full + kGroups;     // [kGroups] ... and has been copied out by the group's four warps
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(qs_empty + kGroups);
  static_assert((4 * 8 + 16 + 4 + 1) * 8 <= kCtrlBytes, "barrier block");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  PP_P(long long pw[26] = {0}; const long long pp_t0 = clock64();)
  const int2 tl = p.tiles[blockIdx.x];
  const int u = tl.x, q0 = tl.y;
  const int h0 = blockIdx.y * kHeadsPerCta;
  const int P = p.utt_P[u], L = p.utt_L[u];
  const int nkt = (L + kBK - 1) / kBK;
  const int N = nkt * kHeadsPerGroup;              // tile units per group: i = g * nkt + t

  // Softmax thread = one query row of one group.
  const int G = warp >> 2, wq = warp & 3;
  const int row = wq * 32 + lane;

  if (threadIdx.x == kSoftmaxThreads) {
    ptx::prefetch_tmap(&p.tm_k);
    ptx::prefetch_tmap(&p.tm_v);
    ptx::prefetch_tmap(&p.tm_q);
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(&k_full[i], 1);
      ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1);
      ptx::mbar_init(&v_empty[i], 1);
    }
    for (int g = 0; g < kGroups; ++g) {
      uint64_t* b = gb + 8 * g;
      ptx::mbar_init(&b[0], 1); ptx::mbar_init(&b[1], 1);      // s_full
      ptx::mbar_init(&b[2], 4); ptx::mbar_init(&b[3], 4);      // p_full: one arrival per softmax warp of the group
      ptx::mbar_init(&b[4], 4);                                // q_full
      ptx::mbar_init(&b[5], 1);                                // o_full
      ptx::mbar_init(&b[6], 4);                                // o_empty
      ptx::mbar_init(&b[7], 1);                                // all_done
      ptx::mbar_init(&qs_full[g], 1);
      ptx::mbar_init(&qs_empty[g], 4);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc<kTmemCols>(tmem_holder);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_holder;
  PP_P(pw[19] = clock64() - pp_t0;)

  if (warp == kSoftmaxWarps + kGroups) {
    if (lane == 0) {
      // ------------------------------------------------ TMA producer: Q tiles (one per head and group, through a
      // single shared-memory buffer per group that the softmax warps copy into tensor memory) and the K / V tiles of
      // unit n = 2 i + G, in issue order
      auto load_q_tile = [&](int grp, int g) {
        uint8_t* d = sQ + grp * q_bytes<W>();
        const int col = (h0 + grp * kHeadsPerGroup + g) * kDH;
        ptx::mbar_expect_tx(&qs_full[grp], q_bytes<W>());
        ptx::tma_load_2d(d, &p.tm_q, &qs_full[grp], col, P + q0);
        if (W == 2) ptx::tma_load_2d(d + kBQ * 128, &p.tm_q, &qs_full[grp], col + p.lo_off, P + q0);
      };
      load_q_tile(0, 0);
      load_q_tile(1, 0);
      int q_next[kGroups] = {1, 1};            // next head whose Q tile is still to be fetched, per group
      for (int n = 0; n < 2 * N; ++n) {
        // a group's buffer is refilled as soon as its warps have copied the previous head out (never waited for here)
#pragma unroll
        for (int gq = 0; gq < kGroups; ++gq)
          if (q_next[gq] < kHeadsPerGroup && ptx::mbar_try_wait(&qs_empty[gq], (q_next[gq] - 1) & 1)) load_q_tile(gq, q_next[gq]++);
        const int grp = n & 1, i = n >> 1;
        const int g = i / nkt, t = i - g * nkt;
        const int col = (h0 + grp * kHeadsPerGroup + g) * kDH;
        const int st = n % kStages;
        const uint32_t ph = ((n / kStages) - 1) & 1;
        uint8_t* d = sKV + st * kStageBytes;
        if (n >= kStages) PP_T(16, ptx::mbar_wait(&k_empty[st], ph))
        ptx::mbar_expect_tx(&k_full[st], W * kKVBytes);
        ptx::tma_load_2d(d, &p.tm_k, &k_full[st], col, P + t * kBK);
        if (W == 2) ptx::tma_load_2d(d + kKVBytes, &p.tm_k, &k_full[st], col + p.lo_off, P + t * kBK);
        if (n >= kStages) PP_T(17, ptx::mbar_wait(&v_empty[st], ph))
        ptx::mbar_expect_tx(&v_full[st], W * kKVBytes);
        ptx::tma_load_2d(d + W * kKVBytes, &p.tm_v, &v_full[st], col, P + t * kBK);
        if (W == 2) ptx::tma_load_2d(d + 3 * kKVBytes, &p.tm_v, &v_full[st], col + p.lo_off, P + t * kBK);
      }
#pragma unroll
      for (int gq = 0; gq < kGroups; ++gq)
        while (q_next[gq] < kHeadsPerGroup) {
          ptx::mbar_wait(&qs_empty[gq], (q_next[gq] - 1) & 1);
          load_q_tile(gq, q_next[gq]++);
        }
      PP_P(atomicAdd(&g_pp_stall[16], (unsigned long long)pw[16]); atomicAdd(&g_pp_stall[17], (unsigned long long)pw[17]);
           atomicAdd(&g_pp_stall[18], (unsigned long long)(clock64() - pp_t0));)
    }
  } else if (warp >= kSoftmaxWarps) {
    // ------------------------------------------------ the MMAs of group (warp - kSoftmaxWarps): whole warp converged, one
    // elected lane issues.  One issuing warp per group, on different schedulers: an M128 N64 K16 instruction needs ~12
    // scalar instructions around it (descriptor arithmetic, moves to uniform registers, elect), which one warp sharing
    // its scheduler with two softmax warps could not issue faster than ~70 clk per MMA against the 32 clk the tensor
    // pipe needs.  The groups' products interleave in the pipe in whatever order the two warps get there.
    const int my = warp - kSoftmaxWarps;
    constexpr uint32_t id_s = idesc(kBQ, kBK, 0);   // S = Q K^T: both operands K-major
    constexpr uint32_t id_o = idesc(kBQ, kDH, 1);   // O = P V  : V is MN-major ([key][d] rows of 128 bytes)
    auto issue_s = [&](int grp, int i) {
      const int n = 2 * i + grp;
      uint64_t* b = gb + 8 * grp;
      const int g = i / nkt;
      if (i - g * nkt == 0) PP_T(0, ptx::mbar_wait(&b[4], g & 1))       // the head's Q is in tensor memory
      PP_T(1, ptx::mbar_wait(&k_full[n % kStages], (n / kStages) & 1))
      ptx::tc_fence_after();
      const uint32_t base = tmem + grp * kGroupCols;
      const uint32_t qa = base + kColQ;
      const uint32_t kb = ptx::smem_u32(sKV + (n % kStages) * kStageBytes);
      const uint32_t d = base + (i & 1) * kBK;
      uint32_t acc = 0;
#pragma unroll
      for (int pr = 0; pr < (W == 2 ? 3 : 1); ++pr) {
        const uint32_t ta = qa + (pr == 2 ? 32 : 0);
        const uint64_t db = ptx::make_desc_sw128(kb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
        for (int k = 0; k < kDH / 16; ++k) {
          ptx::mma_bf16_ts_elect(d, ta + 8 * k, db + 2 * k, id_s, acc);
          acc = 1;
        }
      }
      ptx::mma_commit_elect(&b[i & 1]);
      ptx::mma_commit_elect(&k_empty[n % kStages]);
    };
    auto issue_pv = [&](int grp, int i) {
      const int n = 2 * i + grp;
      uint64_t* b = gb + 8 * grp;
      const int g = i / nkt, t = i - g * nkt;
      PP_T(2, ptx::mbar_wait(&v_full[n % kStages], (n / kStages) & 1))
      PP_T(3, ptx::mbar_wait(&b[2 + (i & 1)], (i >> 1) & 1))             // P of this unit is in tensor memory
      if (t == 0 && g > 0) PP_T(4, ptx::mbar_wait(&b[6], (g - 1) & 1))   // the previous head's O has been read out
      ptx::tc_fence_after();
      const uint32_t base = tmem + grp * kGroupCols;
      const uint32_t vb = ptx::smem_u32(sKV + (n % kStages) * kStageBytes) + W * kKVBytes;
      const uint32_t pa = base + (i & 1) * kBK;
      const uint32_t d = base + kColO;
      uint32_t acc = t > 0;
#pragma unroll
      for (int pr = 0; pr < (W == 2 ? 3 : 1); ++pr) {
        const uint32_t ta = pa + (pr == 2 ? 32 : 0);
        const uint64_t db = ptx::make_desc_sw128(vb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
        for (int k = 0; k < kBK / 16; ++k) {
          ptx::mma_bf16_ts_elect(d, ta + 8 * k, db + 128 * k, id_o, acc);
          acc = 1;
        }
      }
      ptx::mma_commit_elect(&b[5]);
      ptx::mma_commit_elect(&v_empty[n % kStages]);
      if (i == N - 1) ptx::mma_commit_elect(&b[7]);
    };
    issue_s(my, 0);
    for (int i = 0; i < N; ++i) {
      if (i + 1 < N) issue_s(my, i + 1);
      issue_pv(my, i);
    }
    PP_P(if (lane == 0) {
      for (int k = 0; k < 5; ++k) atomicAdd(&g_pp_stall[k], (unsigned long long)pw[k]);
      atomicAdd(&g_pp_stall[5], (unsigned long long)(clock64() - pp_t0));
      if (my == 0) atomicAdd(&g_pp_stall[6], 1ull);
    })
  } else {
    // ------------------------------------------------ softmax: one thread per query row, group G
    uint64_t* b = gb + 8 * G;
    const uint32_t t_row = tmem + (static_cast<uint32_t>(wq * 32) << 16) + G * kGroupCols;
    const float kLog2e = 1.4426950408889634f;
    const int hg = h0 + G * kHeadsPerGroup;
    // Head g's Q row: shared memory (TMA tile, 128-byte swizzle: 16-byte chunk c of row r sits at chunk c ^ (r & 7)) ->
    // registers -> this row's 32 (+32) Q columns of tensor memory.  The buffer is released as soon as the registers
    // hold the row, so the producer refills it with the next head's tile long before that one is needed.
    auto stage_q = [&](int g) {
      ptx::mbar_wait(&qs_full[G], g & 1);
      const uint8_t* qrow = sQ + G * q_bytes<W>() + row * 128;
#pragma unroll
      for (int w = 0; w < W; ++w) {
        uint32_t qreg[32];
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          const uint4 x = *reinterpret_cast<const uint4*>(qrow + w * (kBQ * 128) + ((c ^ (row & 7)) << 4));
          qreg[4 * c] = x.x; qreg[4 * c + 1] = x.y; qreg[4 * c + 2] = x.z; qreg[4 * c + 3] = x.w;
        }
        tmem_st_32x32(t_row + kColQ + 32 * w, qreg);
      }
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&qs_empty[G]);
    };
    stage_q(0);
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&b[4]);

    // Head epilogue: the caller has made sure the head's last P V has completed.
    auto finish_head = [&](int g, float l) {
      ptx::tc_fence_after();
      PP_P(long long pp_e = clock64();)
      uint32_t ov[64];
      tmem_ld_32x32(t_row + kColO, ov);
      tmem_ld_32x32(t_row + kColO + 32, ov + 32);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&b[6]);
      PP_P(pw[20] += clock64() - pp_e; pp_e = clock64();)
      const int col = (hg + g) * kDH;
      // Normalise, split to bf16 hi | lo.  A lane holds one row (128 + 128 bytes); writing it as it stands would touch
      // 32 different rows per store instruction (32 L1 wavefronts of 16 bytes).  The 16-byte pieces are transposed across
      // each lane quad first, four at a time, so that a store instruction covers 8 rows x 64 contiguous bytes.
      const float inv = 1.0f / l;
      const int rq = lane & ~3, pq = lane & 3;          // after the transpose: piece pq of rows rq .. rq + 3
      __nv_bfloat16* obase = p.out + (size_t)(P + q0 + wq * 32 + rq) * p.ld + col + pq * 8;
      bool rok[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) rok[j] = q0 + wq * 32 + rq + j < L;
#pragma unroll
      for (int blk = 0; blk < 2; ++blk) {               // 4 pieces = 32 head-dim values per block
        uint4 hi[4], lo[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float a[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) a[j] = __uint_as_float(ov[32 * blk + 8 * c + j]) * inv;
          if (W == 2) {   // hi truncated (PRMT), lo = bf16(x - hi) rounded: the pair is what the next GEMM's split product needs
            hi[c].x = pack2_trunc(a[0], a[1]); hi[c].y = pack2_trunc(a[2], a[3]); hi[c].z = pack2_trunc(a[4], a[5]); hi[c].w = pack2_trunc(a[6], a[7]);
          } else {
            hi[c].x = pack2_rn(a[0], a[1]); hi[c].y = pack2_rn(a[2], a[3]); hi[c].z = pack2_rn(a[4], a[5]); hi[c].w = pack2_rn(a[6], a[7]);
          }
          if (W == 2) {
            lo[c].x = pack2_rn(a[0] - __uint_as_float(hi[c].x << 16), a[1] - __uint_as_float(hi[c].x & 0xffff0000u));
            lo[c].y = pack2_rn(a[2] - __uint_as_float(hi[c].y << 16), a[3] - __uint_as_float(hi[c].y & 0xffff0000u));
            lo[c].z = pack2_rn(a[4] - __uint_as_float(hi[c].z << 16), a[5] - __uint_as_float(hi[c].z & 0xffff0000u));
            lo[c].w = pack2_rn(a[6] - __uint_as_float(hi[c].w << 16), a[7] - __uint_as_float(hi[c].w & 0xffff0000u));
          }
        }
        quad_transpose(hi, lane);
        if (W == 2) quad_transpose(lo, lane);
        PP_P(pw[21] += clock64() - pp_e; pp_e = clock64();)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (rok[j]) {
            __nv_bfloat16* dst = obase + (size_t)j * p.ld + blk * 32;
            *reinterpret_cast<uint4*>(dst) = hi[j];
            if (W == 2) *reinterpret_cast<uint4*>(dst + p.lo_off) = lo[j];
          }
        }
        PP_P(pw[22] += clock64() - pp_e; pp_e = clock64();)
      }
      // The guard row after the utterance is kept at zero (it is read, times P = 0, as a key / value row of the last
      // tile, so it must stay finite through all layers).
      if (q0 + row == L || (row == kBQ - 1 && q0 + kBQ == L)) {
        __nv_bfloat16* dst = p.out + (size_t)(P + L) * p.ld + col;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          *reinterpret_cast<uint4*>(dst + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
          if (W == 2) *reinterpret_cast<uint4*>(dst + p.lo_off + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
    };

    float m = -INFINITY, l = 0.f, l_prev = 0.f;
    int g = 0, t = 0;
#pragma unroll 1
    for (int i = 0; i < N; ++i) {
      const bool last_tile = t == nkt - 1;
      const bool next_q = last_tile && g + 1 < kHeadsPerGroup;      // the next head's Q replaces this one's after its last S
      if (t == 0) { l_prev = l; l = 0.f; }
      PP_T(8, ptx::mbar_wait(&b[i & 1], (i >> 1) & 1))
      ptx::tc_fence_after();
      uint32_t sv[64];
      const uint32_t sp = t_row + (i & 1) * kBK;
      PP_P(long long pp_c = clock64();)
      tmem_ld_32x32(sp, sv);
      tmem_ld_32x32(sp + 32, sv + 32);
      ptx::tmem_ld_wait();
      PP_P(pw[9] += clock64() - pp_c; pp_c = clock64();)
      if (next_q) {
        // s_full of the head's last tile: every S product of the head has completed, Q may be replaced
        stage_q(g + 1);
        ptx::tmem_st_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&b[4]);
      }
      PP_P(pw[15] += clock64() - pp_c; pp_c = clock64();)
      float* s = reinterpret_cast<float*>(sv);
      if ((t + 1) * kBK > L) {           // the head's last tile may be partial: keys >= L are masked
#pragma unroll
        for (int j = 0; j < kBK; ++j)
          if (t * kBK + j >= L) s[j] = -INFINITY;
      }
      float mx8[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) mx8[j] = fmaxf(fmaxf(s[j], s[8 + j]), fmaxf(s[16 + j], s[24 + j]));
#pragma unroll
      for (int j = 0; j < 8; ++j) mx8[j] = fmaxf(mx8[j], fmaxf(fmaxf(s[32 + j], s[40 + j]), fmaxf(s[48 + j], s[56 + j])));
      const float mx = fmaxf(fmaxf(fmaxf(mx8[0], mx8[1]), fmaxf(mx8[2], mx8[3])),
                             fmaxf(fmaxf(mx8[4], mx8[5]), fmaxf(mx8[6], mx8[7]))) * kLog2e;
      PP_P(pw[23] += clock64() - pp_c;)
      if (t == 0) {
        m = mx;                          // a head's first P V overwrites O: nothing to rescale
      } else {
        const bool need = mx > m + p.slack;
        if (__any_sync(0xffffffffu, need)) {
          // rare: rescale this row's O in tensor memory (P_{i-1} V_{i-1} must have landed; P_i V_i cannot have been
          // issued yet, so o_full is in phase i - 1 or i and the parity is unambiguous)
          ptx::mbar_wait(&b[5], (i - 1) & 1);
          ptx::tc_fence_after();
          const float corr = need ? ex2(m - mx) : 1.0f;
#pragma unroll
          for (int hlf = 0; hlf < 2; ++hlf) {
            uint32_t ov[32];
            tmem_ld_32x32(t_row + kColO + 32 * hlf, ov);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) ov[j] = __float_as_uint(__uint_as_float(ov[j]) * corr);
            tmem_st_32x32(t_row + kColO + 32 * hlf, ov);
          }
          l *= corr;
          if (need) m = mx;
        }
      }
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < kBK; ++j) {
        s[j] = ex2(fmaf(s[j], kLog2e, -m));
        sum4[j & 3] += s[j];
      }
      l += (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
      PP_P(pw[10] += clock64() - pp_c; pp_c = clock64();)
      // P row over the S tile it came from: hi pairs in columns [0, 32), lo pairs in [32, 64) (even key in the low
      // half-word).  Every S value of the row is already in registers.
      {
        if (W == 2) {
          // hi = the truncated bf16 of P, lo = the truncated bf16 of (P - hi): hi + lo carries P to 2^-16 with two PRMT,
          // two LOP and two FADD per pair and nothing on the conversion / MUFU unit, which the ex2 stream saturates
          uint32_t pk[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) pk[j] = pack2_trunc(s[2 * j], s[2 * j + 1]);
          tmem_st_32x32(sp, pk);
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float a = s[2 * j], c = s[2 * j + 1];
            pk[j] = pack2_trunc(a - __uint_as_float(__float_as_uint(a) & 0xffff0000u),
                                c - __uint_as_float(__float_as_uint(c) & 0xffff0000u));
          }
          tmem_st_32x32(sp + 32, pk);
        } else {
          uint32_t ph[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) ph[j] = pack2_rn(s[2 * j], s[2 * j + 1]);
          tmem_st_32x32(sp, ph);
        }
      }
      // head g - 1 is complete once P_{i-1} V_{i-1} has landed; waited for BEFORE this tile's p_full arrival (see the
      // rescale note above for why the parity is unambiguous only then)
      const bool head_done = t == 0 && i > 0;
      if (head_done) PP_T(12, ptx::mbar_wait(&b[5], (i - 1) & 1))
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&b[2 + (i & 1)]);
      PP_P(pw[11] += clock64() - pp_c;)
      if (head_done) PP_T(13, finish_head(g - 1, l_prev))
      if (++t == nkt) { t = 0; ++g; }
    }
    PP_T(12, ptx::mbar_wait(&b[7], 0))
    PP_T(13, finish_head(kHeadsPerGroup - 1, l))
    PP_P(if (threadIdx.x == 0) {
      for (int k = 8; k < 16; ++k) atomicAdd(&g_pp_stall[k], (unsigned long long)pw[k]);
      for (int k = 20; k < 24; ++k) atomicAdd(&g_pp_stall[k], (unsigned long long)pw[k]);
      atomicAdd(&g_pp_stall[14], (unsigned long long)(clock64() - pp_t0));
      atomicAdd(&g_pp_stall[19], (unsigned long long)pw[19]);
    })
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<kTmemCols>(tmem);
  }
}

}  // namespace attn_pp
}  // namespace fy
