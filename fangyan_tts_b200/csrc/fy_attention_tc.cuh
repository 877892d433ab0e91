// Variable-length self-attention on tcgen05 (S3 encoder, row A9), split-bf16 mode.  FY_ATTN_TC=0 selects the
// mma.sync kernel of fy_attention.cuh instead.
//
// One CTA = 128 query rows of one utterance, walked over kHeadsPerCta heads.  Both GEMMs of the flash loop run on the
// 5th-gen tensor cores with fp32 accumulators in TMEM, and BOTH take their A operand from tensor memory (".ts" form):
// Q is constant over a head's key tiles and P is produced by the softmax threads, so neither ever touches shared
// memory -- the only operand bytes the shared-memory port serves are the K and V tiles (2 KB per MMA instead of 6 KB,
// which is what held the all-shared-memory version at ~100 clk per M128 N64 K16 instruction against a 32 clk floor).
//   warp 9, one thread : TMA loads of the K / V 64-key tiles (hi and lo halves) into a kStages ring
//   warp 8, one thread : every tcgen05.mma
//                        S_t = Q K_t^T   (M 128, N 64, K 64; three products hi*hi + hi*lo + lo*hi)
//                        O_t = P_t V_t   (M 128, N 64, K 64 keys; V is consumed MN-major, i.e. exactly as the
//                                         [key][d] tile TMA delivers it -- no transpose anywhere)
//   warps 0-7, 2 threads per query row (32 keys each): tcgen05.ld of the S row, running max / sum in registers (the
//                        two halves meet through shared memory once per tile, no shuffles), P = 2^(s log2e - m)
//                        packed to bf16 hi | lo and stored to TMEM with tcgen05.st, then O_{t-1} is read back from
//                        TMEM and folded into the register accumulator.  The same threads stage Q: each reads its
//                        64 + 64 bytes of the next head's query row from global memory at the start of a head and
//                        stores them to TMEM under the first tile's softmax.
// S, P and Q are double-buffered in TMEM, so S_{t+1} is computed while the softmax of tile t runs.
// Rows past the utterance inside a tile belong to the next utterance or a guard row: keys are masked, queries are
// computed and not stored, V rows meet P = 0; all of it is finite data of the same buffers (the kernel keeps the
// guard rows of its own output at zero for that reason).
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace attn_tc {

constexpr int kBQ = 128, kBK = 64, kDH = 64, kThreads = 320, kStages = 4;
constexpr int kSoftmaxThreads = 256;       // two threads per query row: keys / head-dim columns [0, 32) and [32, 64)
constexpr int kKVBytes = kBK * 128;        // one [64][64] bf16 tile
constexpr int kStageBytes = 4 * kKVBytes;  // K hi, K lo, V hi, V lo
constexpr int kHeadsPerCta = 4;           // heads walked by one CTA (launch, TMEM allocation, pipeline fill paid once)
constexpr int kTileBytes = kStages * kStageBytes;
constexpr int kCtrlBytes = 256 + 2 * 2 * kBQ * 4;      // barriers + TMEM holder, then the row max / sum exchange
constexpr int kSmemBytes = kTileBytes + kCtrlBytes + 1024;   // > half an SM's shared memory: one CTA per SM (TMEM is whole)
constexpr int kAlignSlack = 1024;
static_assert(kSmemBytes > 114 * 1024 && kSmemBytes <= 227 * 1024, "shared memory budget");
// Tensor-memory columns: S0 [0,64) S1 [64,128) O [128,192) | P0 [192,256) P1 [256,320) | Q0 [320,384) Q1 [384,448);
// a P / Q buffer is 32 columns of bf16 pairs (64 values) for the hi halves followed by 32 for the lo halves.
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kColO = 128, kColP = 192, kColQ = 320;

// kind::f16 instruction descriptor, bf16 A (K-major) x bf16 B, fp32 D; b_mn = 1: B is MN-major.
__host__ __device__ constexpr uint32_t idesc(uint32_t m, uint32_t n, uint32_t b_mn) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (b_mn << 16) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ uint32_t pack_lo(uint32_t hi, float a, float b) {
  return pack2(a - __uint_as_float(hi << 16), b - __uint_as_float(hi & 0xffff0000u));
}

struct Params {
  CUtensorMap tm_kv;   // [rows, 2 D] bf16, box 64 cols x 64 rows; k and v share the geometry, bases differ -> two maps
  CUtensorMap tm_v;
  const __nv_bfloat16* q;
  __nv_bfloat16* out;
  int ld, lo_off, rows;
  const int2* tiles;   // (utterance, q0) per blockIdx.x, q0 a multiple of 128
  const int* utt_P;
  const int* utt_L;
};

#ifdef FY_STALLPROF
// Debug build only (tools/attn_stalls.py): cycles each role spends blocked on each barrier, summed over CTAs.
__device__ unsigned long long g_attn_stall[16];
#define FY_AW(slot, ...) { const long long aw_a = clock64(); __VA_ARGS__; aw[slot] += clock64() - aw_a; }
#define FY_AP(...) __VA_ARGS__
#else
#define FY_AW(slot, ...) { __VA_ARGS__; }
#define FY_AP(...)
#endif

__global__ void __launch_bounds__(kThreads, 1) attention_tc_kernel(const __grid_constant__ Params p) {
  extern __shared__ uint8_t smem_raw[];
  const uint32_t align_pad = (1024u - (ptx::smem_u32(smem_raw) & 1023u)) & 1023u;
  uint8_t* sKV = smem_raw + align_pad;               // stage s: K hi, K lo, V hi, V lo
  uint64_t* bars = reinterpret_cast<uint64_t*>(sKV + kStages * kStageBytes);
  uint64_t* k_full = bars;                    // [kStages]   K and V of a stage are tracked apart: K frees up as soon
  uint64_t* k_empty = bars + kStages;         // [kStages]   as S_t has been computed, a whole tile before P_t V_t, so
  uint64_t* v_full = bars + 2 * kStages;      // [kStages]   its slot is refilled that much earlier
  uint64_t* v_empty = bars + 3 * kStages;     // [kStages]
  uint64_t* s_full = bars + 4 * kStages;      // [2]
  uint64_t* s_empty = s_full + 2;             // [2]
  uint64_t* p_full = s_full + 4;              // [2]
  uint64_t* q_full = s_full + 6;              // [2]
  uint64_t* o_full = s_full + 8;              // 1
  uint64_t* o_empty = s_full + 9;             // 1
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(s_full + 10);
  static_assert((4 * kStages + 11) * 8 <= 256, "barrier block");
  float* xch = reinterpret_cast<float*>(bars + 32);   // [2 (tile parity)][2 (half)][128 rows]: row max / row sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int2 tl = p.tiles[blockIdx.x];
  const int u = tl.x, q0 = tl.y;
  const int h0 = blockIdx.y * kHeadsPerCta;        // this CTA walks heads h0 .. h0 + kHeadsPerCta - 1
  const int P = p.utt_P[u], L = p.utt_L[u];
  const int nkt = (L + kBK - 1) / kBK;
  const int total = nkt * kHeadsPerCta;            // global tile index T = g * nkt + t: every ring / parity is keyed on T

  if (threadIdx.x == kSoftmaxThreads) {
    ptx::prefetch_tmap(&p.tm_kv);
    ptx::prefetch_tmap(&p.tm_v);
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(&k_full[i], 1);
      ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1);
      ptx::mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&s_full[i], 1);
      ptx::mbar_init(&s_empty[i], 8);
      ptx::mbar_init(&p_full[i], 8);
      ptx::mbar_init(&q_full[i], 8);
    }
    ptx::mbar_init(o_full, 1);
    ptx::mbar_init(o_empty, 8);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc<kTmemCols>(tmem_holder);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_holder;

  if (warp == 9) {
    if (lane == 0) {
      // ------------------------------------------------ TMA producer: K / V tiles (ring)
      FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
      for (int T = 0; T < total; ++T) {
        const int g = T / nkt, t = T - g * nkt;
        const int col = (h0 + g) * kDH;
        const int st = T % kStages;
        const uint32_t ph = ((T / kStages) - 1) & 1;
        uint8_t* d = sKV + st * kStageBytes;
        if (T >= kStages) FY_AW(0, ptx::mbar_wait(&k_empty[st], ph))
        ptx::mbar_expect_tx(&k_full[st], 2 * kKVBytes);
        ptx::tma_load_2d(d, &p.tm_kv, &k_full[st], col, P + t * kBK);
        ptx::tma_load_2d(d + kKVBytes, &p.tm_kv, &k_full[st], col + p.lo_off, P + t * kBK);
        if (T >= kStages) FY_AW(1, ptx::mbar_wait(&v_empty[st], ph))
        ptx::mbar_expect_tx(&v_full[st], 2 * kKVBytes);
        ptx::tma_load_2d(d + 2 * kKVBytes, &p.tm_v, &v_full[st], col, P + t * kBK);
        ptx::tma_load_2d(d + 3 * kKVBytes, &p.tm_v, &v_full[st], col + p.lo_off, P + t * kBK);
      }
      FY_AP(for (int i = 0; i < 3; ++i) atomicAdd(&g_attn_stall[i], (unsigned long long)aw[i]);
            atomicAdd(&g_attn_stall[3], (unsigned long long)(clock64() - aw_t0));)
    }
  } else if (warp == 8) {
    {
      // ------------------------------------------------ every MMA.  The whole warp runs this loop (waits included) and
      // one elected lane issues each instruction: see ptx::mma_bf16_ts_elect.
      FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
      constexpr uint32_t id_s = idesc(kBQ, kBK, 0);   // S = Q K^T: both operands K-major
      constexpr uint32_t id_o = idesc(kBQ, kDH, 1);   // O = P V  : V is MN-major ([key][d] rows of 128 bytes)
      auto issue_s = [&](int T) {
        const int g = T / nkt;
        if (T == g * nkt) {                            // first tile of a head: its Q has been stored to TMEM
          FY_AW(9, ptx::mbar_wait(&q_full[g & 1], (g >> 1) & 1))
          ptx::tc_fence_after();
        }
        const uint32_t qa = tmem + kColQ + (g & 1) * 64;
        const uint32_t kb = ptx::smem_u32(sKV + (T % kStages) * kStageBytes);
        const uint32_t d = tmem + (T & 1) * kBK;
        uint32_t acc = 0;
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) {
          const uint32_t ta = qa + (pr == 2 ? 32 : 0);
          const uint64_t db = ptx::make_desc_sw128(kb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) {   // 16 head-dim values per step: 8 TMEM columns of Q, +32 bytes in K rows
            ptx::mma_bf16_ts_elect(d, ta + 8 * k, db + 2 * k, id_s, acc);
            acc = 1;
          }
        }
        ptx::mma_commit_elect(&s_full[T & 1]);
        ptx::mma_commit_elect(&k_empty[T % kStages]);
      };
      auto issue_pv = [&](int T) {
        const uint32_t vb = ptx::smem_u32(sKV + (T % kStages) * kStageBytes) + 2 * kKVBytes;
        const uint32_t pa = tmem + kColP + (T & 1) * 64;
        const uint32_t d = tmem + kColO;
        uint32_t acc = 0;
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) {
          const uint32_t ta = pa + (pr == 2 ? 32 : 0);
          const uint64_t db = ptx::make_desc_sw128(vb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {   // 16 keys per step: 8 TMEM columns of P, +16 rows (2048 B) in V
            ptx::mma_bf16_ts_elect(d, ta + 8 * k, db + 128 * k, id_o, acc);
            acc = 1;
          }
        }
        ptx::mma_commit_elect(o_full);
        ptx::mma_commit_elect(&v_empty[T % kStages]);
      };
      FY_AW(4, ptx::mbar_wait(&k_full[0], 0))
      ptx::tc_fence_after();
      issue_s(0);
      for (int T = 0; T < total; ++T) {
        if (T + 1 < total) {
          FY_AW(4, ptx::mbar_wait(&k_full[(T + 1) % kStages], ((T + 1) / kStages) & 1))
          if (T + 1 >= 2) FY_AW(5, ptx::mbar_wait(&s_empty[(T + 1) & 1], (((T + 1) >> 1) - 1) & 1))
          ptx::tc_fence_after();
          issue_s(T + 1);
        }
        FY_AW(6, ptx::mbar_wait(&v_full[T % kStages], (T / kStages) & 1))
        FY_AW(7, ptx::mbar_wait(&p_full[T & 1], (T >> 1) & 1))
        if (T >= 1) FY_AW(8, ptx::mbar_wait(o_empty, (T - 1) & 1))
        ptx::tc_fence_after();
        issue_pv(T);
      }
      FY_AP(if (lane == 0) {
        for (int i = 4; i < 10; ++i) atomicAdd(&g_attn_stall[i], (unsigned long long)aw[i]);
        atomicAdd(&g_attn_stall[10], (unsigned long long)(clock64() - aw_t0));
        atomicAdd(&g_attn_stall[15], 1ull);
      })
    }
  } else {
    // ------------------------------------------------ softmax: two threads per query row (warps w and w + 4)
    const int wq = warp & 3, half = warp >> 2;
    const int row = wq * 32 + lane;
    const uint32_t t_row = tmem + (static_cast<uint32_t>(wq * 32) << 16);
    const float kLog2e = 1.4426950408889634f;
    constexpr int HC = kBK / 2;     // 32 keys (and 32 head-dim columns) per thread
    FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
    float o[HC];
    // Q staging: this thread owns head-dim values [32 half, 32 half + 32) of its query row, hi and lo halves
    const bool q_row_ok = P + q0 + row < p.rows;
    const __nv_bfloat16* q_src = p.q + (size_t)(q_row_ok ? P + q0 + row : 0) * p.ld + h0 * kDH + half * HC;
    uint32_t qh[16], ql[16];
    auto load_q = [&](int g) {
      const uint4* a = reinterpret_cast<const uint4*>(q_src + g * kDH);
      const uint4* b = reinterpret_cast<const uint4*>(q_src + g * kDH + p.lo_off);
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const uint4 x = q_row_ok ? __ldg(a + c) : make_uint4(0u, 0u, 0u, 0u);
        const uint4 y = q_row_ok ? __ldg(b + c) : make_uint4(0u, 0u, 0u, 0u);
        qh[4 * c] = x.x; qh[4 * c + 1] = x.y; qh[4 * c + 2] = x.z; qh[4 * c + 3] = x.w;
        ql[4 * c] = y.x; ql[4 * c + 1] = y.y; ql[4 * c + 2] = y.z; ql[4 * c + 3] = y.w;
      }
    };
    auto store_q = [&](int g) {   // followed by tmem_st_wait + arrive by the caller
      const uint32_t qa = t_row + kColQ + (g & 1) * 64 + half * 16;
      ptx::tmem_st_32x16(qa, qh);
      ptx::tmem_st_32x16(qa + 32, ql);
    };
    load_q(0);
    store_q(0);
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&q_full[0]);
    if (kHeadsPerCta > 1) load_q(1);   // in flight until the first tile's P has been written
    auto fold_o = [&](int T) {   // o += O_T (o is already expressed relative to the maximum O_T was computed with)
      FY_AW(12, ptx::mbar_wait(o_full, T & 1))
      ptx::tc_fence_after();
      uint32_t ov[32];
      ptx::tmem_ld_32x32(t_row + kColO + half * HC, ov);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < HC; ++j) o[j] += __uint_as_float(ov[j]);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(o_empty);
    };
    // Head epilogue: add the two halves' row sums (exchanged through shared memory between warps w and w + 4 only:
    // named barrier 1 + wq, 64 threads), normalise, split to bf16 hi | lo and store.
    auto finish_head = [&](int g, float l) {
      const int col = (h0 + g) * kDH + half * HC;
      float* xl = xch + (g & 1) * 2 * kBQ;   // slot g & 1: its next use is two pair barriers away
      xl[half * kBQ + row] = l;
      FY_AW(13, asm volatile("bar.sync %0, 64;\n" ::"r"(1 + wq) : "memory"))
      l += xl[(half ^ 1) * kBQ + row];
      if (q0 + row < L) {
        const float inv = 1.0f / l;
        __nv_bfloat16* dst = p.out + (size_t)(P + q0 + row) * p.ld + col;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          float a[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) a[j] = o[8 * c + j] * inv;
          uint4 hi, lo;
          hi.x = pack2(a[0], a[1]); hi.y = pack2(a[2], a[3]); hi.z = pack2(a[4], a[5]); hi.w = pack2(a[6], a[7]);
          lo.x = pack_lo(hi.x, a[0], a[1]); lo.y = pack_lo(hi.y, a[2], a[3]);
          lo.z = pack_lo(hi.z, a[4], a[5]); lo.w = pack_lo(hi.w, a[6], a[7]);
          *reinterpret_cast<uint4*>(dst + 8 * c) = hi;
          *reinterpret_cast<uint4*>(dst + p.lo_off + 8 * c) = lo;
        }
      }
      // The guard row after the utterance is written as zeros: unlike the cp.async kernel this one sees the rows
      // behind an utterance inside its last K/V tile (times P = 0), so they must stay finite through all layers, and
      // a guard row of the attention output left holding stale workspace bytes would turn into NaN in the next
      // projection.
      if (q0 + row == L || (row == kBQ - 1 && q0 + kBQ == L)) {
        __nv_bfloat16* dst = p.out + (size_t)(P + L) * p.ld + col;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          *reinterpret_cast<uint4*>(dst + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
          *reinterpret_cast<uint4*>(dst + p.lo_off + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
    };
    // One flat loop over the CTA's tiles T = g * nkt + t.  O_{T-1} is folded one tile late (after P_T has been handed
    // to the tensor core), ALSO across a head boundary: the previous head's last fold, normalisation and store run
    // under the first S / P V products of the next head instead of leaving the tensor pipe idle.
#pragma unroll
    for (int j = 0; j < HC; ++j) o[j] = 0.f;
    float m = -INFINITY, l = 0.f, l_prev = 0.f;
    int g = 0, t = 0;
#pragma unroll 1
    for (int T = 0; T < total; ++T) {
      if (t == 0) { l_prev = l; l = 0.f; m = -INFINITY; }
      FY_AW(11, ptx::mbar_wait(&s_full[T & 1], (T >> 1) & 1))
      ptx::tc_fence_after();
      // the whole 64-key row is read: the other half only for the row maximum (no exchange, no barrier in the loop)
      uint32_t sv[32], sw[32];
      ptx::tmem_ld_32x32(t_row + (T & 1) * kBK + (half ^ 1) * HC, sw);
      ptx::tmem_ld_32x32(t_row + (T & 1) * kBK + half * HC, sv);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&s_empty[T & 1]);
      float s[HC];
#pragma unroll
      for (int j = 0; j < HC; ++j) s[j] = __uint_as_float(sv[j]);
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      if ((t + 1) * kBK > L) {
#pragma unroll
        for (int j = 0; j < HC; ++j) {
          if (t * kBK + half * HC + j >= L) s[j] = -INFINITY;
          if (t * kBK + (half ^ 1) * HC + j < L) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(sw[j]));
        }
      } else {
#pragma unroll
        for (int j = 0; j < HC; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], __uint_as_float(sw[j]));
      }
#pragma unroll
      for (int j = 0; j < HC; ++j) mx4[j & 3] = fmaxf(mx4[j & 3], s[j]);
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      const float m_new = fmaxf(m, mx * kLog2e);     // finite: every tile holds at least one valid key
      const float corr = ex2(m - m_new);             // 0 on the first tile of a head (m = -inf)
      m = m_new;
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < HC; ++j) {
        s[j] = ex2(fmaf(s[j], kLog2e, -m_new));
        sum4[j & 3] += s[j];
      }
      l = l * corr + ((sum4[0] + sum4[1]) + (sum4[2] + sum4[3]));   // partial (this half's keys)
      // P row -> tensor memory: keys [32 half, 32 half + 32) are columns [16 half, 16 half + 16) of the hi and of the
      // lo block of buffer T & 1 (two bf16 per column, even key in the low half-word)
      {
        uint32_t ph[16], pl[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          ph[j] = pack2(s[2 * j], s[2 * j + 1]);
          pl[j] = pack_lo(ph[j], s[2 * j], s[2 * j + 1]);
        }
        const uint32_t pa = t_row + kColP + (T & 1) * 64 + half * 16;
        ptx::tmem_st_32x16(pa, ph);
        ptx::tmem_st_32x16(pa + 32, pl);
      }
      // Q of the next head goes into the buffer head g - 1 used: s_full of this tile implies every earlier product
      // (all S of head g - 1) has completed.
      const bool stage_q = t == 0 && g + 1 < kHeadsPerCta;
      if (stage_q) store_q(g + 1);
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(&p_full[T & 1]);
        if (stage_q) ptx::mbar_arrive(&q_full[(g + 1) & 1]);
      }
      if (stage_q && g + 2 < kHeadsPerCta) load_q(g + 2);
      // Fold O_{T-1} (long finished: its MMA ran under this tile's softmax) into the running accumulator, one tile
      // late so that the tensor-core latency of P V is never waited for; then rescale for this tile's maximum.
      if (T > 0) fold_o(T - 1);
      if (t == 0) {
        if (T > 0) finish_head(g - 1, l_prev);
#pragma unroll
        for (int j = 0; j < HC; ++j) o[j] = 0.f;
      } else {
#pragma unroll
        for (int j = 0; j < HC; ++j) o[j] *= corr;
      }
      if (++t == nkt) { t = 0; ++g; }
    }
    fold_o(total - 1);
    finish_head(kHeadsPerCta - 1, l);
    FY_AP(if (threadIdx.x == 0) {
      for (int i = 11; i < 14; ++i) atomicAdd(&g_attn_stall[i], (unsigned long long)aw[i]);
      atomicAdd(&g_attn_stall[14], (unsigned long long)(clock64() - aw_t0));
    })
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 0) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc<kTmemCols>(tmem);
  }
}

}  // namespace attn_tc
}  // namespace fy
