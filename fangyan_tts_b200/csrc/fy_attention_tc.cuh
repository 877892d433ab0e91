// Variable-length self-attention on tcgen05 (S3 encoder, row A9), split-bf16 mode.  FY_ATTN_TC=1 selects it.
//
// One CTA = 128 query rows of one (utterance, head).  Both GEMMs of the flash loop run on the 5th-gen tensor cores
// with fp32 accumulators in TMEM; the CUDA cores only do the softmax:
//   warp 5, one thread : TMA loads (Q of each head into two alternating buffers; K/V 64-key tiles into a 3-stage ring)
//   warp 4, one thread : every tcgen05.mma
//                        S_t = Q K_t^T   (M 128, N 64, K 64; three products hi*hi + hi*lo + lo*hi)
//                        O_t = P_t V_t   (M 128, N 64, K 64 keys; V is consumed MN-major, i.e. exactly as the
//                                         [key][d] tile TMA delivers it -- no transpose anywhere)
//   warps 0-7, 2 threads per query row (32 keys each): tcgen05.ld of the S row, running max / sum in registers (the
//                        two halves meet through shared memory once per tile, no shuffles), P = 2^(s log2e - m)
//                        written to shared memory as bf16 hi | lo in the K-major 128-byte-swizzled layout the second
//                        MMA reads, then O_t is read back from TMEM and folded into the register accumulator.
// S is double-buffered in TMEM and P in shared memory, so S_{t+1} is computed while the softmax of tile t runs.
// Rows past the utterance inside a tile belong to the next utterance or a guard row: keys are masked, queries are
// computed and not stored, V rows meet P = 0; all of it is finite data of the same buffers (the kernel keeps the
// guard rows of its own output at zero for that reason).
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace attn_tc {

constexpr int kBQ = 128, kBK = 64, kDH = 64, kThreads = 320, kStages = 3;
constexpr int kSoftmaxThreads = 256;       // two threads per query row: keys / head-dim columns [0, 32) and [32, 64)
constexpr int kQBytes = kBQ * 128;         // one [128][64] bf16 tile
constexpr int kKVBytes = kBK * 128;        // one [64][64] bf16 tile
constexpr int kStageBytes = 4 * kKVBytes;  // K hi, K lo, V hi, V lo
constexpr int kPBytes = kBQ * 128;         // one [128][64] bf16 probability tile
constexpr int kHeadsPerCta = 4;           // heads walked by one CTA (launch, TMEM allocation, pipeline fill paid once)
constexpr int kTileBytes = 4 * kQBytes + kStages * kStageBytes + 4 * kPBytes;
constexpr int kCtrlBytes = 256 + 2 * 2 * kBQ * 4;      // barriers + TMEM holder, then the row max / sum exchange
constexpr int kSmemBytes = 227 * 1024;                  // everything the SM has; what is left over is alignment slack
constexpr int kAlignSlack = kSmemBytes - kTileBytes - kCtrlBytes;
static_assert(kStages == 3, "barrier layout");
static_assert(kAlignSlack >= 512, "shared memory budget");
constexpr uint32_t kTmemCols = 256;        // S0 [0,64) S1 [64,128) O [128,192)

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
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory"); }

struct Params {
  CUtensorMap tm_q;    // [rows, 2 D] bf16, box 64 cols x 128 rows
  CUtensorMap tm_kv;   // same tensor geometry (k and v share ld), box 64 cols x 64 rows; bases differ -> two maps
  CUtensorMap tm_v;
  __nv_bfloat16* out;
  int ld, lo_off;
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
  if (align_pad > static_cast<uint32_t>(kAlignSlack)) __trap();   // dynamic shared memory starts 1024-aligned in practice
  uint8_t* smem = smem_raw + align_pad;
  uint8_t* sQ = smem;                              // buffer g & 1 of head g: hi | lo
  uint8_t* sKV = sQ + 4 * kQBytes;                 // stage s: K hi, K lo, V hi, V lo
  uint8_t* sP = sKV + kStages * kStageBytes;       // buffer b: P hi, P lo
  uint64_t* bars = reinterpret_cast<uint64_t*>(sP + 4 * kPBytes);
  uint64_t* q_full = bars;            // [2]
  uint64_t* q_empty = bars + 2;       // [2]
  uint64_t* k_full = bars + 4;        // [kStages]   K and V of a stage are tracked apart: K frees up as soon as S_t
  uint64_t* k_empty = bars + 7;       // [kStages]   has been computed, a whole tile before P_t V_t, so its slot is
  uint64_t* v_full = bars + 10;       // [kStages]   refilled that much earlier
  uint64_t* v_empty = bars + 13;      // [kStages]
  uint64_t* s_full = bars + 16;       // [2]
  uint64_t* s_empty = bars + 18;      // [2]
  uint64_t* p_full = bars + 20;       // [2]
  uint64_t* o_full = bars + 22;       // 1
  uint64_t* o_empty = bars + 23;      // 1
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(bars + 24);
  float* xch = reinterpret_cast<float*>(bars + 32);   // [2 (tile parity)][2 (half)][128 rows]: row max / row sum exchange

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int2 tl = p.tiles[blockIdx.x];
  const int u = tl.x, q0 = tl.y;
  const int h0 = blockIdx.y * kHeadsPerCta;        // this CTA walks heads h0 .. h0 + kHeadsPerCta - 1
  const int P = p.utt_P[u], L = p.utt_L[u];
  const int nkt = (L + kBK - 1) / kBK;
  const int total = nkt * kHeadsPerCta;            // global tile index T = g * nkt + t: every ring / parity is keyed on T

  if (threadIdx.x == kSoftmaxThreads) {
    ptx::prefetch_tmap(&p.tm_q);
    ptx::prefetch_tmap(&p.tm_kv);
    ptx::prefetch_tmap(&p.tm_v);
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(&k_full[i], 1);
      ptx::mbar_init(&k_empty[i], 1);
      ptx::mbar_init(&v_full[i], 1);
      ptx::mbar_init(&v_empty[i], 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&q_full[i], 1);
      ptx::mbar_init(&q_empty[i], 1);
      ptx::mbar_init(&s_full[i], 1);
      ptx::mbar_init(&s_empty[i], 8);
      ptx::mbar_init(&p_full[i], 8);
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
      // ------------------------------------------------ TMA producer: Q per head (2 buffers), K/V tiles (ring)
      FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
      for (int T = 0; T < total; ++T) {
        const int g = T / nkt, t = T - g * nkt;
        const int col = (h0 + g) * kDH;
        if (t == 0) {
          if (g >= 2) FY_AW(2, ptx::mbar_wait(&q_empty[g & 1], ((g >> 1) - 1) & 1))   // every S of head g - 2 has completed
          uint8_t* d = sQ + (g & 1) * 2 * kQBytes;
          ptx::mbar_expect_tx(&q_full[g & 1], 2 * kQBytes);
          ptx::tma_load_2d(d, &p.tm_q, &q_full[g & 1], col, P + q0);
          ptx::tma_load_2d(d + kQBytes, &p.tm_q, &q_full[g & 1], col + p.lo_off, P + q0);
        }
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
    if (lane == 0) {
      // ------------------------------------------------ every MMA
      FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
      constexpr uint32_t id_s = idesc(kBQ, kBK, 0);   // S = Q K^T: both operands K-major
      constexpr uint32_t id_o = idesc(kBQ, kDH, 1);   // O = P V  : V is MN-major ([key][d] rows of 128 bytes)
      auto issue_s = [&](int T) {
        const int g = T / nkt;
        if (T == g * nkt) FY_AW(9, ptx::mbar_wait(&q_full[g & 1], (g >> 1) & 1))   // first tile of a head: its Q has landed
        const uint32_t q_hi = ptx::smem_u32(sQ + (g & 1) * 2 * kQBytes), q_lo = q_hi + kQBytes;
        const uint32_t kb = ptx::smem_u32(sKV + (T % kStages) * kStageBytes);
        const uint32_t d = tmem + (T & 1) * kBK;
        uint32_t acc = 0;
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) {
          const uint64_t da = ptx::make_desc_sw128(pr == 2 ? q_lo : q_hi);
          const uint64_t db = ptx::make_desc_sw128(kb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
          for (int k = 0; k < kDH / 16; ++k) {
            ptx::mma_bf16_ss(d, da + 2 * k, db + 2 * k, id_s, acc);
            acc = 1;
          }
        }
        ptx::mma_commit(&s_full[T & 1]);
        ptx::mma_commit(&k_empty[T % kStages]);
        if (T == g * nkt + nkt - 1) ptx::mma_commit(&q_empty[g & 1]);   // last S of the head: its Q buffer frees up
      };
      auto issue_pv = [&](int T) {
        const uint32_t vb = ptx::smem_u32(sKV + (T % kStages) * kStageBytes) + 2 * kKVBytes;
        const uint32_t pb = ptx::smem_u32(sP + (T & 1) * 2 * kPBytes);
        const uint32_t d = tmem + 2 * kBK;
        uint32_t acc = 0;
#pragma unroll
        for (int pr = 0; pr < 3; ++pr) {
          const uint64_t da = ptx::make_desc_sw128(pb + (pr == 2 ? kPBytes : 0));
          const uint64_t db = ptx::make_desc_sw128(vb + (pr == 1 ? kKVBytes : 0));
#pragma unroll
          for (int k = 0; k < kBK / 16; ++k) {   // 16 keys per step: +32 bytes in P rows, +16 rows (2048 B) in V
            ptx::mma_bf16_ss(d, da + 2 * k, db + 128 * k, id_o, acc);
            acc = 1;
          }
        }
        ptx::mma_commit(o_full);
        ptx::mma_commit(&v_empty[T % kStages]);
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
      FY_AP(for (int i = 4; i < 10; ++i) atomicAdd(&g_attn_stall[i], (unsigned long long)aw[i]);
            atomicAdd(&g_attn_stall[10], (unsigned long long)(clock64() - aw_t0));
            atomicAdd(&g_attn_stall[15], 1ull);)
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
    auto fold_o = [&](int T) {   // o += O_T (o is already expressed relative to the maximum O_T was computed with)
      FY_AW(12, ptx::mbar_wait(o_full, T & 1))
      ptx::tc_fence_after();
      uint32_t ov[32];
      ptx::tmem_ld_32x32(t_row + 2 * kBK + half * HC, ov);
      ptx::tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < HC; ++j) o[j] += __uint_as_float(ov[j]);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(o_empty);
    };
#pragma unroll 1
    for (int g = 0; g < kHeadsPerCta; ++g) {
      const int col = (h0 + g) * kDH + half * HC;
#pragma unroll
      for (int j = 0; j < HC; ++j) o[j] = 0.f;
      float m = -INFINITY, l = 0.f;
#pragma unroll 1
      for (int t = 0; t < nkt; ++t) {
        const int T = g * nkt + t;
        FY_AW(11, ptx::mbar_wait(&s_full[T & 1], (T >> 1) & 1))
        ptx::tc_fence_after();
        uint32_t sv[32];
        ptx::tmem_ld_32x32(t_row + (T & 1) * kBK + half * HC, sv);
        ptx::tmem_ld_wait();
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&s_empty[T & 1]);
        float s[HC];
#pragma unroll
        for (int j = 0; j < HC; ++j) s[j] = __uint_as_float(sv[j]);
        if ((t + 1) * kBK > L) {
#pragma unroll
          for (int j = 0; j < HC; ++j)
            if (t * kBK + half * HC + j >= L) s[j] = -INFINITY;
        }
        float mx = s[0];
#pragma unroll
        for (int j = 1; j < HC; ++j) mx = fmaxf(mx, s[j]);
        // row maximum over both halves: exchanged through shared memory (slot T & 1, one named barrier per tile)
        float* xm = xch + (T & 1) * 2 * kBQ;
        xm[half * kBQ + row] = mx;
        FY_AW(13, asm volatile("bar.sync 1, %0;\n" ::"n"(kSoftmaxThreads) : "memory"))
        mx = fmaxf(mx, xm[(half ^ 1) * kBQ + row]);
        const float m_new = fmaxf(m, mx * kLog2e);     // finite: every tile holds at least one valid key
        const float corr = ex2(m - m_new);
        m = m_new;
        float sum = 0.f;
#pragma unroll
        for (int j = 0; j < HC; ++j) {
          s[j] = ex2(fmaf(s[j], kLog2e, -m_new));
          sum += s[j];
        }
        l = l * corr + sum;                            // partial (this half's keys); the halves are added at the end
        // P row -> shared memory, K-major SW128: chunk c (8 keys) of row r at r * 128 + ((c ^ (r & 7)) << 4)
        uint8_t* ph = sP + (T & 1) * 2 * kPBytes + row * 128;
#pragma unroll
        for (int cc = 0; cc < 4; ++cc) {
          const int c = half * 4 + cc;
          uint4 hi, lo;
          hi.x = pack2(s[8 * cc], s[8 * cc + 1]); hi.y = pack2(s[8 * cc + 2], s[8 * cc + 3]);
          hi.z = pack2(s[8 * cc + 4], s[8 * cc + 5]); hi.w = pack2(s[8 * cc + 6], s[8 * cc + 7]);
          lo.x = pack_lo(hi.x, s[8 * cc], s[8 * cc + 1]); lo.y = pack_lo(hi.y, s[8 * cc + 2], s[8 * cc + 3]);
          lo.z = pack_lo(hi.z, s[8 * cc + 4], s[8 * cc + 5]); lo.w = pack_lo(hi.w, s[8 * cc + 6], s[8 * cc + 7]);
          const int off = (c ^ (row & 7)) << 4;
          *reinterpret_cast<uint4*>(ph + off) = hi;
          *reinterpret_cast<uint4*>(ph + kPBytes + off) = lo;
        }
        fence_async_smem();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(&p_full[T & 1]);
        // Fold O_{t-1} (long finished: its MMA ran under this tile's softmax) into the running accumulator, one tile
        // late so that the tensor-core latency of P V is never waited for; then rescale for this tile's maximum.
        if (t > 0) fold_o(T - 1);
#pragma unroll
        for (int j = 0; j < HC; ++j) o[j] *= corr;
      }
      fold_o(g * nkt + nkt - 1);
      // row sum = both halves (slot g & 1 of the exchange buffer; the next use of that slot is two barriers away)
      float* xl = xch + (g & 1) * 2 * kBQ;
      xl[half * kBQ + row] = l;
      asm volatile("bar.sync 1, %0;\n" ::"n"(kSoftmaxThreads) : "memory");
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
    }
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
