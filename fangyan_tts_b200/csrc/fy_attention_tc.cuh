// Variable-length self-attention on tcgen05 (S3 encoder, row A9), split-bf16 mode.  FY_ATTN_TC=0 selects the
// mma.sync kernel of fy_attention.cuh instead.
//
// One CTA = 128 query rows of one utterance, walked over kHeadsPerCta heads.  Both GEMMs of the flash loop run on the
// 5th-gen tensor cores with fp32 accumulators in TMEM, and BOTH take their A operand from tensor memory (".ts" form):
// Q is constant over a head's key tiles and P is produced by the softmax threads, so neither ever touches shared
// memory -- the only operand bytes the shared-memory port serves are the K and V tiles (2 KB per MMA instead of 6 KB,
// which is what held the all-shared-memory version at ~100 clk per M128 N64 K16 instruction against a 32 clk floor).
//   warp 17, one thread: TMA loads of the K / V 64-key tiles (hi and lo halves) into a kStages ring
//   warp 16 (one elected lane): every tcgen05.mma
//                        S_t = Q K_t^T   (M 128, N 64, K 64; three products hi*hi + hi*lo + lo*hi)
//                        O_t = P_t V_t   (M 128, N 64, K 64 keys; V is consumed MN-major, i.e. exactly as the
//                                         [key][d] tile TMA delivers it -- no transpose anywhere)
//                        O accumulates in TMEM over a head's tiles (two buffers, alternating by head).
//   warps 0-15, FOUR threads per query row (16 keys each; the role-stall counters showed the softmax warps to be
//                        latency-bound at two warps per scheduler, and the four warps of a row quarter share one
//                        scheduler, so waiting for each other costs no issue slots): tcgen05.ld of the thread's 16
//                        scores, partial row maximum exchanged through shared memory (one 128-thread named barrier
//                        per row quarter), P = 2^(s log2e - m) packed to bf16 hi | lo and stored to TMEM with
//                        tcgen05.st.  The running maximum m only moves when the new one exceeds it by more than 2^8
//                        (P stays <= 256, which the hi | lo split carries exactly); only then is O rescaled in TMEM
//                        (tcgen05.ld / st of the thread's 16 columns) -- about once per head instead of a read-back
//                        every tile.  The same threads stage Q: each reads its 32 + 32 bytes of the next head's query
//                        row from global memory and stores them to TMEM under a first tile's softmax.
// S, P, Q and O are double-buffered in TMEM (all 512 columns): S_{t+1} is computed while the softmax of tile t runs,
// and a head is normalised and stored under the first products of the next head.
// Rows past the utterance inside a tile belong to the next utterance or a guard row: keys are masked, queries are
// computed and not stored, V rows meet P = 0; all of it is finite data of the same buffers (the kernel keeps the
// guard rows of its own output at zero for that reason).
#pragma once
#include "fy_common.cuh"
#include "fy_ptx.cuh"

namespace fy {
namespace attn_tc {

constexpr int kBQ = 128, kBK = 64, kDH = 64, kStages = 4;
constexpr int kParts = 4;                  // threads per query row: keys / head-dim columns [16 part, 16 part + 16)
constexpr int kSoftmaxWarps = 4 * kParts, kSoftmaxThreads = 32 * kSoftmaxWarps;
constexpr int kThreads = kSoftmaxThreads + 64;
constexpr int kKVBytes = kBK * 128;        // one [64][64] bf16 tile
constexpr int kStageBytes = 4 * kKVBytes;  // K hi, K lo, V hi, V lo
constexpr int kHeadsPerCta = 4;           // heads walked by one CTA (launch, TMEM allocation, pipeline fill paid once)
constexpr int kTileBytes = kStages * kStageBytes;
constexpr int kCtrlBytes = 256 + 2 * 2 * kParts * kBQ * 4;   // barriers + TMEM holder, then the row max and sum exchange
constexpr int kSmemBytes = kTileBytes + kCtrlBytes + 1024;   // > half an SM's shared memory: one CTA per SM (TMEM is whole)
constexpr int kAlignSlack = 1024;
static_assert(kSmemBytes > 114 * 1024 && kSmemBytes <= 227 * 1024, "shared memory budget");
// Tensor-memory columns: S0 [0,64) S1 [64,128) | O0 [128,192) O1 [192,256) | P0 [256,320) P1 [320,384) |
// Q0 [384,448) Q1 [448,512); a P / Q buffer is 32 columns of bf16 pairs (64 values) for the hi halves followed by 32
// for the lo halves.
constexpr uint32_t kTmemCols = 512;
constexpr uint32_t kColO = 128, kColP = 256, kColQ = 384;
constexpr float kRescaleSlack = 8.0f;      // log2 units the running maximum may lag behind before O is rescaled
                                           // (Params::slack; FY_ATTN_SLACK=0 forces the rescale path in tests)

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
  float slack;         // kRescaleSlack
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
  uint64_t* o_empty = s_full + 8;             // [2]  O buffer g & 1 has been read out (head g normalised)
  uint64_t* o_full = s_full + 10;             // 1    one phase per P V product group
  uint64_t* all_done = s_full + 11;           // 1    every product of the CTA has completed
  uint32_t* tmem_holder = reinterpret_cast<uint32_t*>(s_full + 12);
  static_assert((4 * kStages + 13) * 8 <= 256, "barrier block");
  float* xmax = reinterpret_cast<float*>(bars + 32);   // [2 (tile parity)][kParts][128 rows]: partial row maxima
  float* xsum = xmax + 2 * kParts * kBQ;               // [2 (head parity)][kParts][128 rows]: partial row sums

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int2 tl = p.tiles[blockIdx.x];
  const int u = tl.x, q0 = tl.y;
  const int h0 = blockIdx.y * kHeadsPerCta;        // this CTA walks heads h0 .. h0 + kHeadsPerCta - 1
  const int P = p.utt_P[u], L = p.utt_L[u];
  const int nkt = (L + kBK - 1) / kBK;
  const int total = nkt * kHeadsPerCta;            // global tile index T = g * nkt + t: every ring / parity is keyed on T

  // Softmax threads: four per query row (warps w, w+4, w+8, w+12 own row quarter w).  Q staging: a thread owns head-dim
  // values [16 part, 16 part + 16) of its query row, hi and lo halves; the first head's loads are issued here so that
  // they land under the barrier initialisation and the TMEM allocation.
  const int wq = warp & 3, part = warp >> 2;
  const int row = wq * 32 + lane;
  constexpr int PC = kBK / kParts;     // 16 keys (and 16 head-dim columns) per thread
  const bool q_row_ok = warp < kSoftmaxWarps && P + q0 + row < p.rows;
  const __nv_bfloat16* q_src = p.q + (size_t)(q_row_ok ? P + q0 + row : 0) * p.ld + h0 * kDH + (part & 3) * PC;
  uint32_t qh[8], ql[8];
  auto load_q = [&](int g) {
    const uint4* a = reinterpret_cast<const uint4*>(q_src + g * kDH);
    const uint4* b = reinterpret_cast<const uint4*>(q_src + g * kDH + p.lo_off);
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      const uint4 x = q_row_ok ? __ldg(a + c) : make_uint4(0u, 0u, 0u, 0u);
      const uint4 y = q_row_ok ? __ldg(b + c) : make_uint4(0u, 0u, 0u, 0u);
      qh[4 * c] = x.x; qh[4 * c + 1] = x.y; qh[4 * c + 2] = x.z; qh[4 * c + 3] = x.w;
      ql[4 * c] = y.x; ql[4 * c + 1] = y.y; ql[4 * c + 2] = y.z; ql[4 * c + 3] = y.w;
    }
  };
  if (warp < kSoftmaxWarps) load_q(0);

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
      ptx::mbar_init(&s_empty[i], kSoftmaxWarps);
      ptx::mbar_init(&p_full[i], kSoftmaxWarps);
      ptx::mbar_init(&q_full[i], kSoftmaxWarps);
    }
    ptx::mbar_init(&o_empty[0], kSoftmaxWarps);
    ptx::mbar_init(&o_empty[1], kSoftmaxWarps);
    ptx::mbar_init(o_full, 1);
    ptx::mbar_init(all_done, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 0) ptx::tmem_alloc<kTmemCols>(tmem_holder);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *tmem_holder;

  if (warp == kSoftmaxWarps + 1) {
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
  } else if (warp == kSoftmaxWarps) {
    {
      // ------------------------------------------------ every MMA.  The whole warp runs this loop (waits included) and
      // one elected lane issues each instruction: see ptx::mma_bf16_ts_elect.
      FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
      constexpr uint32_t id_s = idesc(kBQ, kBK, 0);   // S = Q K^T: both operands K-major
      constexpr uint32_t id_o = idesc(kBQ, kDH, 1);   // O = P V  : V is MN-major ([key][d] rows of 128 bytes)
      auto issue_s = [&](int T, int g, int t) {
        if (t == 0) {                                  // first tile of a head: its Q has been stored to TMEM
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
      auto issue_pv = [&](int T, int g, int t) {
        const uint32_t vb = ptx::smem_u32(sKV + (T % kStages) * kStageBytes) + 2 * kKVBytes;
        const uint32_t pa = tmem + kColP + (T & 1) * 64;
        const uint32_t d = tmem + kColO + (g & 1) * 64;
        uint32_t acc = t > 0;                          // a head's first tile overwrites its O buffer
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
      issue_s(0, 0, 0);
      int g = 0, t = 0;        // of tile T
      int g1 = 0, t1 = 0;      // of tile T + 1
      for (int T = 0; T < total; ++T) {
        if (++t1 == nkt) { t1 = 0; ++g1; }
        if (T + 1 < total) {
          FY_AW(4, ptx::mbar_wait(&k_full[(T + 1) % kStages], ((T + 1) / kStages) & 1))
          if (T + 1 >= 2) FY_AW(5, ptx::mbar_wait(&s_empty[(T + 1) & 1], (((T + 1) >> 1) - 1) & 1))
          ptx::tc_fence_after();
          issue_s(T + 1, g1, t1);
        }
        FY_AW(6, ptx::mbar_wait(&v_full[T % kStages], (T / kStages) & 1))
        FY_AW(7, ptx::mbar_wait(&p_full[T & 1], (T >> 1) & 1))
        if (t == 0 && g >= 2) FY_AW(8, ptx::mbar_wait(&o_empty[g & 1], ((g >> 1) - 1) & 1))   // head g - 2 read out
        ptx::tc_fence_after();
        issue_pv(T, g, t);
        g = g1; t = t1;
      }
      ptx::mma_commit_elect(all_done);
      FY_AP(if (lane == 0) {
        for (int i = 4; i < 10; ++i) atomicAdd(&g_attn_stall[i], (unsigned long long)aw[i]);
        atomicAdd(&g_attn_stall[10], (unsigned long long)(clock64() - aw_t0));
        atomicAdd(&g_attn_stall[15], 1ull);
      })
    }
  } else {
    // ------------------------------------------------ softmax: four threads per query row (warps w, w+4, w+8, w+12)
    const uint32_t t_row = tmem + (static_cast<uint32_t>(wq * 32) << 16);
    const float kLog2e = 1.4426950408889634f;
    FY_AP(long long aw[16] = {0}; const long long aw_t0 = clock64();)
    auto store_q = [&](int g) {   // followed by tmem_st_wait + arrive by the caller
      const uint32_t qa = t_row + kColQ + (g & 1) * 64 + part * 8;
      ptx::tmem_st_32x8(qa, qh);
      ptx::tmem_st_32x8(qa + 32, ql);
    };
    store_q(0);                            // loaded before the prologue barrier
    ptx::tmem_st_wait();
    ptx::tc_fence_before();
    __syncwarp();
    if (lane == 0) ptx::mbar_arrive(&q_full[0]);
    // Head epilogue (the caller has waited for the head's last P V): read this thread's 16 O columns, add the four
    // partial row sums (shared memory + the row quarter's named barrier 1 + wq, 128 threads), normalise, split to
    // bf16 hi | lo and store.
    auto finish_head = [&](int g, float l) {
      const int col = (h0 + g) * kDH + part * PC;
      ptx::tc_fence_after();
      uint32_t ov[16];
      ptx::tmem_ld_32x16(t_row + kColO + (g & 1) * 64 + part * PC, ov);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&o_empty[g & 1]);
      float* xl = xsum + (g & 1) * kParts * kBQ;   // slot g & 1: its next use is two barriers away
      xl[part * kBQ + row] = l;
      FY_AW(13, asm volatile("bar.sync %0, 128;\n" ::"r"(1 + wq) : "memory"))
      const float lt = (xl[row] + xl[kBQ + row]) + (xl[2 * kBQ + row] + xl[3 * kBQ + row]);
      if (q0 + row < L) {
        const float inv = 1.0f / lt;
        __nv_bfloat16* dst = p.out + (size_t)(P + q0 + row) * p.ld + col;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          float a[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) a[j] = __uint_as_float(ov[8 * c + j]) * inv;
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
        for (int c = 0; c < 2; ++c) {
          *reinterpret_cast<uint4*>(dst + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
          *reinterpret_cast<uint4*>(dst + p.lo_off + 8 * c) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
    };
    // One flat loop over the CTA's tiles T = g * nkt + t.  A head's normalisation and store are deferred to the first
    // tile of the next head (after its P has been handed to the tensor core), so they run under S / P V products
    // instead of leaving the tensor pipe idle.  Inside the loop o_full is only waited for BEFORE this warp's p_full
    // arrival of the current tile T >= 1: P_T V_T cannot have been issued yet and s_full of tile T implies that
    // P_{T-2} V_{T-2} has completed (S_T is issued after it), so the barrier is in phase T - 1 or T and the parity of
    // T - 1 is unambiguous.  After the loop that argument no longer holds, hence the separate all_done barrier.
    float m = -INFINITY, l = 0.f, l_prev = 0.f;
    int g = 0, t = 0;
#pragma unroll 1
    for (int T = 0; T < total; ++T) {
      const bool stage_q = t == 0 && g + 1 < kHeadsPerCta;
      if (stage_q) load_q(g + 1);          // lands under this tile's softmax; stored after P below
      if (t == 0) { l_prev = l; l = 0.f; }
      FY_AW(11, ptx::mbar_wait(&s_full[T & 1], (T >> 1) & 1))
      ptx::tc_fence_after();
      uint32_t sv[16];
      ptx::tmem_ld_32x16(t_row + (T & 1) * kBK + part * PC, sv);
      ptx::tmem_ld_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(&s_empty[T & 1]);
      float s[PC];
#pragma unroll
      for (int j = 0; j < PC; ++j) s[j] = __uint_as_float(sv[j]);
      if ((t + 1) * kBK > L) {           // the head's last tile may be partial: keys >= L are masked
#pragma unroll
        for (int j = 0; j < PC; ++j)
          if (t * kBK + part * PC + j >= L) s[j] = -INFINITY;
      }
      float mx4[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) mx4[j] = fmaxf(fmaxf(s[j], s[4 + j]), fmaxf(s[8 + j], s[12 + j]));
      // row maximum over the four parts: slot T & 1 of the exchange buffer (its next use is two barriers away)
      float* xm = xmax + (T & 1) * kParts * kBQ;
      xm[part * kBQ + row] = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      FY_AW(13, asm volatile("bar.sync %0, 128;\n" ::"r"(1 + wq) : "memory"))
      const float mx = fmaxf(fmaxf(xm[row], xm[kBQ + row]), fmaxf(xm[2 * kBQ + row], xm[3 * kBQ + row])) * kLog2e;
      // Running maximum (finite: every tile holds a valid key).  The four threads of a row see the same mx and take
      // the same decisions.
      if (t == 0) {
        m = mx;                          // a head's first P V overwrites O: nothing to rescale
      } else {
        const bool need = mx > m + p.slack;
        if (__any_sync(0xffffffffu, need)) {
          // rare: rescale this thread's 16 O columns in tensor memory (P_{T-1} V_{T-1} must have landed)
          FY_AW(12, ptx::mbar_wait(o_full, (T - 1) & 1))
          ptx::tc_fence_after();
          const float corr = need ? ex2(m - mx) : 1.0f;
          uint32_t ov[16];
          const uint32_t oa = t_row + kColO + (g & 1) * 64 + part * PC;
          ptx::tmem_ld_32x16(oa, ov);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 16; ++j) ov[j] = __float_as_uint(__uint_as_float(ov[j]) * corr);
          ptx::tmem_st_32x16(oa, ov);
          l *= corr;
          if (need) m = mx;
        }
      }
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int j = 0; j < PC; ++j) {
        s[j] = ex2(fmaf(s[j], kLog2e, -m));
        sum4[j & 3] += s[j];
      }
      l += (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);   // partial (this thread's keys); added up in finish_head
      // P row -> tensor memory: keys [16 part, 16 part + 16) are columns [8 part, 8 part + 8) of the hi and of the lo
      // block of buffer T & 1 (two bf16 per column, even key in the low half-word)
      {
        uint32_t ph[8], pl[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          ph[j] = pack2(s[2 * j], s[2 * j + 1]);
          pl[j] = pack_lo(ph[j], s[2 * j], s[2 * j + 1]);
        }
        const uint32_t pa = t_row + kColP + (T & 1) * 64 + part * 8;
        ptx::tmem_st_32x8(pa, ph);
        ptx::tmem_st_32x8(pa + 32, pl);
      }
      // Q of the next head goes into the buffer head g - 1 used: s_full of this tile implies every earlier product
      // (all S of head g - 1) has completed.
      if (stage_q) store_q(g + 1);
      const bool head_done = t == 0 && T > 0;           // head g - 1 is complete once P_{T-1} V_{T-1} has landed
      if (head_done) FY_AW(12, ptx::mbar_wait(o_full, (T - 1) & 1))
      ptx::tmem_st_wait();
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(&p_full[T & 1]);
        if (stage_q) ptx::mbar_arrive(&q_full[(g + 1) & 1]);
      }
      if (head_done) finish_head(g - 1, l_prev);
      if (++t == nkt) { t = 0; ++g; }
    }
    ptx::mbar_wait(all_done, 0);
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
