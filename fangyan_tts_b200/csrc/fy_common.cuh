// Shared host-side declarations for the fangyan_b200 library (context, errors, kernel launch protos).
#pragma once
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <map>
#include <string>
#include <vector>

#include "../../include/fangyan_b200.h"

namespace fy {

// ---------------------------------------------------------------- errors
void set_error(const char* fmt, ...);
#define FY_CHECK(cond, code, ...)  \
  do {                             \
    if (!(cond)) {                 \
      fy::set_error(__VA_ARGS__);  \
      return (code);               \
    }                              \
  } while (0)
#define FY_CUDA(expr)                                                                    \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      fy::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return FY_ERR_CUDA;                                                                \
    }                                                                                    \
  } while (0)
#define FY_TRY(expr)             \
  do {                           \
    int _r = (expr);             \
    if (_r != FY_OK) return _r;  \
  } while (0)

// ---------------------------------------------------------------- GEMM
constexpr int kMaxSeg = 32;
struct GemmSeg {
  int a_row_off;  // added to the tile's first row for the A load (may be negative: TMA zero-fills)
  int a_col;      // first K column in A
  int b_col;      // first K column in B
  int nkb;        // number of 64-wide K blocks
  // Fused split-bf16 segment (a_lo != 0): per K block the operands A_hi, A_lo, B_hi, B_lo are loaded ONCE into two
  // consecutive ring slots ((A_hi, B_hi), (A_lo, B_lo)) and feed three products hi*hi + hi*lo + lo*hi, instead of
  // three separate segments that fetch six tiles.  a_lo / b_lo = column distance from the hi to the lo half.
  // A GEMM uses either fused segments only or none (ring-slot parity), and only kernels with an even ring depth.
  int a_lo = 0;
  int b_lo = 0;
};

// EPI_QKV_ROPE: the fused q|k|v projection of the S3 encoder.  Column block n0 / rope_D selects q, k or v; q and k get
// bias, rotary embedding (rotate-half inside each 64-wide head) and the d^-1/4 scale, v gets bias only; all three are
// written as bf16 hi | lo operands for the attention / FSMN kernels (out = q, out2 = k, out3 = v).
enum EpiMode { EPI_BF16 = 0, EPI_F32_ADD = 1, EPI_F32_STORE = 2, EPI_QKV_ROPE = 3 };
// Operand types: bf16 and fp16 use 64-element K blocks (kind::f16), tf32 reads fp32 data in 32-element K blocks.
enum GemmOp { OP_BF16 = 0, OP_TF32 = 1, OP_F16 = 2 };
enum EpiAct { ACT_NONE = 0, ACT_GELU = 1, ACT_RELU = 2 };

struct GemmEpi {
  int mode = EPI_BF16;
  int act = ACT_NONE;
  void* out = nullptr;   // bf16 or float
  int ldo = 0;           // leading dimension of out (elements)
  int col_off = 0;       // column offset of this GEMM's N range inside out
  int lo_col_off = -1;   // EPI_BF16 only: if >= 0 also write the bf16 residual (x - bf16(x)) at this column offset
  const float* scale = nullptr;    // [N] per-column multiplier (folded BatchNorm) or null
  const float* bias = nullptr;     // [N] per-column addend or null
  const int* row_idx = nullptr;    // [M] or null: row_idx == -1 -> row written as zero; -2 -> row not written
  const float* row_gate = nullptr; // [*, N] or null: multiply by row_gate[row_idx[row] * N + n] (CAM mask)
  const __nv_bfloat16* resid = nullptr;  // EPI_BF16 only: bf16 tensor (same rows / ldo / col_off as out) added before act
  int out_f16 = 0;                       // EPI_BF16 only: out (and resid) hold IEEE fp16 instead of bf16
  const float* resid_f32 = nullptr;      // fp32 modes: fp32 tensor (same rows / ldo / col_off as out) added before act
  int round_tf32 = 0;                    // fp32 modes: round the stored value to tf32 (operand of a later tf32 GEMM)
  // EPI_QKV_ROPE only
  void* out2 = nullptr;
  void* out3 = nullptr;
  const float* rope_cos = nullptr;   // [positions, 64], both halves equal
  const float* rope_sin = nullptr;
  const int* rope_pos = nullptr;     // [M] position of each row inside its utterance (-1 on guard rows)
  int rope_D = 0;                    // d_model (columns per projection)
  float rope_scale = 1.0f;
  // output row remap: out_row = (row / remap_q) * remap_fp + row % remap_q + remap_add   (remap_q == 0: identity)
  int remap_q = 0, remap_fp = 0, remap_add = 0;
};

struct GemmArgs {
  CUtensorMap tmap_a;
  CUtensorMap tmap_b;
  int M = 0, N = 0;
  int op = 0;    // GemmOp: operand type of A and B
  int nseg = 0;
  GemmSeg seg[kMaxSeg];
  GemmEpi epi;
  // Optional A-side transform (fp16 operands, single-CTA kernels, no fused segments): every A tile is rewritten in shared
  // memory as relu(a * a_scale[k] + a_shift[k]) (k = the tile's A column) between its TMA arrival and its MMA, by the two
  // otherwise idle control warps.  This is how CAMPPlus folds BatchNorm + ReLU of the growing concat buffer into the
  // bottleneck / transit GEMMs instead of a separate pass over the buffer.  Tables: fp16, padded to a multiple of 64.
  const __half* a_scale = nullptr;
  const __half* a_shift = nullptr;
};

// Encode a 2-D bf16 row-major [rows, cols] tensor map (ld in elements) with a {64, box_rows} box, 128B swizzle.
int make_tmap_bf16(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
int make_tmap_f32(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows);
int gemm_block_n(int N);  // tile width the launcher picks for this N (32, 128 or 256)
int launch_gemm(const GemmArgs& a, int num_sms, cudaStream_t st);

// ---------------------------------------------------------------- context
struct Tensor {
  void* ptr = nullptr;   // device
  size_t bytes = 0;
  std::vector<int64_t> shape;
};

struct Ctx {
  int device = 0;
  int num_sms = 148;
  int precision = FY_PREC_BF16X3;
  cudaStream_t stream = nullptr;
  // front-end tables (device)
  float* fb_window = nullptr;      // [400] povey
  float* fb_mel_w = nullptr;       // packed mel weights
  int* fb_mel_off = nullptr;       // [80] first fft bin
  int* fb_mel_len = nullptr;       // [80]
  int* fb_mel_ptr = nullptr;       // [80] offset into fb_mel_w
  int fb_mel_nw = 0, wh_mel_nw = 0; // packed weights in fb_mel_w / wh_mel_w
  float2* fb_tw = nullptr;         // [256] twiddles for N=512
  float* wh_window = nullptr;      // [400] hann periodic
  float* wh_mel_w = nullptr;
  int* wh_mel_off = nullptr;
  int* wh_mel_len = nullptr;
  int* wh_mel_ptr = nullptr;
  float2* wh_tw = nullptr;         // [400] twiddles for N=400
  float* mm_window = nullptr;      // [1920] hann periodic (matcha 24 kHz mel)
  float* mm_mel_w = nullptr;
  int* mm_mel_off = nullptr;
  int* mm_mel_len = nullptr;
  int* mm_mel_ptr = nullptr;
  float2* mm_tw = nullptr;         // [1920] twiddles for N=1920
  // model weights (device, owned)
  std::map<std::string, Tensor> cam_w;
  std::map<std::string, Tensor> s3_w;
  void* cam_plan = nullptr;
  void* s3_plan = nullptr;
  std::map<std::pair<int, int>, float*> resample_kern;  // (orig, new) reduced -> device filter bank
  std::vector<void*> owned;        // misc device allocations freed in destroy
  // optional per-kernel-class timing (fy_profile_*): CUDA events recorded on the launch stream
  struct ProfRec { int kind; double flops; cudaEvent_t a, b; };
  bool prof = false;
  std::vector<ProfRec> recs;
  long long launches = 0;          // kernels launched since the last fy_profile_read
};

enum ProfKind { PK_GEMM = 0, PK_ATTN = 1, PK_OTHER = 2, PK_COUNT = 3 };
int prof_begin(Ctx* c, int kind);                 // returns a record index or -1 when profiling is off
void prof_end(Ctx* c, int idx, double flops);
int run_gemm(Ctx* c, const GemmArgs& g, double alg_flops);  // launch_gemm + launch counting + profiling

}  // namespace fy
