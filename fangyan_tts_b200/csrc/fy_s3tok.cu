// S3 speech tokenizer v3 encoder + FSQ (rows A9, A10): weight preparation, per-call planning, launch sequence.
// Replaces ort_session.run() on speech_tokenizer_v3.onnx (tools/extract_speech_token.py:40-48).
//
// Activation layout: utterances are PACKED along the row (time) axis, never padded.  Utterance u owns token rows
// [P_u, P_u + L_u) with one spare row between utterances; its conv1 outputs sit at rows 2*P_u + 1 + t and its
// mel frames at rows 4*P_u + 3 + t.  With that alignment both stride-2 convolutions are plain GEMMs over a
// "two rows per row" view of their input (taps 0,1 = the two halves of view row g, tap 2 = first half of row g+1),
// and the spare rows -- always written as zero -- are the convolutions' zero padding.
#include <math.h>
#include <string.h>

#include <algorithm>
#include <initializer_list>

#include "fy_attention.cuh"
#include "fy_attention_pp.cuh"
#include "fy_attention_tc.cuh"
#include "fy_common.cuh"
#include "fy_plan.cuh"
#include "fy_s3_kernels.cuh"

namespace fy {
namespace {

struct LayerW {
  float *ln1_g, *ln1_b, *ln2_g, *ln2_b;
  __nv_bfloat16 *wqkv, *wo, *w1, *w2;  // [N, K*W] hi | lo
  float *bqkv, *bo, *b1, *b2;
  float* fsmn;                          // [KS][D]
  CUtensorMap m_qkv, m_o, m_1, m_2;
};

struct S3Plan {
  fy_s3_config cfg;
  int W = 1;  // 2 when split-bf16
  __nv_bfloat16 *wc1 = nullptr, *wc2 = nullptr;
  float *bc1 = nullptr, *bc2 = nullptr;
  CUtensorMap m_c1, m_c2;
  std::vector<LayerW> layers;
  float *pd_w = nullptr, *pd_b = nullptr;
  float *rope_cos = nullptr, *rope_sin = nullptr;
  int rope_len = 0;
  std::vector<void*> owned;
  bool attn_configured = false;   // cudaFuncSetAttribute done for this context's device
  float* stage = nullptr;         // fp32 staging of one weight matrix during fy_load_s3tok
  size_t stage_cap = 0;
};

const fy_tensor* find(const fy_tensor* t, int n, const std::string& name) {
  for (int i = 0; i < n; ++i)
    if (name == t[i].name) return &t[i];
  return nullptr;
}

int64_t numel(const fy_tensor* t) {
  int64_t n = 1;
  for (int i = 0; i < t->ndim; ++i) n *= t->shape[i];
  return n;
}

// Full-shape check (an element count alone lets a transposed square matrix through).
bool shape_is(const fy_tensor* t, std::initializer_list<int64_t> want) {
  if (t->ndim != (int)want.size()) return false;
  int i = 0;
  for (int64_t w : want)
    if (t->shape[i++] != w) return false;
  return true;
}
#define FY_SHAPE(t, ...)                                                                                           \
  FY_CHECK(shape_is(t, {__VA_ARGS__}), FY_ERR_ARG, "fy_load_s3tok: tensor '%s' has the wrong shape (rank %d, first dims %lld x %lld)", \
           (t)->name, (t)->ndim, (long long)(t)->shape[0], (long long)((t)->ndim > 1 ? (t)->shape[1] : 1))

template <typename T>
int dev_upload(S3Plan* p, const std::vector<T>& h, T** d) {
  FY_CUDA(cudaMalloc(d, h.size() * sizeof(T)));
  p->owned.push_back(*d);
  FY_CUDA(cudaMemcpy(*d, h.data(), h.size() * sizeof(T), cudaMemcpyHostToDevice));
  return FY_OK;
}

// fp32 [N, K] -> bf16 [N, K*W]: hi at [0, K), lo = bf16(x - hi) at [K, 2K) when W == 2.  The split runs on the device
// (one streaming kernel per matrix over a staged fp32 copy): a 12-layer encoder has 236 M weights, which the host
// loop this replaces took seconds to convert.
__global__ void split_weight_kernel(const float* __restrict__ w, long long n, int K, int W, __nv_bfloat16* __restrict__ out) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / K;
    const int k = (int)(i - r * K);
    const float x = w[i];
    const __nv_bfloat16 hi = __float2bfloat16_rn(x);
    out[r * K * W + k] = hi;
    if (W == 2) out[r * K * 2 + K + k] = __float2bfloat16_rn(x - __bfloat162float(hi));
  }
}

int upload_weight(S3Plan* p, const float* w, int64_t N, int64_t K, __nv_bfloat16** d) {
  const size_t n = (size_t)N * K;
  if (p->stage_cap < n) {
    if (p->stage) cudaFree(p->stage);
    p->stage = nullptr;
    FY_CUDA(cudaMalloc(&p->stage, n * sizeof(float)));
    p->stage_cap = n;
  }
  FY_CUDA(cudaMalloc(d, n * p->W * sizeof(__nv_bfloat16)));
  p->owned.push_back(*d);
  FY_CUDA(cudaMemcpy(p->stage, w, n * sizeof(float), cudaMemcpyHostToDevice));
  split_weight_kernel<<<1184, 256>>>(p->stage, (long long)n, (int)K, p->W, *d);
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

#define NEED(var, name)                                                           \
  const fy_tensor* var = find(t, n, name);                                        \
  FY_CHECK(var != nullptr, FY_ERR_ARG, "fy_load_s3tok: missing tensor '%s'", std::string(name).c_str())

int upload_f32(S3Plan* p, const fy_tensor* t, int64_t expect, float** d) {
  FY_CHECK(numel(t) == expect, FY_ERR_ARG, "fy_load_s3tok: tensor '%s' has %lld elements, expected %lld", t->name,
           (long long)numel(t), (long long)expect);
  std::vector<float> h(t->h_data, t->h_data + expect);
  return dev_upload(p, h, d);
}

// Conv1d weight [Cout, Cin, 3] -> [Cout, 3*Cin] tap-major.
std::vector<float> conv_tap_major(const float* w, int64_t co, int64_t ci, int64_t ks) {
  std::vector<float> r((size_t)co * ci * ks);
  for (int64_t o = 0; o < co; ++o)
    for (int64_t c = 0; c < ci; ++c)
      for (int64_t k = 0; k < ks; ++k) r[(o * ks + k) * ci + c] = w[(o * ci + c) * ks + k];
  return r;
}

// Three-product segment list for a K-range: hi*hi (+ hi*lo + lo*hi when split).
// a_col/b_col address the hi halves; lo halves sit a_lo / b_lo columns further.
// With a 256-wide tile (g.N set before the call, a multiple of 256) the split is one FUSED segment: each K block's
// four operand tiles are fetched once for the three products (FY_GEMM_FUSED_SPLIT=0 restores three segments).
void add_products(GemmArgs& g, int W, int a_row_off, int a_col, int b_col, int K, int a_lo, int b_lo) {
  const int nkb = K / 64;
  static const bool fused_ok = [] { const char* e = getenv("FY_GEMM_FUSED_SPLIT"); return !(e && e[0] == '0'); }();
  if (W == 2 && fused_ok && g.N > 0 && gemm_block_n(g.N) == 256) {
    g.seg[g.nseg++] = GemmSeg{a_row_off, a_col, b_col, nkb, a_lo, b_lo};
    return;
  }
  g.seg[g.nseg++] = GemmSeg{a_row_off, a_col, b_col, nkb};
  if (W == 2) {
    g.seg[g.nseg++] = GemmSeg{a_row_off, a_col, b_col + b_lo, nkb};
    g.seg[g.nseg++] = GemmSeg{a_row_off, a_col + a_lo, b_col, nkb};
  }
}

size_t align_up(size_t x, size_t a = 1024) { return (x + a - 1) / a * a; }

struct Layout {  // per-call sizes derived from the frame counts
  int n_utts = 0, Rtok = 0, R1 = 0, Rmel = 0, n_attn_tiles = 0, Tmax = 0;
  size_t o_meta, o_mel, o_c1, o_x, o_xn, o_q, o_k, o_v, o_att, o_h, total;
  // meta sub-offsets (ints)
  size_t m_P, m_L, m_t64, m_t128, m_row1, m_rutt, m_rpos, m_tiles, m_tiles128, meta_ints;
};

int token_len(int T) { const int l1 = (T - 1) / 2 + 1; return (l1 - 1) / 2 + 1; }

// Upper-bound layout for workspace sizing: rows from totals only.
Layout make_layout(const fy_s3_config& c, int W, int n_utts, int rtok, int n_tiles) {
  Layout l;
  l.n_utts = n_utts;
  l.Rtok = rtok;
  l.R1 = 2 * rtok + 2;
  l.Rmel = 2 * l.R1 + 2;
  l.n_attn_tiles = n_tiles;
  const size_t D = c.d_model;
  l.m_P = 0;
  l.m_L = l.m_P + n_utts;
  l.m_t64 = l.m_L + n_utts;
  l.m_t128 = l.m_t64 + n_utts;
  l.m_row1 = l.m_t128 + n_utts;
  l.m_rutt = l.m_row1 + l.R1;
  l.m_rpos = l.m_rutt + rtok;
  l.m_tiles = l.m_rpos + rtok;
  l.m_tiles += l.m_tiles & 1;  // int2 alignment
  l.m_tiles128 = l.m_tiles + 2 * (size_t)n_tiles;            // 128-row query tiles (tcgen05 attention), <= n_tiles
  l.meta_ints = l.m_tiles128 + 2 * (size_t)n_tiles;
  size_t o = 0;
  l.o_meta = o; o = align_up(o + l.meta_ints * 4);
  l.o_mel = o; o = align_up(o + (size_t)l.Rmel * c.n_mels * W * 2);
  l.o_c1 = o; o = align_up(o + (size_t)l.R1 * D * W * 2);
  l.o_x = o; o = align_up(o + (size_t)rtok * D * 4);
  l.o_xn = o; o = align_up(o + (size_t)rtok * D * W * 2);
  l.o_q = o; o = align_up(o + (size_t)rtok * D * W * 2);
  l.o_k = o; o = align_up(o + (size_t)rtok * D * W * 2);
  l.o_v = o; o = align_up(o + (size_t)rtok * D * W * 2);
  l.o_att = o; o = align_up(o + (size_t)rtok * D * W * 2);
  l.o_h = o; o = align_up(o + (size_t)rtok * c.d_ff * W * 2);
  l.total = o;
  return l;
}

}  // namespace

void s3_free_plan(Ctx* c) {
  S3Plan* p = reinterpret_cast<S3Plan*>(c->s3_plan);
  if (!p) return;
  for (void* d : p->owned) cudaFree(d);
  if (p->stage) cudaFree(p->stage);
  delete p;
  c->s3_plan = nullptr;
}

}  // namespace fy

using namespace fy;

extern "C" {

int fy_load_s3tok(fy_ctx* ctx, const fy_tensor* t, int n, const fy_s3_config* cfg) {
  FY_CHECK(ctx && t && cfg && n > 0, FY_ERR_ARG, "fy_load_s3tok: NULL argument");
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  FY_CUDA(cudaSetDevice(c->device));
  FY_CHECK(cfg->d_model % 256 == 0 && cfg->d_ff % 256 == 0 && cfg->n_mels % 64 == 0, FY_ERR_ARG,
           "fy_load_s3tok: d_model/d_ff must be multiples of 256 and n_mels of 64");
  FY_CHECK(cfg->d_model / cfg->n_head == 64, FY_ERR_ARG, "fy_load_s3tok: head_dim must be 64 (got %d)",
           cfg->d_model / cfg->n_head);
  FY_CHECK(cfg->fsq_dim == 8, FY_ERR_ARG, "fy_load_s3tok: fsq_dim must be 8");
  FY_CHECK(cfg->fsmn_kernel == s3k::kFsmnMaxK && cfg->d_model % 64 == 0, FY_ERR_ARG,
           "fy_load_s3tok: the FSMN kernel is instantiated for fsmn_kernel == %d only", s3k::kFsmnMaxK);
  s3_free_plan(c);
  S3Plan* p = new S3Plan();
  c->s3_plan = p;
  p->cfg = *cfg;
  p->W = c->precision == FY_PREC_BF16X3 ? 2 : 1;
  const int64_t D = cfg->d_model, F = cfg->d_ff, NM = cfg->n_mels, KS = cfg->fsmn_kernel;
  {
    NEED(w, "conv1.weight"); NEED(b, "conv1.bias");
    FY_SHAPE(w, D, NM, 3);
    auto r = conv_tap_major(w->h_data, D, NM, 3);
    FY_TRY(upload_weight(p, r.data(), D, 3 * NM, &p->wc1));
    FY_TRY(upload_f32(p, b, D, &p->bc1));
    FY_TRY(make_tmap_bf16(&p->m_c1, p->wc1, D, 3 * NM * p->W, 3 * NM * p->W, gemm_block_n((int)D)));
  }
  {
    NEED(w, "conv2.weight"); NEED(b, "conv2.bias");
    FY_SHAPE(w, D, D, 3);
    auto r = conv_tap_major(w->h_data, D, D, 3);
    FY_TRY(upload_weight(p, r.data(), D, 3 * D, &p->wc2));
    FY_TRY(upload_f32(p, b, D, &p->bc2));
    FY_TRY(make_tmap_bf16(&p->m_c2, p->wc2, D, 3 * D * p->W, 3 * D * p->W, gemm_block_n((int)D)));
  }
  p->layers.resize(cfg->n_layer);
  for (int i = 0; i < cfg->n_layer; ++i) {
    LayerW& L = p->layers[i];
    const std::string pre = "blocks." + std::to_string(i) + ".";
    NEED(g1, pre + "attn_ln.weight"); NEED(be1, pre + "attn_ln.bias");
    NEED(g2, pre + "mlp_ln.weight"); NEED(be2, pre + "mlp_ln.bias");
    NEED(wq, pre + "attn.query.weight"); NEED(bq, pre + "attn.query.bias");
    NEED(wk, pre + "attn.key.weight");
    NEED(wv, pre + "attn.value.weight"); NEED(bv, pre + "attn.value.bias");
    NEED(wo, pre + "attn.out.weight"); NEED(bo, pre + "attn.out.bias");
    NEED(fs, pre + "attn.fsmn.weight");
    NEED(w1, pre + "fc1.weight"); NEED(b1, pre + "fc1.bias");
    NEED(w2, pre + "fc2.weight"); NEED(b2, pre + "fc2.bias");
    FY_TRY(upload_f32(p, g1, D, &L.ln1_g)); FY_TRY(upload_f32(p, be1, D, &L.ln1_b));
    FY_TRY(upload_f32(p, g2, D, &L.ln2_g)); FY_TRY(upload_f32(p, be2, D, &L.ln2_b));
    FY_SHAPE(wq, D, D); FY_SHAPE(wk, D, D); FY_SHAPE(wv, D, D); FY_SHAPE(wo, D, D);
    std::vector<float> qkv((size_t)3 * D * D), bqkv((size_t)3 * D, 0.f);
    memcpy(qkv.data(), wq->h_data, D * D * 4);
    memcpy(qkv.data() + D * D, wk->h_data, D * D * 4);
    memcpy(qkv.data() + 2 * D * D, wv->h_data, D * D * 4);
    memcpy(bqkv.data(), bq->h_data, D * 4);
    memcpy(bqkv.data() + 2 * D, bv->h_data, D * 4);
    FY_TRY(upload_weight(p, qkv.data(), 3 * D, D, &L.wqkv));
    FY_TRY(dev_upload(p, bqkv, &L.bqkv));
    FY_TRY(upload_weight(p, wo->h_data, D, D, &L.wo));
    FY_TRY(upload_f32(p, bo, D, &L.bo));
    FY_SHAPE(w1, F, D); FY_SHAPE(w2, D, F);
    FY_TRY(upload_weight(p, w1->h_data, F, D, &L.w1));
    FY_TRY(upload_f32(p, b1, F, &L.b1));
    FY_TRY(upload_weight(p, w2->h_data, D, F, &L.w2));
    FY_TRY(upload_f32(p, b2, D, &L.b2));
    FY_SHAPE(fs, D, 1, KS);
    std::vector<float> ft((size_t)KS * D);
    for (int64_t ch = 0; ch < D; ++ch)
      for (int64_t j = 0; j < KS; ++j) ft[j * D + ch] = fs->h_data[ch * KS + j];
    FY_TRY(dev_upload(p, ft, &L.fsmn));
    const int W = p->W;
    FY_TRY(make_tmap_bf16(&L.m_qkv, L.wqkv, 3 * D, D * W, D * W, gemm_block_n((int)(3 * D))));
    FY_TRY(make_tmap_bf16(&L.m_o, L.wo, D, D * W, D * W, gemm_block_n((int)D)));
    FY_TRY(make_tmap_bf16(&L.m_1, L.w1, F, D * W, D * W, gemm_block_n((int)F)));
    FY_TRY(make_tmap_bf16(&L.m_2, L.w2, D, F * W, F * W, gemm_block_n((int)D)));
  }
  {
    NEED(w, "project_down.weight"); NEED(b, "project_down.bias");
    FY_SHAPE(w, 8, D);
    FY_TRY(upload_f32(p, w, 8 * D, &p->pd_w));
    FY_TRY(upload_f32(p, b, 8, &p->pd_b));
  }
  {
    const int DH = 64;
    const fy_tensor* rc = find(t, n, "rope.cos");
    const fy_tensor* rs = find(t, n, "rope.sin");
    if (rc && rs) {
      FY_CHECK(rc->ndim == 2 && rc->shape[1] == DH && numel(rc) == numel(rs), FY_ERR_ARG, "rope tables must be [n, 64]");
      p->rope_len = (int)rc->shape[0];
      FY_TRY(upload_f32(p, rc, numel(rc), &p->rope_cos));
      FY_TRY(upload_f32(p, rs, numel(rs), &p->rope_sin));
    } else {
      p->rope_len = 2048;
      std::vector<float> hc((size_t)p->rope_len * DH), hs((size_t)p->rope_len * DH);
      for (int pos = 0; pos < p->rope_len; ++pos)
        for (int d = 0; d < DH / 2; ++d) {
          const float inv = 1.0f / powf(cfg->rope_theta, (float)(2 * d) / DH);
          const float a = (float)pos * inv;
          hc[(size_t)pos * DH + d] = hc[(size_t)pos * DH + d + DH / 2] = cosf(a);
          hs[(size_t)pos * DH + d] = hs[(size_t)pos * DH + d + DH / 2] = sinf(a);
        }
      FY_TRY(dev_upload(p, hc, &p->rope_cos));
      FY_TRY(dev_upload(p, hs, &p->rope_sin));
    }
  }
  FY_CUDA(cudaDeviceSynchronize());
  if (p->stage) { cudaFree(p->stage); p->stage = nullptr; p->stage_cap = 0; }
  FY_CUDA(cudaFuncSetAttribute(attn_tc::attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               attn_tc::kSmemBytes));
  FY_CUDA(cudaFuncSetAttribute(attn_pp::attention_pp_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               attn_pp::smem_bytes<1>()));
  FY_CUDA(cudaFuncSetAttribute(attn_pp::attention_pp_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               attn_pp::smem_bytes<2>()));
  p->attn_configured = true;
  FY_CUDA(cudaFuncSetAttribute(attn::attention_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               attn::smem_bytes<true>()));
  FY_CUDA(cudaFuncSetAttribute(attn::attention_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                               attn::smem_bytes<false>()));
  return FY_OK;
}

size_t fy_s3tok_workspace_bytes(fy_ctx* ctx, int n_utts, int total_frames) {
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  if (!c || !c->s3_plan || n_utts <= 0) return 0;
  S3Plan* p = reinterpret_cast<S3Plan*>(c->s3_plan);
  // tokens <= T/4 + 1 per utterance, plus one spare row each; attention tiles <= tokens/64 + 1 per utterance
  const int rtok = total_frames / 4 + 3 * n_utts + 1;
  const int tiles = rtok / 64 + n_utts + 1;
  return make_layout(p->cfg, p->W, n_utts, rtok, tiles).total;
}

int fy_s3tok_encode_fsq(fy_ctx* ctx, const float* mel, const int32_t* frame_off, const int32_t* h_frame_off,
                        int n_utts, int32_t* tokens, const int32_t* tok_off, float* fsq_pre, void* ws,
                        size_t ws_bytes) {
  FY_CHECK(ctx && mel && frame_off && h_frame_off && tokens && tok_off && ws, FY_ERR_ARG,
           "fy_s3tok_encode_fsq: NULL argument");
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  S3Plan* p = reinterpret_cast<S3Plan*>(c->s3_plan);
  FY_CHECK(p != nullptr, FY_ERR_STATE, "fy_s3tok_encode_fsq: call fy_load_s3tok first");
  FY_CHECK(n_utts > 0 && n_utts <= 65535, FY_ERR_ARG, "fy_s3tok_encode_fsq: n_utts must be in [1, 65535]");
  FY_CHECK((reinterpret_cast<uintptr_t>(ws) & 255) == 0, FY_ERR_ARG, "fy_s3tok_encode_fsq: workspace must be 256-byte aligned");
  const fy_s3_config& cf = p->cfg;
  const int W = p->W, D = cf.d_model, F = cf.d_ff, NM = cf.n_mels;
  cudaStream_t st = c->stream;

  // ---- plan the packed layout on the host
  std::vector<int> L(n_utts);
  int row = 0, tiles = 0, Tmax = 0;
  double vtok = 0, vc1 = 0, attn_flops = 0;  // valid rows, for algorithmic FLOP accounting
  for (int u = 0; u < n_utts; ++u) {
    const int T = h_frame_off[u + 1] - h_frame_off[u];
    FY_CHECK(T > 0, FY_ERR_ARG, "fy_s3tok_encode_fsq: utterance %d has %d mel frames", u, T);
    L[u] = token_len(T);
    FY_CHECK(L[u] <= p->rope_len, FY_ERR_ARG, "utterance %d has %d tokens, rope table holds %d", u, L[u], p->rope_len);
    row += L[u] + 1;
    vtok += L[u];
    vc1 += (T - 1) / 2 + 1;
    attn_flops += 4.0 * L[u] * L[u] * D;
    tiles += (L[u] + 63) / 64;
    Tmax = std::max(Tmax, T);
  }
  Layout lay = make_layout(cf, W, n_utts, row, tiles);
  FY_CHECK(lay.total <= ws_bytes, FY_ERR_NOMEM, "fy_s3tok_encode_fsq: workspace %zu < %zu bytes", ws_bytes, lay.total);
  // ---- metadata: built on the device from frame_off (fy_plan.cuh); the host only needed the totals above
  int tiles128 = 0;
  for (int u = 0; u < n_utts; ++u) tiles128 += (L[u] + 127) / 128;
  uint8_t* base = reinterpret_cast<uint8_t*>(ws);
  int* d_meta = reinterpret_cast<int*>(base + lay.o_meta);
  plan::s3_scan_kernel<<<1, plan::kScanThreads, 0, st>>>(frame_off, n_utts, d_meta + lay.m_P, d_meta + lay.m_L,
                                                         d_meta + lay.m_t64, d_meta + lay.m_t128);
  {
    const int top = std::max(std::max(lay.Rtok, lay.R1), tiles);
    plan::s3_fill_kernel<<<std::min((top + 255) / 256, 4 * c->num_sms), 256, 0, st>>>(
        frame_off, n_utts, d_meta + lay.m_P, d_meta + lay.m_L, d_meta + lay.m_t64, d_meta + lay.m_t128, lay.Rtok, lay.R1,
        tiles, tiles128, d_meta + lay.m_row1, d_meta + lay.m_rutt, d_meta + lay.m_rpos,
        reinterpret_cast<int2*>(d_meta + lay.m_tiles), reinterpret_cast<int2*>(d_meta + lay.m_tiles128));
  }
  FY_CUDA(cudaGetLastError());
  c->launches += 2;
  const int* d_P = d_meta + lay.m_P;
  const int* d_L = d_meta + lay.m_L;
  const int* d_row1 = d_meta + lay.m_row1;
  const int* d_rutt = d_meta + lay.m_rutt;
  const int* d_rpos = d_meta + lay.m_rpos;
  const int2* d_tiles = reinterpret_cast<const int2*>(d_meta + lay.m_tiles);
  __nv_bfloat16* b_mel = reinterpret_cast<__nv_bfloat16*>(base + lay.o_mel);
  __nv_bfloat16* b_c1 = reinterpret_cast<__nv_bfloat16*>(base + lay.o_c1);
  float* b_x = reinterpret_cast<float*>(base + lay.o_x);
  __nv_bfloat16* b_xn = reinterpret_cast<__nv_bfloat16*>(base + lay.o_xn);
  __nv_bfloat16* b_q = reinterpret_cast<__nv_bfloat16*>(base + lay.o_q);
  __nv_bfloat16* b_k = reinterpret_cast<__nv_bfloat16*>(base + lay.o_k);
  __nv_bfloat16* b_v = reinterpret_cast<__nv_bfloat16*>(base + lay.o_v);
  __nv_bfloat16* b_att = reinterpret_cast<__nv_bfloat16*>(base + lay.o_att);
  __nv_bfloat16* b_h = reinterpret_cast<__nv_bfloat16*>(base + lay.o_h);
  const int Rtok = lay.Rtok, R1 = lay.R1, Rmel = lay.Rmel;

  // ---- mel -> bf16 rows
  s3k::zero_gap_rows_kernel<<<n_utts + 1, 128, 0, st>>>(b_mel, (int)(NM * W * 2), frame_off, d_P, n_utts, Rmel);
  {
    dim3 grid((Tmax + 31) / 32, n_utts);
    FY_CHECK(NM == 128, FY_ERR_ARG, "mel_to_rows supports n_mels == 128 only");
    if (W == 2) s3k::mel_to_rows_kernel<true><<<grid, 256, 0, st>>>(mel, frame_off, d_P, b_mel);
    else s3k::mel_to_rows_kernel<false><<<grid, 256, 0, st>>>(mel, frame_off, d_P, b_mel);
    FY_CUDA(cudaGetLastError());
  }
  // ---- conv1: [R1, 3*NM] x [D, 3*NM]^T, bias + GELU, masked rows -> c1 (bf16 hi|lo)
  {
    GemmArgs g;
    FY_TRY(make_tmap_bf16(&g.tmap_a, b_mel, Rmel / 2, 2 * NM * W, 2 * NM * W, 128));
    g.tmap_b = p->m_c1;
    g.M = R1; g.N = D;
    const int rw = NM * W;  // one mel row in the 2-row view
    for (int tap = 0; tap < 3; ++tap)
      add_products(g, W, tap / 2, (tap % 2) * rw, tap * NM, NM, NM, 3 * NM);
    g.epi.mode = EPI_BF16; g.epi.act = ACT_GELU; g.epi.out = b_c1; g.epi.ldo = D * W;
    g.epi.lo_col_off = W == 2 ? D : -1; g.epi.bias = p->bc1; g.epi.row_idx = d_row1;
    FY_TRY(run_gemm(c, g, 2.0 * vc1 * D * 3 * NM));
  }
  // ---- conv2: [Rtok, 3*D] x [D, 3*D]^T, bias + GELU, masked rows -> x (fp32)
  {
    GemmArgs g;
    FY_TRY(make_tmap_bf16(&g.tmap_a, b_c1, R1 / 2, 2 * D * W, 2 * D * W, 128));
    g.tmap_b = p->m_c2;
    g.M = Rtok; g.N = D;
    const int rw = D * W;
    for (int tap = 0; tap < 3; ++tap)
      add_products(g, W, tap / 2, (tap % 2) * rw, tap * D, D, D, 3 * D);
    g.epi.mode = EPI_F32_STORE; g.epi.act = ACT_GELU; g.epi.out = b_x; g.epi.ldo = D;
    g.epi.bias = p->bc2; g.epi.row_idx = d_rutt;
    FY_TRY(run_gemm(c, g, 2.0 * vtok * D * 3 * D));
  }
  // ---- A-side tensor maps shared by all layers
  CUtensorMap m_xn, m_att, m_h;
  FY_TRY(make_tmap_bf16(&m_xn, b_xn, Rtok, D * W, D * W, 128));
  FY_TRY(make_tmap_bf16(&m_att, b_att, Rtok, D * W, D * W, 128));
  FY_TRY(make_tmap_bf16(&m_h, b_h, Rtok, F * W, F * W, 128));
  FY_CHECK(D == 1280, FY_ERR_ARG, "layernorm kernel is instantiated for d_model == 1280 only (got %d)", D);
  // Attention kernel: 2 (default) = two-group ping-pong tcgen05 kernel (both precisions), 1 = the single-group tcgen05
  // kernel (split-bf16 only), 0 = mma.sync.  TMA views of k, v and the 128-row tile list are shared by the two tcgen05
  // kernels.
  static const int attn_env = [] { const char* e = getenv("FY_ATTN_TC"); return e ? atoi(e) : 2; }();
  const bool heads_ok = cf.d_model / cf.n_head == 64 && cf.n_head % 4 == 0;
  const bool attn_pp_on = attn_env >= 2 && heads_ok;
  const bool attn_tc_on = attn_env == 1 && W == 2 && heads_ok;
  static const float attn_slack = [] { const char* e = getenv("FY_ATTN_SLACK"); return e ? (float)atof(e) : attn_tc::kRescaleSlack; }();
  attn_tc::Params tc_params;
  attn_pp::Params pp_params;
  if (attn_tc_on) {
    FY_TRY(make_tmap_bf16(&tc_params.tm_kv, b_k, Rtok, D * W, D * W, 64));
    FY_TRY(make_tmap_bf16(&tc_params.tm_v, b_v, Rtok, D * W, D * W, 64));
    tc_params.q = b_q; tc_params.out = b_att; tc_params.ld = D * W; tc_params.lo_off = D; tc_params.rows = Rtok;
    tc_params.slack = attn_slack;
    tc_params.tiles = reinterpret_cast<const int2*>(d_meta + lay.m_tiles128);
    tc_params.utt_P = d_P; tc_params.utt_L = d_L;
  }
  if (attn_pp_on) {
    FY_TRY(make_tmap_bf16(&pp_params.tm_k, b_k, Rtok, D * W, D * W, 64));
    FY_TRY(make_tmap_bf16(&pp_params.tm_v, b_v, Rtok, D * W, D * W, 64));
    FY_TRY(make_tmap_bf16(&pp_params.tm_q, b_q, Rtok, D * W, D * W, 128));
    pp_params.q = b_q; pp_params.out = b_att; pp_params.ld = D * W; pp_params.lo_off = D; pp_params.rows = Rtok;
    pp_params.slack = attn_slack;
    pp_params.tiles = reinterpret_cast<const int2*>(d_meta + lay.m_tiles128);
    pp_params.utt_P = d_P; pp_params.utt_L = d_L;
  }
  const float qk_scale = powf(64.0f, -0.25f);
  const int ln_grid = (Rtok + 7) / 8;
  for (int i = 0; i < cf.n_layer; ++i) {
    const LayerW& Lw = p->layers[i];
    if (W == 2) s3k::layernorm_kernel<true, 1280><<<ln_grid, 256, 0, st>>>(b_x, Lw.ln1_g, Lw.ln1_b, cf.ln_eps, Rtok, b_xn);
    else s3k::layernorm_kernel<false, 1280><<<ln_grid, 256, 0, st>>>(b_x, Lw.ln1_g, Lw.ln1_b, cf.ln_eps, Rtok, b_xn);
    {
      GemmArgs g;
      g.tmap_a = m_xn; g.tmap_b = Lw.m_qkv; g.M = Rtok; g.N = 3 * D;
      add_products(g, W, 0, 0, 0, D, D, D);
      g.epi.mode = EPI_QKV_ROPE; g.epi.out = b_q; g.epi.out2 = b_k; g.epi.out3 = b_v; g.epi.ldo = D * W;
      g.epi.lo_col_off = W == 2 ? D : -1; g.epi.bias = Lw.bqkv;
      g.epi.rope_cos = p->rope_cos; g.epi.rope_sin = p->rope_sin; g.epi.rope_pos = d_rpos; g.epi.rope_D = D;
      g.epi.rope_scale = qk_scale;
      FY_TRY(run_gemm(c, g, 2.0 * vtok * 3 * D * D));
    }
    {
      dim3 fg(tiles, D / s3k::kFsmnCols);
      if (W == 2) s3k::fsmn_kernel<31, true><<<fg, 128, 0, st>>>(b_v, D * W, D, d_tiles, d_P, d_L, Lw.fsmn, D, b_x);
      else s3k::fsmn_kernel<31, false><<<fg, 128, 0, st>>>(b_v, D * W, D, d_tiles, d_P, d_L, Lw.fsmn, D, b_x);
    }
    if (attn_pp_on) {
      const int pi = prof_begin(c, PK_ATTN);
      dim3 grid(tiles128, cf.n_head / attn_pp::kHeadsPerCta);
      if (W == 2) attn_pp::attention_pp_kernel<2><<<grid, attn_pp::kThreads, attn_pp::smem_bytes<2>(), st>>>(pp_params);
      else attn_pp::attention_pp_kernel<1><<<grid, attn_pp::kThreads, attn_pp::smem_bytes<1>(), st>>>(pp_params);
      prof_end(c, pi, attn_flops);
      c->launches += 4;
    } else if (attn_tc_on) {
      const int pi = prof_begin(c, PK_ATTN);
      dim3 grid(tiles128, cf.n_head / attn_tc::kHeadsPerCta);
      attn_tc::attention_tc_kernel<<<grid, attn_tc::kThreads, attn_tc::kSmemBytes, st>>>(tc_params);
      prof_end(c, pi, attn_flops);
      c->launches += 4;
    } else {
      const int pi = prof_begin(c, PK_ATTN);
      dim3 grid(tiles, cf.n_head);
      if (W == 2) attn::attention_kernel<true><<<grid, attn::kThreads, attn::smem_bytes<true>(), st>>>(b_q, b_k, b_v, b_att, D * W, D, d_tiles, d_P, d_L);
      else attn::attention_kernel<false><<<grid, attn::kThreads, attn::smem_bytes<false>(), st>>>(b_q, b_k, b_v, b_att, D * W, D, d_tiles, d_P, d_L);
      prof_end(c, pi, attn_flops);
      c->launches += 4;  // layernorm x2, fsmn, attention
    }
    FY_CUDA(cudaGetLastError());
    {
      GemmArgs g;
      g.tmap_a = m_att; g.tmap_b = Lw.m_o; g.M = Rtok; g.N = D;
      add_products(g, W, 0, 0, 0, D, D, D);
      g.epi.mode = EPI_F32_ADD; g.epi.out = b_x; g.epi.ldo = D; g.epi.bias = Lw.bo;
      FY_TRY(run_gemm(c, g, 2.0 * vtok * D * D));
    }
    const float eps2 = cf.mlp_ln_eps > 0.f ? cf.mlp_ln_eps : cf.ln_eps;
    if (W == 2) s3k::layernorm_kernel<true, 1280><<<ln_grid, 256, 0, st>>>(b_x, Lw.ln2_g, Lw.ln2_b, eps2, Rtok, b_xn);
    else s3k::layernorm_kernel<false, 1280><<<ln_grid, 256, 0, st>>>(b_x, Lw.ln2_g, Lw.ln2_b, eps2, Rtok, b_xn);
    {
      GemmArgs g;
      g.tmap_a = m_xn; g.tmap_b = Lw.m_1; g.M = Rtok; g.N = F;
      add_products(g, W, 0, 0, 0, D, D, D);
      g.epi.mode = EPI_BF16; g.epi.act = ACT_GELU; g.epi.out = b_h; g.epi.ldo = F * W;
      g.epi.lo_col_off = W == 2 ? F : -1; g.epi.bias = Lw.b1;
      FY_TRY(run_gemm(c, g, 2.0 * vtok * F * D));
    }
    {
      GemmArgs g;
      g.tmap_a = m_h; g.tmap_b = Lw.m_2; g.M = Rtok; g.N = D;
      add_products(g, W, 0, 0, 0, F, F, F);
      g.epi.mode = EPI_F32_ADD; g.epi.out = b_x; g.epi.ldo = D; g.epi.bias = Lw.b2;
      FY_TRY(run_gemm(c, g, 2.0 * vtok * D * F));
    }
  }
  c->launches += 3;  // zero_gap_rows, mel_to_rows, fsq
  s3k::fsq_kernel<<<(Rtok + 7) / 8, 256, 0, st>>>(b_x, d_rutt, d_rpos, tok_off, p->pd_w, p->pd_b, Rtok, D, cf.fsq_scale,
                                                 tokens, fsq_pre);
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

}  // extern "C"

#ifdef FY_STALLPROF
// Debug build only: the ping-pong attention kernel's counters (32 x u64).
extern "C" int fy_debug_pp_stalls(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out, fy::attn_pp::g_pp_stall, sizeof(unsigned long long) * 32) != cudaSuccess) return 1;
  if (reset) {
    unsigned long long z[32] = {0};
    if (cudaMemcpyToSymbol(fy::attn_pp::g_pp_stall, z, sizeof(z)) != cudaSuccess) return 1;
  }
  return 0;
}
// Debug build only: copy the attention role-stall counters to the host (16 x u64) and optionally clear them.
extern "C" int fy_debug_attn_stalls(unsigned long long* out, int reset) {
  cudaDeviceSynchronize();
  if (cudaMemcpyFromSymbol(out, fy::attn_tc::g_attn_stall, sizeof(unsigned long long) * 16) != cudaSuccess) return 1;
  if (reset) {
    unsigned long long z[16] = {0};
    if (cudaMemcpyToSymbol(fy::attn_tc::g_attn_stall, z, sizeof(z)) != cudaSuccess) return 1;
  }
  return 0;
}
#endif
