// C-ABI entry points: context lifetime, errors, the generic GEMM and the speaker mean (row A7).
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "fy_common.cuh"

namespace fy {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int frontend_init(Ctx* c);   // fy_frontend.cu
void s3_free_plan(Ctx* c);   // fy_s3tok.cu
void cam_free_plan(Ctx* c);  // fy_campplus.cu

int prof_begin(Ctx* c, int kind) {
  if (!c->prof) return -1;
  Ctx::ProfRec r;
  r.kind = kind;
  r.flops = 0;
  if (cudaEventCreate(&r.a) != cudaSuccess || cudaEventCreate(&r.b) != cudaSuccess) return -1;
  cudaEventRecord(r.a, c->stream);
  c->recs.push_back(r);
  return (int)c->recs.size() - 1;
}
void prof_end(Ctx* c, int idx, double flops) {
  if (idx < 0) return;
  c->recs[idx].flops = flops;
  cudaEventRecord(c->recs[idx].b, c->stream);
}
int run_gemm(Ctx* c, const GemmArgs& g, double alg_flops) {
  const int pi = prof_begin(c, PK_GEMM);
  const int r = launch_gemm(g, c->num_sms, c->stream);
  prof_end(c, pi, alg_flops);
  c->launches += 1;
  return r;
}

namespace {

__global__ void speaker_sum_kernel(const float* __restrict__ emb, const int* __restrict__ spk, int n_utts, int n_spk,
                                   float* __restrict__ sum, int* __restrict__ count) {
  // one block per speaker, 192 threads: deterministic accumulation in utterance order.
  const int s = blockIdx.x;
  const int d = threadIdx.x;
  float acc = 0.0f;
  int cnt = 0;
  for (int u = 0; u < n_utts; ++u) {
    if (spk[u] == s) {
      acc += emb[(size_t)u * 192 + d];
      ++cnt;
    }
  }
  sum[(size_t)s * 192 + d] += acc;
  if (d == 0) count[s] += cnt;
}

__global__ void speaker_finish_kernel(const float* __restrict__ sum, const int* __restrict__ count, int n_spk,
                                      float* __restrict__ mean) {
  const int s = blockIdx.x;
  const int d = threadIdx.x;
  const int c = count[s];
  mean[(size_t)s * 192 + d] = c > 0 ? sum[(size_t)s * 192 + d] / (float)c : 0.0f;
}

}  // namespace
}  // namespace fy

using namespace fy;

extern "C" {

const char* fy_version(void) { return "fangyan_b200 0.1 (sm_100a)"; }
const char* fy_last_error(void) { return g_err; }

int fy_create(int device, int precision, fy_ctx** out) {
  FY_CHECK(out != nullptr, FY_ERR_ARG, "fy_create: out is NULL");
  FY_CHECK(precision == FY_PREC_BF16 || precision == FY_PREC_BF16X3, FY_ERR_ARG, "fy_create: bad precision %d",
           precision);
  int n = 0;
  FY_CUDA(cudaGetDeviceCount(&n));
  FY_CHECK(device >= 0 && device < n, FY_ERR_ARG, "fy_create: device %d not present (%d visible)", device, n);
  FY_CUDA(cudaSetDevice(device));
  cudaDeviceProp prop;
  FY_CUDA(cudaGetDeviceProperties(&prop, device));
  FY_CHECK(prop.major == 10, FY_ERR_STATE, "fy_create: device %d is sm_%d%d; this library is built for sm_100a only",
           device, prop.major, prop.minor);
  Ctx* c = new Ctx();
  c->device = device;
  c->num_sms = prop.multiProcessorCount;
  c->precision = precision;
  int r = frontend_init(c);
  if (r != FY_OK) {
    delete c;
    return r;
  }
  *out = reinterpret_cast<fy_ctx*>(c);
  return FY_OK;
}

void fy_destroy(fy_ctx* ctx) {
  if (!ctx) return;
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  cudaSetDevice(c->device);
  cudaDeviceSynchronize();
  s3_free_plan(c);
  cam_free_plan(c);
  for (void* p : c->owned) cudaFree(p);
  for (auto& kv : c->cam_w) cudaFree(kv.second.ptr);
  for (auto& kv : c->s3_w) cudaFree(kv.second.ptr);
  delete c;
}

int fy_set_stream(fy_ctx* ctx, void* cuda_stream) {
  FY_CHECK(ctx != nullptr, FY_ERR_ARG, "fy_set_stream: ctx is NULL");
  reinterpret_cast<Ctx*>(ctx)->stream = reinterpret_cast<cudaStream_t>(cuda_stream);
  return FY_OK;
}

int fy_profile_enable(fy_ctx* ctx, int on) {
  FY_CHECK(ctx != nullptr, FY_ERR_ARG, "fy_profile_enable: ctx is NULL");
  reinterpret_cast<Ctx*>(ctx)->prof = on != 0;
  return FY_OK;
}

// out[0..2] = ms per kind (gemm, attention, other), out[3..5] = algorithmic FLOPs per kind,
// out[6..8] = timed launches per kind, out[9] = all kernel launches since the last read.  Synchronises the stream.
int fy_profile_read(fy_ctx* ctx, double* out, int n_out) {
  FY_CHECK(ctx && out && n_out >= 10, FY_ERR_ARG, "fy_profile_read: bad argument");
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  FY_CUDA(cudaStreamSynchronize(c->stream));
  for (int i = 0; i < 10; ++i) out[i] = 0.0;
  for (auto& r : c->recs) {
    float ms = 0.f;
    cudaEventElapsedTime(&ms, r.a, r.b);
    out[r.kind] += ms;
    out[3 + r.kind] += r.flops;
    out[6 + r.kind] += 1.0;
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  c->recs.clear();
  out[9] = (double)c->launches;
  c->launches = 0;
  return FY_OK;
}

int fy_gemm_bf16(fy_ctx* ctx, const void* a, const void* b, float* out, int M, int N, int K) {
  FY_CHECK(ctx && a && b && out, FY_ERR_ARG, "fy_gemm_bf16: NULL argument");
  FY_CHECK(M > 0 && N > 0 && K > 0 && K % 8 == 0 && N % 32 == 0, FY_ERR_ARG,
           "fy_gemm_bf16: need K %% 8 == 0 and N %% 32 == 0 (M=%d N=%d K=%d)", M, N, K);
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  GemmArgs g;
  FY_TRY(make_tmap_bf16(&g.tmap_a, a, M, K, K, 128));
  FY_TRY(make_tmap_bf16(&g.tmap_b, b, N, K, K, gemm_block_n(N)));
  g.M = M;
  g.N = N;
  g.nseg = 1;
  g.seg[0] = GemmSeg{0, 0, 0, (K + 63) / 64};
  g.epi.mode = EPI_F32_STORE;
  g.epi.out = out;
  g.epi.ldo = N;
  return run_gemm(c, g, 2.0 * M * N * K);
}

int fy_speaker_sum(fy_ctx* ctx, const float* emb, const int32_t* spk_id, int n_utts, int n_spk, float* sum,
                   int32_t* count) {
  FY_CHECK(ctx && emb && spk_id && sum && count, FY_ERR_ARG, "fy_speaker_sum: NULL argument");
  FY_CHECK(n_utts >= 0 && n_spk > 0, FY_ERR_ARG, "fy_speaker_sum: bad sizes");
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  if (n_utts == 0) return FY_OK;
  speaker_sum_kernel<<<n_spk, 192, 0, c->stream>>>(emb, spk_id, n_utts, n_spk, sum, count);
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

int fy_speaker_mean_finish(fy_ctx* ctx, const float* sum, const int32_t* count, int n_spk, float* mean) {
  FY_CHECK(ctx && sum && count && mean && n_spk > 0, FY_ERR_ARG, "fy_speaker_mean_finish: bad argument");
  Ctx* c = reinterpret_cast<Ctx*>(ctx);
  speaker_finish_kernel<<<n_spk, 192, 0, c->stream>>>(sum, count, n_spk, mean);
  FY_CUDA(cudaGetLastError());
  return FY_OK;
}

}  // extern "C"
