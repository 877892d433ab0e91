#include "fy_common.cuh"
namespace fy { void s3_free_plan(Ctx*) {} }
using namespace fy;
extern "C" {
int fy_load_s3tok(fy_ctx*, const fy_tensor*, int, const fy_s3_config*) { set_error("not implemented"); return FY_ERR_STATE; }
size_t fy_s3tok_workspace_bytes(fy_ctx*, int, int) { return 0; }
int fy_s3tok_encode_fsq(fy_ctx*, const float*, const int32_t*, const int32_t*, int, int32_t*, const int32_t*, float*, void*, size_t) { set_error("not implemented"); return FY_ERR_STATE; }
}
