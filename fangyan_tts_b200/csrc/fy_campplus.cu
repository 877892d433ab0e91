#include "fy_common.cuh"
namespace fy { void cam_free_plan(Ctx*) {} }
using namespace fy;
extern "C" {
int fy_load_campplus(fy_ctx*, const fy_tensor*, int) { set_error("not implemented"); return FY_ERR_STATE; }
size_t fy_campplus_workspace_bytes(fy_ctx*, int, int) { return 0; }
int fy_campplus_forward(fy_ctx*, const float*, const int32_t*, const int32_t*, int, float*, void*, size_t) { set_error("not implemented"); return FY_ERR_STATE; }
}
