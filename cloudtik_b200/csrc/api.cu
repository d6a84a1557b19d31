// C-ABI glue: error state, version, interaction dispatch.
#include <stdarg.h>

#include "common.cuh"

namespace b200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int interact_fwd_simt(const void*, void*, int64_t, int64_t, int64_t, int64_t, int, cudaStream_t, int self = 0);
int interact_bwd_simt(const void*, const void*, void*, int64_t, int64_t, int64_t, int64_t, int, cudaStream_t, int self = 0);
bool interact_tc_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr);
int interact_fwd_tc(const void*, void*, int64_t, int64_t, int64_t, int64_t, cudaStream_t);
int interact_bwd_tc(const void*, const void*, void*, int64_t, int64_t, int64_t, int64_t, cudaStream_t);
bool interact_tf32_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr);
int interact_fwd_tf32(const void*, void*, int64_t, int64_t, int64_t, int64_t, cudaStream_t);
bool interact_f32_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, const void* R, int64_t ldr);
int interact_fwd_f32(const void*, void*, int64_t, int64_t, int64_t, int64_t, cudaStream_t);
int interact_bwd_f32(const void*, const void*, void*, int64_t, int64_t, int64_t, int64_t, cudaStream_t);

}  // namespace b200

using namespace b200;

extern "C" const char* b200_last_error(void) { return g_err; }
extern "C" int b200_version(void) { return 100; }

static int check_interact(const void* a, const void* b, int64_t B, int64_t F, int64_t D, int64_t ldr, int dtype,
                          int self = 0) {
  B200_REQUIRE(B >= 0 && F >= 1 && D >= 1, B200_EINVAL, "interact: bad shape");
  B200_REQUIRE(ldr >= D + F * (F - 1) / 2 + (self ? F : 0), B200_EINVAL,
               "interact: row stride smaller than D + number of pairs");
  B200_REQUIRE(dtype == B200_F32 || dtype == B200_BF16, B200_EINVAL, "interact: bad dtype");
  B200_REQUIRE(B == 0 || (a && b), B200_EINVAL, "interact: null pointer");
  return B200_OK;
}

extern "C" int b200_interact_fwd(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr,
                                 int dtype, b200_stream_t stream) {
  int rc = check_interact(d_T, d_R, B, F, D, ldr, dtype);
  if (rc || B == 0) return rc;
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (interact_tc_supported(B, F, D, dtype, d_T, ldr)) return interact_fwd_tc(d_T, d_R, B, F, D, ldr, st);
  if (interact_tf32_supported(B, F, D, dtype, d_T, ldr)) return interact_fwd_tf32(d_T, d_R, B, F, D, ldr, st);
  if (interact_f32_supported(B, F, D, dtype, d_T, d_R, ldr)) return interact_fwd_f32(d_T, d_R, B, F, D, ldr, st);
  return interact_fwd_simt(d_T, d_R, B, F, D, ldr, dtype, st);
}

extern "C" int b200_interact_bwd(const void* d_T, const void* d_dR, void* d_dT, int64_t B, int64_t F,
                                 int64_t D, int64_t ldr, int dtype, b200_stream_t stream) {
  int rc = check_interact(d_T, d_dR, B, F, D, ldr, dtype);
  if (rc || B == 0) return rc;
  B200_REQUIRE(d_dT, B200_EINVAL, "interact_bwd: null output");
  cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
  if (interact_tc_supported(B, F, D, dtype, d_T, ldr)) return interact_bwd_tc(d_T, d_dR, d_dT, B, F, D, ldr, st);
  if (interact_f32_supported(B, F, D, dtype, d_T, d_dR, ldr) && interact_f32_supported(B, F, D, dtype, d_dT, d_dR, ldr))
    return interact_bwd_f32(d_T, d_dR, d_dT, B, F, D, ldr, st);
  return interact_bwd_simt(d_T, d_dR, d_dT, B, F, D, ldr, dtype, st);
}

extern "C" int b200_interact_fwd_ex(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr,
                                    int dtype, int flags, b200_stream_t stream) {
  B200_REQUIRE((flags & ~B200_INTERACT_ITSELF) == 0, B200_EINVAL, "interact: unknown flag");
  if (flags == 0) return b200_interact_fwd(d_T, d_R, B, F, D, ldr, dtype, stream);
  int rc = check_interact(d_T, d_R, B, F, D, ldr, dtype, 1);
  if (rc || B == 0) return rc;
  return interact_fwd_simt(d_T, d_R, B, F, D, ldr, dtype, reinterpret_cast<cudaStream_t>(stream), 1);
}

extern "C" int b200_interact_bwd_ex(const void* d_T, const void* d_dR, void* d_dT, int64_t B, int64_t F,
                                    int64_t D, int64_t ldr, int dtype, int flags, b200_stream_t stream) {
  B200_REQUIRE((flags & ~B200_INTERACT_ITSELF) == 0, B200_EINVAL, "interact: unknown flag");
  if (flags == 0) return b200_interact_bwd(d_T, d_dR, d_dT, B, F, D, ldr, dtype, stream);
  int rc = check_interact(d_T, d_dR, B, F, D, ldr, dtype, 1);
  if (rc || B == 0) return rc;
  B200_REQUIRE(d_dT, B200_EINVAL, "interact_bwd: null output");
  return interact_bwd_simt(d_T, d_dR, d_dT, B, F, D, ldr, dtype, reinterpret_cast<cudaStream_t>(stream), 1);
}
