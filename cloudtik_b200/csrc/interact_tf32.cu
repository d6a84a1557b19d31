// Pairwise dot-product interaction in fp32 on tcgen05 tensor cores: every operand is split into two TF32 numbers
// and three products are accumulated (3xTF32), which keeps the 1e-5 parity bound of the reference's default
// precision (reference P/dlrm_s_pytorch.py:315-333 runs fp32 unless --bf16; a single TF32 product is good to 1e-3).
//
//   x = hi + lo,  hi = tf32(x) (round to nearest),  lo = tf32(x - hi),   x*y ~ hi*hi' + hi*lo' + lo*hi'
//   (the dropped lo*lo' and the rounding of lo are ~2^-22 |xy| each, unbiased)
//
// Forward, same structure as interact_tc.cu (persistent, one CTA per SM, warp-specialised):
//   warp 0      TMA producer: 3-D tensor map over fp32 T[B, F, D]; one box = (32 d, 32 f, 4 samples) = 16 KB,
//               128-byte swizzle, rows f >= F and samples >= B zero-filled.  One pipeline stage = one box.
//   warps 6..9  converters: overwrite the landed box with hi in place and write lo next to it (same swizzled layout,
//               the split is element-wise), fence.proxy.async, arrive.
//   warp 1      MMA issuer: per stage 4 k-steps (K = 8 floats = 32 B) x 3 products, tcgen05.mma kind::tf32,
//               M = N = 128 (4 samples x 32 padded features; the four 32x32 diagonal blocks are the Gram matrices),
//               fp32 accumulators in TMEM, D / 32 stages per tile.
//   warps 2..5  epilogue: tcgen05.ld of the diagonal block, lower-triangle gather + x pass-through staged in shared
//               memory, coalesced row store.
// Not the default fp32 path: see interact_tf32_supported() for the measured reason (shared-memory operand bandwidth).
#include <stdlib.h>

#include <algorithm>

#include "tc_common.cuh"

namespace b200 {

constexpr int TF_NS = 4;                             // samples per tile
constexpr int TF_FP = 32;                            // padded features per sample
constexpr int TF_BOX_BYTES = TF_NS * TF_FP * 128;    // 16 KB: (32 fp32) x 32 x 4
constexpr int TF_STAGE_BYTES = 2 * TF_BOX_BYTES;     // hi (in place of the raw box) + lo
constexpr int TF_STAGES = 6;
constexpr int TF_ACC_COLS = 128;
constexpr int TF_ACC_STAGES = 4;
constexpr int TF_EPI_WARPS = 4;
constexpr int TF_CONV_WARPS = 4;

// x = hi + lo with hi = x rounded to nearest TF32 and lo = (x - hi) rounded to nearest TF32 (x - hi is exact in
// fp32).  Both halves are exact TF32 numbers, so the tensor core uses them as they are; rounding (not truncating)
// keeps the errors of the dropped bits unbiased -- with truncation they add up linearly over the D terms of a dot
// product (measured: mean error 9e-6 at D = 128 against 2e-6 with rounding).
__device__ __forceinline__ void split_tf32(uint32_t x, uint32_t& hi, uint32_t& lo) {
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(__uint_as_float(x)));
  const float r = __uint_as_float(x) - __uint_as_float(hi);
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
}

struct TfParams {
  const float* T;
  float* R;
  int64_t B;
  int64_t ldr;
  int F, D;
  int n_tiles;
};

template <int KB>  // D / 32
__global__ void __launch_bounds__(320, 1)
    interact_fwd_tf32_kernel(const __grid_constant__ CUtensorMap tmap, const TfParams p) {
  constexpr int D = KB * 32;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* tiles = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)TF_STAGES * TF_STAGE_BYTES);
  uint64_t* conv = full + TF_STAGES;
  uint64_t* empty = conv + TF_STAGES;
  uint64_t* acc_full = empty + TF_STAGES;
  uint64_t* acc_empty = acc_full + TF_ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + TF_ACC_STAGES);
  float* stg = reinterpret_cast<float*>(tmem_slot + 4);  // [4 warps][ldr rounded up to 4]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    if (lane == 0) {
      for (int s = 0; s < TF_STAGES; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&conv[s], TF_CONV_WARPS);
        mbar_init(&empty[s], 1);
      }
      for (int s = 0; s < TF_ACC_STAGES; ++s) {
        mbar_init(&acc_full[s], 1);
        mbar_init(&acc_empty[s], TF_EPI_WARPS);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, TF_ACC_COLS * TF_ACC_STAGES);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------- TMA producer: one box (32 columns of d) per stage
    if (elect_one()) {
      int st = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
          mbar_wait_bounded(&empty[st], ph ^ 1);
          mbar_expect_tx(&full[st], TF_BOX_BYTES);
          tma_load_3d(tiles + (size_t)st * TF_STAGE_BYTES, &tmap, kb * 32, 0, tile * TF_NS, &full[st]);
          if (++st == TF_STAGES) { st = 0; ph ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer
    constexpr uint32_t idesc = make_idesc(128, 128, 0, 0, /*TF32*/ 2);
    int st = 0, as = 0;
    uint32_t ph = 0, aph = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      mbar_wait_bounded(&acc_empty[as], aph ^ 1);
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait_bounded(&conv[st], ph);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t hi = smem_u32(tiles + (size_t)st * TF_STAGE_BYTES);
          const uint32_t lo = hi + TF_BOX_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // UMMA_K = 8 fp32 = 32 bytes inside the 128-byte swizzle row
            const uint64_t dh = smem_desc(hi + k * 32, 16, 1024, LAYOUT_SW128);
            const uint64_t dl = smem_desc(lo + k * 32, 16, 1024, LAYOUT_SW128);
            umma_tf32(tmem_base + as * TF_ACC_COLS, dh, dh, idesc, (kb | k) != 0);
            umma_tf32(tmem_base + as * TF_ACC_COLS, dh, dl, idesc, 1);
            umma_tf32(tmem_base + as * TF_ACC_COLS, dl, dh, idesc, 1);
          }
          umma_commit(&empty[st]);  // stage free once these MMAs have read it
          if (kb == KB - 1) umma_commit(&acc_full[as]);
        }
        __syncwarp();
        if (++st == TF_STAGES) { st = 0; ph ^= 1; }
      }
      if (++as == TF_ACC_STAGES) { as = 0; aph ^= 1; }
    }
  } else if (warp < 6) {
    // ---------------- epilogue: warp q owns TMEM lanes [32q, 32q+32) = sample q of the tile
    const int q = warp & 3;
    const int F = p.F;
    const int npairs = F * (F - 1) / 2;
    const int W = D + npairs;
    float* my = stg + (size_t)(warp - 2) * ((p.ldr + 3) & ~3);
    const bool vec_ok = (p.ldr % 4) == 0 && (reinterpret_cast<uintptr_t>(p.R) & 15) == 0;
    int as = 0;
    uint32_t aph = 0;
    // x pass-through (feature 0) straight from global, prefetched one tile ahead (D/4 <= 32 16-byte chunks)
    auto load_x = [&](int tile) -> uint4 {
      const int64_t b = (int64_t)tile * TF_NS + q;
      if (tile < p.n_tiles && b < p.B && lane < D / 4)
        return __ldg(reinterpret_cast<const uint4*>(p.T + b * (int64_t)F * D) + lane);
      return make_uint4(0, 0, 0, 0);
    };
    uint4 xv = load_x(blockIdx.x);
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t b = (int64_t)tile * TF_NS + q;
      const uint4 xn = load_x(tile + gridDim.x);
      mbar_wait_bounded(&acc_full[as], aph);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + as * TF_ACC_COLS + 32 * q, v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      if (b < p.B) {
        if (lane < D / 4) reinterpret_cast<uint4*>(my)[lane] = xv;
        if (lane < F) {  // lane i holds Z[i][0..31]; keep j < i
          const int rowoff = D + lane * (lane - 1) / 2;
#pragma unroll
          for (int j = 0; j < 31; ++j)
            if (j < lane) my[rowoff + j] = __uint_as_float(v[j]);
        }
        for (int c = W + lane; c < p.ldr; c += 32) my[c] = 0.f;
        __syncwarp();
        float* out = p.R + b * p.ldr;
        if (vec_ok) {
          for (int c = lane; c < p.ldr / 4; c += 32) st_cs_v4(reinterpret_cast<uint4*>(out) + c, reinterpret_cast<uint4*>(my)[c]);
        } else {
          for (int c = lane; c < p.ldr; c += 32) out[c] = my[c];
        }
        __syncwarp();
      }
      xv = xn;
      if (++as == TF_ACC_STAGES) { as = 0; aph ^= 1; }
    }
  } else {
    // ---------------- converters (warps 6..9): raw fp32 box -> hi in place, lo behind it
    const int t = threadIdx.x - 192;  // 0..127
    int st = 0;
    uint32_t ph = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
#pragma unroll
      for (int kb = 0; kb < KB; ++kb) {
        mbar_wait_bounded(&full[st], ph);
        uint4* hi = reinterpret_cast<uint4*>(tiles + (size_t)st * TF_STAGE_BYTES);
        uint4* lo = reinterpret_cast<uint4*>(tiles + (size_t)st * TF_STAGE_BYTES + TF_BOX_BYTES);
#pragma unroll
        for (int i = 0; i < TF_BOX_BYTES / 16 / 128; ++i) {
          const int o = t + i * 128;
          const uint4 x = hi[o];
          uint4 h, l;
          split_tf32(x.x, h.x, l.x);
          split_tf32(x.y, h.y, l.y);
          split_tf32(x.z, h.z, l.z);
          split_tf32(x.w, h.w, l.w);
          hi[o] = h;
          lo[o] = l;
        }
        fence_proxy_async();  // generic-proxy smem writes -> visible to the tensor-core (async) proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(&conv[st]);
        if (++st == TF_STAGES) { st = 0; ph ^= 1; }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, TF_ACC_COLS * TF_ACC_STAGES);
}

// ================================================================== host side
static int make_tmap_f32(CUtensorMap* map, const void* T, int64_t B, int64_t F, int64_t D) {
  EncodeTiledFn enc = tc_get_encode();
  B200_REQUIRE(enc != nullptr, B200_ELAUNCH, "interact_tf32: cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)F, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)D * 4, (cuuint64_t)F * D * 4};
  cuuint32_t box[3] = {32, TF_FP, TF_NS};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<void*>(T), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ELAUNCH, "interact_tf32: cuTensorMapEncodeTiled failed (%d)", (int)r);
  return B200_OK;
}

bool interact_tf32_shape_ok(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr);

bool interact_tf32_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr) {
  // Measured on B200, B = 32768, F = 27, D = 128 (tools/tf32_ab.sh): 0.187 ms here against 0.146 ms for the
  // register-tiled CUDA-core kernel (interact_f32.cu).  The tensor pipe is not the limit, shared memory is: with
  // both operands in shared memory a K = 8 TF32 step reads (128 + 128) rows x 32 B = 8 KB per 64-cycle dispatch,
  // i.e. the full 128 B/clk of the SM, for 48 dispatches per tile, and the hi / lo conversion pass and the TMA
  // writes share that port (640 KB per 4-sample tile ~ 2.6 us, against 1.4 us of HBM time).  So fp32 dispatches
  // here only on request (B200_INTERACT_TF32=1, or b200_interact_fwd_tf32 directly); bf16 has no such problem
  // (8 dispatches per tile).
  static const bool on = []() {
    const char* e = getenv("B200_INTERACT_TF32");
    return e != nullptr && e[0] == '1';
  }();
  return on && interact_tf32_shape_ok(B, F, D, dtype, T, ldr);
}

bool interact_tf32_shape_ok(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr) {
  return dtype == B200_F32 && F >= 2 && F <= TF_FP && (D == 32 || D == 64 || D == 128) && B >= 1 &&
         B < ((int64_t)1 << 31) - 8 && (reinterpret_cast<uintptr_t>(T) & 15) == 0 && ldr <= 2048;
}

int interact_fwd_tf32(const void* T, void* R, int64_t B, int64_t F, int64_t D, int64_t ldr, cudaStream_t st) {
  CUtensorMap map;
  int rc = make_tmap_f32(&map, T, B, F, D);
  if (rc) return rc;
  TfParams p{};
  p.T = static_cast<const float*>(T);
  p.R = static_cast<float*>(R);
  p.B = B;
  p.ldr = ldr;
  p.F = (int)F;
  p.D = (int)D;
  p.n_tiles = (int)ceil_div(B, TF_NS);
  const size_t smem = 1024 + (size_t)TF_STAGES * TF_STAGE_BYTES + (3 * TF_STAGES + 2 * TF_ACC_STAGES) * 8 + 16 +
                      TF_EPI_WARPS * (((size_t)ldr + 3) & ~3) * 4;
  const int grid = std::min(p.n_tiles, tc_sm_count());
#define B200_TF32_FWD(KB)                                                                                       \
  {                                                                                                             \
    cudaFuncSetAttribute(interact_fwd_tf32_kernel<KB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    interact_fwd_tf32_kernel<KB><<<grid, 320, smem, st>>>(map, p);                                              \
  }
  if (D == 128) B200_TF32_FWD(4)
  else if (D == 64) B200_TF32_FWD(2)
  else B200_TF32_FWD(1)
#undef B200_TF32_FWD
  B200_CHECK_LAUNCH("interact_fwd_tf32");
  return B200_OK;
}

}  // namespace b200

extern "C" int b200_interact_fwd_tf32(const void* d_T, void* d_R, int64_t B, int64_t F, int64_t D, int64_t ldr,
                                      b200_stream_t stream) {
  B200_REQUIRE(B >= 0 && F >= 2 && D >= 1 && ldr >= D + F * (F - 1) / 2, B200_EINVAL, "interact_fwd_tf32: bad shape");
  if (B == 0) return B200_OK;
  B200_REQUIRE(d_T && d_R, B200_EINVAL, "interact_fwd_tf32: null pointer");
  B200_REQUIRE(b200::interact_tf32_shape_ok(B, F, D, B200_F32, d_T, ldr), B200_EUNSUPPORTED,
               "interact_fwd_tf32: needs F <= 32, D in {32, 64, 128}, 16-byte aligned T");
  return b200::interact_fwd_tf32(d_T, d_R, B, F, D, ldr, reinterpret_cast<cudaStream_t>(stream));
}
