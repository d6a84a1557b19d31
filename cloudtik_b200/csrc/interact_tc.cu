// Pairwise dot-product interaction on tcgen05 tensor cores (sm_100a), bf16 in / fp32 accumulate.
//
// Replaces torch.cat + torch.bmm + tril gather + cat of DLRM_Net.interact_features
// (reference P/dlrm_s_pytorch.py:309-334) and its autograd (two bmm + index_put).
//
// Shared structure of both kernels (persistent, one CTA per SM, warp-specialised):
//   warp 0      TMA producer: 3-D tensor map over T[B, F, D]; one box = (64 d, 32 f, 4 samples),
//               128-byte swizzle, rows f >= F and samples >= B zero-filled by the TMA unit.
//               A stage holds D/64 boxes = the whole [4 x 32, D] tile of four samples.
//   warp 1      MMA issuer (one elected lane): tcgen05.mma kind::f16, accumulators in TMEM.
//   warps 2..5  epilogue: tcgen05.ld the accumulator quarter their warp may touch, write out.
//   warps 6..9  (backward only) build the symmetric dZ tiles from dR in shared memory.
//   warps 10..13 (backward only) a second epilogue set: the two sets take alternate tiles, because the
//               backward epilogue (32 TMEM columns -> packed bf16 -> 32 row stores per sample) is the
//               slowest stage of that pipeline.
// Forward:  Z = A A^t with A = the 128-row tile (4 samples x 32 padded features), K = D; the four
//           32x32 diagonal blocks are the per-sample Gram matrices (the off-diagonal MMA work is
//           free: the kernel is HBM-bound).
// Backward: per sample dT^t[d, f] = sum_g T[g, d] * S[f, g]  (S = dZ + dZ^t): M = 128 d-values,
//           N = 32, K = 32, A = the same smem tile read MN-major, B = S (K-major, no swizzle).
#include <algorithm>

#include "tc_common.cuh"

namespace b200 {

constexpr int TC_NS = 4;              // samples per tile
constexpr int TC_FP = 32;             // padded features per sample
constexpr int TC_BOX_BYTES = TC_NS * TC_FP * 128;  // 16 KB: (64 bf16) x 32 x 4
constexpr int TC_EPI_WARPS = 4;
struct TcParams {
  const __nv_bfloat16* T;
  __nv_bfloat16* R;         // fwd out  [B, ldr]
  const __nv_bfloat16* dR;  // bwd in   [B, ldr]
  __nv_bfloat16* dT;        // bwd out  [B, F, D]
  int64_t B;
  int64_t ldr;
  int F, D;
  int n_tiles;
  int stages;
};

// ================================================================== forward
template <int KB>  // D / 64
__global__ void __launch_bounds__(192, 1)
    interact_fwd_tc_kernel(const __grid_constant__ CUtensorMap tmap, const TcParams p) {
  constexpr int D = KB * 64;
  constexpr int STAGE_BYTES = KB * TC_BOX_BYTES;
  constexpr int ACC_COLS = 128;
  constexpr int ACC_STAGES = 4;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stages = p.stages;
  uint8_t* tiles = smem;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * STAGE_BYTES);
  uint64_t* empty = full + stages;
  uint64_t* acc_full = empty + stages;
  uint64_t* acc_empty = acc_full + ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + ACC_STAGES);
  __nv_bfloat16* stg = reinterpret_cast<__nv_bfloat16*>(tmem_slot + 4);  // [4 warps][ldr_max]

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], 1);
      }
      for (int s = 0; s < ACC_STAGES; ++s) {
        mbar_init(&acc_full[s], 1);
        mbar_init(&acc_empty[s], TC_EPI_WARPS);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, ACC_COLS * ACC_STAGES);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---------------- TMA producer
    if (elect_one()) {
      int st = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        mbar_wait_bounded(&empty[st], ph ^ 1);
        mbar_expect_tx(&full[st], STAGE_BYTES);
        uint8_t* dst = tiles + (size_t)st * STAGE_BYTES;
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) tma_load_3d(dst + kb * TC_BOX_BYTES, &tmap, kb * 64, 0, tile * TC_NS, &full[st]);
        if (++st == stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer
    constexpr uint32_t idesc = make_idesc(128, 128, 0, 0);
    int st = 0, as = 0;
    uint32_t ph = 0, aph = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      mbar_wait_bounded(&acc_empty[as], aph ^ 1);
      mbar_wait_bounded(&full[st], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t base = smem_u32(tiles + (size_t)st * STAGE_BYTES);
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) {
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // UMMA_K = 16 bf16 = 32 bytes inside the 128-byte swizzle row
            const uint64_t d = smem_desc(base + kb * TC_BOX_BYTES + k * 32, 16, 1024, LAYOUT_SW128);
            umma_bf16(tmem_base + as * ACC_COLS, d, d, idesc, (kb | k) != 0);
          }
        }
        umma_commit(&empty[st]);     // smem slot free once these MMAs have read it
        umma_commit(&acc_full[as]);  // accumulator ready for the epilogue
      }
      __syncwarp();
      if (++st == stages) { st = 0; ph ^= 1; }
      if (++as == ACC_STAGES) { as = 0; aph ^= 1; }
    }
  } else {
    // ---------------- epilogue: warp q owns TMEM lanes [32q, 32q+32) = sample q of the tile
    const int q = warp & 3;
    const int F = p.F;
    const int npairs = F * (F - 1) / 2;
    const int W = D + npairs;
    __nv_bfloat16* my = stg + (size_t)(warp - 2) * ((p.ldr + 7) & ~7);
    const bool vec_ok = (p.ldr % 8) == 0 && (reinterpret_cast<uintptr_t>(p.R) & 15) == 0;
    int as = 0;
    uint32_t aph = 0;
    // x pass-through (feature 0) comes straight from global and is prefetched one tile ahead so its
    // latency hides behind the current tile's epilogue (D/8 <= 32 16-byte chunks: one per lane)
    auto load_x = [&](int tile) -> uint4 {
      const int64_t b = (int64_t)tile * TC_NS + q;
      if (tile < p.n_tiles && b < p.B && lane < D / 8)
        return __ldg(reinterpret_cast<const uint4*>(p.T + b * (int64_t)F * D) + lane);
      return make_uint4(0, 0, 0, 0);
    };
    uint4 xv = load_x(blockIdx.x);
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      const int64_t b = (int64_t)tile * TC_NS + q;
      const uint4 xn = load_x(tile + gridDim.x);
      mbar_wait_bounded(&acc_full[as], aph);
      tc_fence_after();
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + as * ACC_COLS + 32 * q, v);
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
      if (b < p.B) {
        if (lane < D / 8) reinterpret_cast<uint4*>(my)[lane] = xv;
        // lane i holds Z[i][0..31]; keep j < i
        if (lane < F) {
          const int rowoff = D + lane * (lane - 1) / 2;
#pragma unroll
          for (int j = 0; j < 31; ++j)
            if (j < lane) my[rowoff + j] = __float2bfloat16_rn(__uint_as_float(v[j]));
        }
        for (int c = W + lane; c < p.ldr; c += 32) my[c] = __float2bfloat16_rn(0.f);
        __syncwarp();
        __nv_bfloat16* out = p.R + b * p.ldr;
        if (vec_ok) {
          for (int c = lane; c < p.ldr / 8; c += 32) st_cs_v4(reinterpret_cast<uint4*>(out) + c, reinterpret_cast<uint4*>(my)[c]);
        } else {
          for (int c = lane; c < p.ldr; c += 32) out[c] = my[c];
        }
        __syncwarp();
      }
      xv = xn;
      if (++as == ACC_STAGES) { as = 0; aph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, ACC_COLS * ACC_STAGES);
}

// ================================================================== backward
// S tile of one sample: [32 f][32 g] bf16, K-major, no swizzle: core matrices of 8 rows x 16 bytes,
// offset(f, g) = (f/8)*512 + (g/8)*128 + (f%8)*16 + (g%8)*2   (LBO = 128 B along K, SBO = 512 B along N)
constexpr int S_BYTES = 2048;
__device__ __forceinline__ uint32_t s_off(int f, int g) {
  return (uint32_t)((f >> 3) * 512 + (g >> 3) * 128 + (f & 7) * 16 + (g & 7) * 2);
}

template <int KB>  // D / 64 (2 or 4)
__global__ void __launch_bounds__(448, 1)
    interact_bwd_tc_kernel(const __grid_constant__ CUtensorMap tmap, const TcParams p) {
  constexpr int D = KB * 64;
  constexpr int MB = D / 128;                    // 128-row accumulator blocks per sample
  constexpr int STAGE_BYTES = KB * TC_BOX_BYTES;
  constexpr int ACC_COLS = TC_NS * 32 * MB;      // TMEM columns per tile
  constexpr int ACC_STAGES = 512 / ACC_COLS >= 4 ? 4 : 2;
  constexpr int S_STAGE_BYTES = TC_NS * S_BYTES;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int stages = p.stages;
  uint8_t* tiles = smem;
  uint8_t* stiles = smem + (size_t)stages * STAGE_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(stiles + (size_t)stages * S_STAGE_BYTES);
  uint64_t* empty = full + stages;
  uint64_t* s_full = empty + stages;
  uint64_t* s_empty = s_full + stages;
  uint64_t* acc_full = s_empty + stages;
  uint64_t* acc_empty = acc_full + ACC_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + ACC_STAGES);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // zero the S tiles once: the diagonal and the padding rows / columns stay zero for ever
  for (int i = threadIdx.x; i < stages * S_STAGE_BYTES / 16; i += blockDim.x)
    reinterpret_cast<uint4*>(stiles)[i] = make_uint4(0, 0, 0, 0);
  if (warp == 0) {
    if (lane == 0) {
      for (int s = 0; s < stages; ++s) {
        mbar_init(&full[s], 1);
        mbar_init(&empty[s], 1);
        mbar_init(&s_full[s], 4);   // one arrive per S-builder warp
        mbar_init(&s_empty[s], 1);
      }
      for (int s = 0; s < ACC_STAGES; ++s) {
        mbar_init(&acc_full[s], 1);
        mbar_init(&acc_empty[s], TC_EPI_WARPS);
      }
      fence_mbar_init();
    }
    __syncwarp();
    tmem_alloc(tmem_slot, 512);
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int F = p.F;
  const int npairs = F * (F - 1) / 2;

  if (warp == 0) {
    // ---------------- TMA producer (same tiles as the forward)
    if (elect_one()) {
      int st = 0;
      uint32_t ph = 0;
      for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
        mbar_wait_bounded(&empty[st], ph ^ 1);
        mbar_expect_tx(&full[st], STAGE_BYTES);
        uint8_t* dst = tiles + (size_t)st * STAGE_BYTES;
#pragma unroll
        for (int kb = 0; kb < KB; ++kb) tma_load_3d(dst + kb * TC_BOX_BYTES, &tmap, kb * 64, 0, tile * TC_NS, &full[st]);
        if (++st == stages) { st = 0; ph ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer: per sample (and per 128-d block) D[128 d, 32 f] += A[d, g] B[f, g]
    constexpr uint32_t idesc = make_idesc(128, 32, /*a MN-major*/ 1, 0);
    int st = 0, as = 0;
    uint32_t ph = 0, aph = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      mbar_wait_bounded(&acc_empty[as], aph ^ 1);
      mbar_wait_bounded(&full[st], ph);
      mbar_wait_bounded(&s_full[st], ph);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t tbase = smem_u32(tiles + (size_t)st * STAGE_BYTES);
        const uint32_t sbase = smem_u32(stiles + (size_t)st * S_STAGE_BYTES);
#pragma unroll
        for (int s = 0; s < TC_NS; ++s) {
#pragma unroll
          for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
            for (int k = 0; k < 2; ++k) {  // K = 32 g-rows = 2 x UMMA_K(16)
              // A: element (d, g) at box(2mb + d/64) + s*4096 + g*128 + (d%64)*2 (128B-swizzled):
              //    MN-major, LBO = one box (next 64 d), SBO = 1024 B (next 8 g rows)
              const uint64_t ad = smem_desc(tbase + (2 * mb) * TC_BOX_BYTES + s * 4096 + k * 2048, TC_BOX_BYTES, 1024,
                                            LAYOUT_SW128);
              const uint64_t bd = smem_desc(sbase + s * S_BYTES + k * 256, 128, 512, LAYOUT_NONE);
              umma_bf16(tmem_base + as * ACC_COLS + (s * MB + mb) * 32, ad, bd, idesc, k != 0);
            }
          }
        }
        umma_commit(&empty[st]);
        umma_commit(&s_empty[st]);
        umma_commit(&acc_full[as]);
      }
      __syncwarp();
      if (++st == stages) { st = 0; ph ^= 1; }
      if (++as == ACC_STAGES) { as = 0; aph ^= 1; }
    }
  } else if (warp < 6 || warp >= 10) {
    // ---------------- epilogue: warp q reads TMEM lanes [32q, 32q+32) = d-values 32q + lane (+128 mb).
    // TWO sets of four warps (2..5 and 10..13) take alternate tiles: the epilogue (TMEM load, bf16 packing, 32
    // row stores per sample) was the slowest stage of the pipeline, not HBM
    const int q = warp & 3;
    const int eset = warp >= 10 ? 1 : 0;
    auto tile_of = [&](int i) { return (int)(blockIdx.x + (int64_t)i * gridDim.x); };
    // dR[:, :D] flows into feature 0; prefetched one tile ahead (TC_NS x MB values per thread)
    float pass_cur[TC_NS][MB], pass_nxt[TC_NS][MB];
    auto load_pass = [&](int tile, float (*dst)[MB]) {
#pragma unroll
      for (int s = 0; s < TC_NS; ++s) {
        const int64_t b = (int64_t)tile * TC_NS + s;
#pragma unroll
        for (int mb = 0; mb < MB; ++mb)
          dst[s][mb] = (tile < p.n_tiles && b < p.B) ? __bfloat162float(p.dR[b * p.ldr + mb * 128 + 32 * q + lane]) : 0.f;
      }
    };
    load_pass(tile_of(eset), pass_cur);
    for (int it = eset; tile_of(it) < p.n_tiles; it += 2) {
      const int tile = tile_of(it);
      const int as = it % ACC_STAGES;
      const uint32_t aph = (uint32_t)(it / ACC_STAGES) & 1u;
      load_pass(tile_of(it + 2), pass_nxt);
      mbar_wait_bounded(&acc_full[as], aph);
      tc_fence_after();
#pragma unroll
      for (int s = 0; s < TC_NS; ++s) {
        const int64_t b = (int64_t)tile * TC_NS + s;
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) {
          uint32_t v[32];
          tmem_ld32(tmem_base + ((uint32_t)(32 * q) << 16) + as * ACC_COLS + (s * MB + mb) * 32, v);
          if (b < p.B) {
            const int d = mb * 128 + 32 * q + lane;
            __nv_bfloat16* o = p.dT + b * (int64_t)F * D + d;
            // lanes 2k / 2k+1 hold adjacent d: swap so that each lane stores one packed bf16x2
            // (even lanes take the even features, odd lanes the odd ones)
            const bool odd = lane & 1;
#pragma unroll
            for (int f = 0; f < 32; f += 2) {
              float a0 = __uint_as_float(v[f]) + (f == 0 ? pass_cur[s][mb] : 0.f);
              float a1 = __uint_as_float(v[f + 1]);
              const float send = odd ? a0 : a1;
              const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
              const int fo = f + (odd ? 1 : 0);
              const float lo = odd ? recv : a0, hi = odd ? a1 : recv;
              if (fo < F) *reinterpret_cast<uint32_t*>(o - (odd ? 1 : 0) + (int64_t)fo * D) = pack_bf16x2(lo, hi);
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[as]);
#pragma unroll
      for (int s = 0; s < TC_NS; ++s)
#pragma unroll
        for (int mb = 0; mb < MB; ++mb) pass_cur[s][mb] = pass_nxt[s][mb];
    }
  } else {
    // ---------------- S builders (warps 6..9): S[i][j] = S[j][i] = dR[b, D + i(i-1)/2 + j]
    const int t = threadIdx.x - 192;  // 0..127
    constexpr int PPT = 4;            // pairs per thread: 4 * 128 >= 32 * 31 / 2
    uint32_t off_ij[PPT], off_ji[PPT];
#pragma unroll
    for (int c = 0; c < PPT; ++c) {
      const int pi = t + c * 128;
      int i = 1, j = 0;
      if (pi < npairs) tri_ij(pi, i, j);
      off_ij[c] = s_off(i, j);
      off_ji[c] = s_off(j, i);
    }
    const __nv_bfloat16 zero = __float2bfloat16_rn(0.f);
    __nv_bfloat16 cur[TC_NS][PPT], nxt[TC_NS][PPT];
    auto load_vals = [&](int tile, __nv_bfloat16 (*dst)[PPT]) {
#pragma unroll
      for (int s = 0; s < TC_NS; ++s) {
        const int64_t b = (int64_t)tile * TC_NS + s;
        const bool live = tile < p.n_tiles && b < p.B;
#pragma unroll
        for (int c = 0; c < PPT; ++c) {
          const int pi = t + c * 128;
          dst[s][c] = (live && pi < npairs) ? p.dR[b * p.ldr + D + pi] : zero;
        }
      }
    };
    load_vals(blockIdx.x, cur);
    int st = 0;
    uint32_t ph = 0;
    for (int tile = blockIdx.x; tile < p.n_tiles; tile += gridDim.x) {
      load_vals(tile + gridDim.x, nxt);  // in flight while this tile's S is written
      mbar_wait_bounded(&s_empty[st], ph ^ 1);
      uint8_t* sb = stiles + (size_t)st * S_STAGE_BYTES;
#pragma unroll
      for (int s = 0; s < TC_NS; ++s) {
#pragma unroll
        for (int c = 0; c < PPT; ++c) {
          if (t + c * 128 < npairs) {
            *reinterpret_cast<__nv_bfloat16*>(sb + s * S_BYTES + off_ij[c]) = cur[s][c];
            *reinterpret_cast<__nv_bfloat16*>(sb + s * S_BYTES + off_ji[c]) = cur[s][c];
          }
        }
      }
      fence_proxy_async();  // generic-proxy smem writes -> visible to the tensor-core (async) proxy
      __syncwarp();
      if (lane == 0) mbar_arrive(&s_full[st]);
#pragma unroll
      for (int s = 0; s < TC_NS; ++s)
#pragma unroll
        for (int c = 0; c < PPT; ++c) cur[s][c] = nxt[s][c];
      if (++st == stages) { st = 0; ph ^= 1; }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, 512);
}

// ================================================================== host side
EncodeTiledFn tc_get_encode() {
  static EncodeTiledFn fn = nullptr;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(ptr);
  }
  return fn;
}

static int make_tmap(CUtensorMap* map, const void* T, int64_t B, int64_t F, int64_t D) {
  EncodeTiledFn enc = tc_get_encode();
  B200_REQUIRE(enc != nullptr, B200_ELAUNCH, "interact_tc: cuTensorMapEncodeTiled entry point not available");
  cuuint64_t dims[3] = {(cuuint64_t)D, (cuuint64_t)F, (cuuint64_t)B};
  cuuint64_t strides[2] = {(cuuint64_t)D * 2, (cuuint64_t)F * D * 2};
  cuuint32_t box[3] = {64, TC_FP, TC_NS};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(T), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  B200_REQUIRE(r == CUDA_SUCCESS, B200_ELAUNCH, "interact_tc: cuTensorMapEncodeTiled failed (%d)", (int)r);
  return B200_OK;
}

int tc_sm_count() {
  static int n = 0;
  if (!n) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

bool interact_tc_supported(int64_t B, int64_t F, int64_t D, int dtype, const void* T, int64_t ldr) {
  return dtype == B200_BF16 && F >= 2 && F <= TC_FP && (D == 128 || D == 256) && B >= 1 &&
         B < ((int64_t)1 << 31) - 8 && (reinterpret_cast<uintptr_t>(T) & 15) == 0 && ldr <= 2048;
}

int interact_fwd_tc(const void* T, void* R, int64_t B, int64_t F, int64_t D, int64_t ldr, cudaStream_t st) {
  CUtensorMap map;
  int rc = make_tmap(&map, T, B, F, D);
  if (rc) return rc;
  TcParams p{};
  p.T = static_cast<const __nv_bfloat16*>(T);
  p.R = static_cast<__nv_bfloat16*>(R);
  p.B = B;
  p.ldr = ldr;
  p.F = (int)F;
  p.D = (int)D;
  p.n_tiles = (int)ceil_div(B, TC_NS);
  const int KB = (int)(D / 64);
  const size_t stage = (size_t)KB * TC_BOX_BYTES;
  p.stages = (int)std::min<size_t>(4, (200 * 1024) / stage);
  const size_t smem = 1024 + p.stages * stage + (2 * p.stages + 8) * 8 + 16 + 4 * (((size_t)ldr + 7) & ~7) * 2;
  const int grid = std::min(p.n_tiles, tc_sm_count());
  if (KB == 2) {
    cudaFuncSetAttribute(interact_fwd_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_fwd_tc_kernel<2><<<grid, 192, smem, st>>>(map, p);
  } else {
    cudaFuncSetAttribute(interact_fwd_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_fwd_tc_kernel<4><<<grid, 192, smem, st>>>(map, p);
  }
  B200_CHECK_LAUNCH("interact_fwd_tc");
  return B200_OK;
}

int interact_bwd_tc(const void* T, const void* dR, void* dT, int64_t B, int64_t F, int64_t D, int64_t ldr,
                    cudaStream_t st) {
  CUtensorMap map;
  int rc = make_tmap(&map, T, B, F, D);
  if (rc) return rc;
  TcParams p{};
  p.T = static_cast<const __nv_bfloat16*>(T);
  p.dR = static_cast<const __nv_bfloat16*>(dR);
  p.dT = static_cast<__nv_bfloat16*>(dT);
  p.B = B;
  p.ldr = ldr;
  p.F = (int)F;
  p.D = (int)D;
  p.n_tiles = (int)ceil_div(B, TC_NS);
  const int KB = (int)(D / 64);
  const size_t stage = (size_t)KB * TC_BOX_BYTES + TC_NS * S_BYTES;
  p.stages = (int)std::min<size_t>(4, (200 * 1024) / stage);
  const size_t smem = 1024 + p.stages * stage + (4 * p.stages + 8) * 8 + 16;
  const int grid = std::min(p.n_tiles, tc_sm_count());
  if (KB == 2) {
    cudaFuncSetAttribute(interact_bwd_tc_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_bwd_tc_kernel<2><<<grid, 448, smem, st>>>(map, p);
  } else {
    cudaFuncSetAttribute(interact_bwd_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    interact_bwd_tc_kernel<4><<<grid, 448, smem, st>>>(map, p);
  }
  B200_CHECK_LAUNCH("interact_bwd_tc");
  return B200_OK;
}

}  // namespace b200
