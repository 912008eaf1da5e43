// conv2 (5x5, SAME, 32 -> 64 on 14x14 maps) on tcgen05: forward, data gradient and filter gradient as
// implicit GEMMs.  Replaces tf.nn.conv2d + bias_add + relu + max_pool of /root/reference/mnist_python_m.py:
// 93-101,114-116 and the conv gradients TF's autodiff generates at m:222.
//
// im2col never touches memory: the A tiles are 4-D TMA boxes of the NHWC activation shifted by the filter tap;
// TMA zero-fills everything outside the image, which is exactly SAME padding.  Boxes are 16 wide on a 14-wide
// map, so 2 of 16 accumulator rows per image row are padding (never stored); in exchange a 2x2 pooling window
// lives inside one warp's 32 TMEM lanes (lane = hh*16 + ww) and the pool is two shuffles.  The whole bf16 filter (100 KB) stays resident in shared
// memory; CTAs are persistent over pixel tiles with double-buffered TMEM accumulators.
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"
#include <stdlib.h>

namespace {

using namespace tfd;
typedef __nv_bfloat16 bf16;

constexpr int kThreads = 192;
constexpr int kFilterBytes = 25 * 4096;  // [800][64] bf16, or one 32-channel half of the fp32 filter

__device__ __forceinline__ void load_filter(uint8_t* filt, const CUtensorMap* tmW, uint64_t* bar, int c0) {
  mbar_arrive_expect_tx(bar, kFilterBytes);
#pragma unroll
  for (int i = 0; i < 4; ++i) tma_load_2d(filt + i * 25600, tmW, bar, c0, i * 200);
}

// ------------------------------------------------------------------------------------------------
// forward / dgrad share one kernel shape.  kFwd: A = in [B,14,14,32] (64 B rows, SW64), N = 64, fused
// bias+ReLU+pool epilogue.  !kFwd: A = dy [B,14,14,64] (128 B rows, SW128), N = 32, fp32 store epilogue.
//
// Tile = 8 rows x 16 columns of ONE image (two tiles per image; rows >= 14 and columns >= 14 are padding).
// Per tile only FIVE TMA boxes are loaded, one per filter column kw: {C, 16 w, 12 h, 1 n} starting at
// (w = kw-2, h = 8*tile-2) -- 8 output rows plus the 4-row halo.  The A operand of tap (kh,kw) is then just a
// descriptor into that box at row offset kh*16 (a multiple of the swizzle atom), so the five row taps of a
// column reuse one box: 5x less L2->SM traffic than one box per tap, which is what bounded this kernel.
// ------------------------------------------------------------------------------------------------
// tf32 arm (TF = 1): activations and filter are fp32 in memory and are read by the tensor core as tf32.  The fp32 filter
// is 200 KB, twice what fits next to the input ring, so a CTA works on ONE HALF of the 64 output channels and keeps
// that half of the filter (25 taps x [32 ci][32 co] fp32 = 100 KB, the same bytes as the whole bf16 filter) resident:
//   forward: the half is an N split -- work item = (pixel tile, co half), N = 32, epilogue writes 32 of the 64 pooled
//            channels; the filter tile of a tap is the MN-major B operand ([32 ci rows][32 co], 32 B swizzle atoms);
//   dgrad:   the half is a K split -- work item = (pixel tile, co half), A = that half of dy's channels, the same
//            filter tile read K-major; the two halves write two partial planes of dx that the consumer
//            (conv1_wgrad) adds, in a fixed order.
// Even CTAs take half 0, odd CTAs half 1, so a CTA loads its filter half once.
template <bool kFwd, int TF>
struct ConvCfg {
  static constexpr int kRowBytes = TF ? 128 : (kFwd ? 64 : 128);
  static constexpr int kBoxBytes = 12 * 16 * kRowBytes;  // 12 KB / 24 KB
  static constexpr int kStages = (kRowBytes == 64) ? 8 : 5;  // ring of kw boxes
  static constexpr int kN = TF ? 32 : (kFwd ? 64 : 32);
  static constexpr int kKSteps = TF ? 4 : (kFwd ? 2 : 4);  // MMAs per tap (UMMA K = 32 B)
  static constexpr int kSmem = kFilterBytes + kStages * kBoxBytes + 1024 + 512;
};

template <bool kFwd, int TF>
__global__ void __launch_bounds__(kThreads, 1)
conv2_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                const float* __restrict__ bias, void* __restrict__ pool_out_v, unsigned char* __restrict__ amax,
                float* __restrict__ dx, int B, int num_tiles, long long dx_plane_stride) {
  using C = ConvCfg<kFwd, TF>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* filt = smem;
  uint8_t* abuf = smem + kFilterBytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(abuf + C::kStages * C::kBoxBytes);
  uint64_t* a_empty = a_full + C::kStages;
  uint64_t* filt_full = a_empty + C::kStages;
  uint64_t* acc_full = filt_full + 1;   // [2]
  uint64_t* acc_empty = acc_full + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // work distribution: bf16 -- CTA b takes tiles b, b+grid, ...; tf32 -- CTA b takes channel half b&1 of tiles b>>1, ...
  const int half = TF ? (int)(blockIdx.x & 1) : 0;
  const int tile0 = TF ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tstride = TF ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    mbar_init(filt_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * C::kN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_wait();
  grid_dep_launch();

  // Producer and MMA warps run their loops with ALL lanes (warp-uniform control flow and addresses, so the shared-memory
  // descriptors live in uniform registers); only the TMA / MMA / commit instructions themselves sit under elect_one().
  // With the loops under `if (lane == 0)` the compiler had to move five registers to the uniform file around every
  // MMA (21 instructions per UTCHMMA), which -- not the tensor pipe -- bounded the kernel.
  if (warp == 0) {
    // the first input boxes go out before the 100 KB filter so the first MMAs are not queued behind it
    int stage = 0;
    uint32_t phase = 0;
    bool filter_sent = false;
    for (int tile = tile0; tile < num_tiles; tile += tstride) {
      const int n = tile >> 1, ht = tile & 1;
      for (int kw = 0; kw < 5; ++kw) {
        mbar_wait(&a_empty[stage], phase ^ 1);
        if (elect_one()) {
          mbar_arrive_expect_tx(&a_full[stage], C::kBoxBytes);
          // tf32 dgrad reads only this CTA's half of dy's channels (the K split)
          tma_load_4d(abuf + stage * C::kBoxBytes, &tmA, &a_full[stage], (TF && !kFwd) ? half * 32 : 0,
                      kFwd ? kw - 2 : 2 - kw, 8 * ht - 2, n);
          if (!filter_sent) load_filter(filt, &tmW, filt_full, TF ? half * 32 : 0);
        }
        __syncwarp();
        filter_sent = true;
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = TF ? make_idesc_tf32(128, C::kN, 0, kFwd ? 1 : 0) : make_idesc_bf16(128, C::kN, 0, kFwd ? 1 : 0);
    // descriptor templates: everything but the start address is fixed per operand
    const uint64_t da0 = (TF || !kFwd) ? make_smem_desc(smem_u32(abuf), 16, 1024, kSwz128)
                                       : make_smem_desc(smem_u32(abuf), 16, 512, kSwz64);
    const uint64_t db0 = TF ? (kFwd ? make_smem_desc(smem_u32(filt), 4096, 512, kSwz128B32)
                                    : make_smem_desc(smem_u32(filt), 16, 1024, kSwz128))
                            : (kFwd ? make_smem_desc(smem_u32(filt), 4096, 1024, kSwz128)
                                    : make_smem_desc(smem_u32(filt), 16, 1024, kSwz128));
    // bytes between consecutive MMAs' K slices inside a filter tap tile: MN-major B steps by k-rows, K-major by 32 B
    constexpr uint32_t kBStep = kFwd ? (TF ? 1024u : 2048u) : 32u;
    mbar_wait(filt_full, 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = tile0; tile < num_tiles; tile += tstride, ++it) {
      const int buf = it & 1;
      mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + buf * C::kN;
      for (int kw = 0; kw < 5; ++kw) {
        mbar_wait(&a_full[stage], phase);
        tc_fence_after();
        if (elect_one()) {
          const uint64_t da_s = desc_advance(da0, (uint32_t)(stage * C::kBoxBytes));
          const uint64_t db_s = desc_advance(db0, (uint32_t)(kw * 4096));
#pragma unroll
          for (int kh = 0; kh < 5; ++kh) {
            // forward: output row hh reads box row hh+kh.  dgrad: dy row hh+2-kh -> box row hh+4-kh.
            // A: [128 px][K slice] K-major (32 B per MMA inside the row).  B: filter tap (kh,kw) at tap*4096.
#pragma unroll
            for (int ks = 0; ks < C::kKSteps; ++ks) {
              const uint64_t da = desc_advance(da_s, (uint32_t)((kFwd ? kh : 4 - kh) * 16 * C::kRowBytes + ks * 32));
              const uint64_t db = desc_advance(db_s, (uint32_t)(kh * 5 * 4096) + ks * kBStep);
              const uint32_t acc = (kw > 0 || kh > 0 || ks > 0) ? 1u : 0u;
              if (TF) umma_tf32(tacc, da, db, idesc, acc);
              else    umma_bf16(tacc, da, db, idesc, acc);
            }
          }
          umma_commit(&a_empty[stage]);
          if (kw == 4) umma_commit(&acc_full[buf]);
        }
        __syncwarp();
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    const int q = warp & 3;  // TMEM lane quadrant = two image rows of the tile
    const int hh = lane >> 4, ww = lane & 15;
    int it = 0;
    for (int tile = tile0; tile < num_tiles; tile += tstride, ++it) {
      const int buf = it & 1;
      const int n = tile >> 1, ht = tile & 1;
      const int h = 8 * ht + 2 * q + hh;  // conv output row of this lane
      mbar_wait(&acc_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + buf * C::kN + ((uint32_t)(q * 32) << 16);
      if (kFwd) {
        const int self = hh * 2 + (ww & 1);
        const int ph = 4 * ht + q;  // pooled row
        const int cbase = half * 32;  // first output channel of this CTA's accumulator columns (tf32: its half)
#pragma unroll 1
        for (int j = 0; j < C::kN / 16; ++j) {
          float v[16];
          tmem_ld16(tacc + j * 16, v);
          uint32_t am[4] = {0, 0, 0, 0};
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = fmaxf(v[i] + __ldg(bias + cbase + j * 16 + i), 0.f);
            int xi = self;
            float y = __shfl_xor_sync(0xffffffffu, x, 1);
            int yi = self ^ 1;
            if (y > x || (y == x && yi < xi)) { x = y; xi = yi; }
            y = __shfl_xor_sync(0xffffffffu, x, 16);
            yi = __shfl_xor_sync(0xffffffffu, xi, 16);
            if (y > x || (y == x && yi < xi)) { x = y; xi = yi; }
            v[i] = x;
            am[i >> 2] |= (uint32_t)xi << ((i & 3) * 8);
          }
          // the four lanes of a pooling window hold the same result: lane `self` == j stores chunk j
          if (j == self && ww < 14 && ph < 7 && n < B) {
            const long long o = ((((long long)n * 7 + ph) * 7) + (ww >> 1)) * 64 + cbase + j * 16;
            if (TF) {
              float* po = static_cast<float*>(pool_out_v) + o;
#pragma unroll
              for (int i = 0; i < 16; i += 4)
                *reinterpret_cast<float4*>(po + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            } else {
              bf16* po = static_cast<bf16*>(pool_out_v) + o;
              uint32_t packed[8];
#pragma unroll
              for (int i = 0; i < 8; ++i) packed[i] = pack_bf16x2(v[2 * i], v[2 * i + 1]);
              *reinterpret_cast<uint4*>(po) = make_uint4(packed[0], packed[1], packed[2], packed[3]);
              *reinterpret_cast<uint4*>(po + 8) = make_uint4(packed[4], packed[5], packed[6], packed[7]);
            }
            *reinterpret_cast<uint4*>(amax + o) = make_uint4(am[0], am[1], am[2], am[3]);
          }
        }
      } else {
        const bool ok = (ww < 14) && (h < 14) && (n < B);
        float* o = dx + (long long)half * dx_plane_stride + ((((long long)n * 14 + h) * 14) + ww) * 32;
#pragma unroll 1
        for (int j = 0; j < 2; ++j) {
          float v[16];
          tmem_ld16(tacc + j * 16, v);
          if (ok) {
#pragma unroll
            for (int i = 0; i < 16; i += 4)
              *reinterpret_cast<float4*>(o + j * 16 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * C::kN);
  }
}

// ------------------------------------------------------------------------------------------------
// One-box variant of the forward / dgrad kernel for operands with 128 B rows (tf32 forward and dgrad, bf16 dgrad).
//
// The kernel above fetches five boxes per tile (one per filter column), 120 KB of L2 -> SM traffic for 128 output
// pixels, and that traffic is what bounds it (~5.5 TB/s aggregate).  Here a tile is 8 columns x 16 rows of one image
// (accumulator row = h*8 + w) and ONE box {C, 16 w, 20 h} = 40 KB holds the inputs of all 25 taps: the A operand of
// tap (kh,kw) is the same box read from a start address moved by (kh*16 + kw) rows.  The 8 rows of a core-matrix group
// are 8 consecutive columns (1024 B contiguous), groups follow at the box's row pitch (SBO = 16 * 128 B).  The start
// is moved by kw rows, which is not a multiple of the 8-row swizzle pattern; measured on B200: the tensor core derives
// the swizzle phase from the address bits (as TMA did when it wrote the box), so the descriptor needs NO base offset --
// with (address >> 7) & 7 in the base-offset field the results are wrong, with 0 they match the oracle.  3x less
// L2 -> SM traffic per tile (15.7 vs 17.2 us forward, 18.3 vs 19.3 us dgrad on the tf32 arm: the kernels are no longer
// traffic-bound after it); the pooling window is still inside one warp (partners at lane ^ 1 and lane ^ 8).
// ------------------------------------------------------------------------------------------------
template <bool kFwd, int TF>
struct BoxCfg {
  static constexpr int kBoxBytes = 20 * 16 * 128;  // 40 KB
  static constexpr int kStages = 3;
  static constexpr int kN = TF ? 32 : 32;           // tf32: a 32-channel half; bf16 dgrad: the 32 input channels
  static constexpr int kKSteps = 4;                 // 128 B of K per tap, 32 B per MMA
  static constexpr int kSmem = kFilterBytes + kStages * kBoxBytes + 1024 + 512;
};

template <bool kFwd, int TF>
__global__ void __launch_bounds__(kThreads, 1)
conv2_box_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmW,
                 const float* __restrict__ bias, void* __restrict__ pool_out_v, unsigned char* __restrict__ amax,
                 float* __restrict__ dx, int B, int num_tiles, long long dx_plane_stride) {
  static_assert(TF == 1 || !kFwd, "the bf16 forward operand has 64 B rows: it stays on the five-box kernel");
  using C = BoxCfg<kFwd, TF>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* filt = smem;
  uint8_t* abuf = smem + kFilterBytes;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(abuf + C::kStages * C::kBoxBytes);
  uint64_t* a_empty = a_full + C::kStages;
  uint64_t* filt_full = a_empty + C::kStages;
  uint64_t* acc_full = filt_full + 1;   // [2]
  uint64_t* acc_empty = acc_full + 2;   // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int half = TF ? (int)(blockIdx.x & 1) : 0;
  const int tile0 = TF ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int tstride = TF ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmW);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&a_full[s], 1); mbar_init(&a_empty[s], 1); }
    mbar_init(filt_full, 1);
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 128); }
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 2 * C::kN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_wait();
  grid_dep_launch();

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    bool filter_sent = false;
    for (int tile = tile0; tile < num_tiles; tile += tstride) {
      const int n = tile >> 1, wt = tile & 1;
      mbar_wait(&a_empty[stage], phase ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&a_full[stage], C::kBoxBytes);
        // box origin: column 8*wt - 2, row -2 (out-of-bounds pixels are zero-filled = SAME padding); the dgrad reads
        // this CTA's half of dy's channels on the tf32 arm, all 64 on the bf16 arm (128 B either way)
        tma_load_4d(abuf + stage * C::kBoxBytes, &tmA, &a_full[stage], (TF && !kFwd) ? half * 32 : 0, 8 * wt - 2, -2, n);
        if (!filter_sent) load_filter(filt, &tmW, filt_full, TF ? half * 32 : 0);
      }
      __syncwarp();
      filter_sent = true;
      if (++stage == C::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = TF ? make_idesc_tf32(128, C::kN, 0, kFwd ? 1 : 0) : make_idesc_bf16(128, C::kN, 0, 0);
    // A: K-major, 128 B rows, 8-row groups 2048 B apart (the box's row pitch).  B: the filter tap tile.
    const uint64_t da0 = make_smem_desc(smem_u32(abuf), 16, 2048, kSwz128);
    const uint64_t db0 = (TF && kFwd) ? make_smem_desc(smem_u32(filt), 4096, 512, kSwz128B32)   // [32 ci][32 co] MN-major
                                      : make_smem_desc(smem_u32(filt), 16, 1024, kSwz128);       // K-major (co = K)
    constexpr uint32_t kBStep = (TF && kFwd) ? 1024u : 32u;
    mbar_wait(filt_full, 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = tile0; tile < num_tiles; tile += tstride, ++it) {
      const int buf = it & 1;
      mbar_wait(&acc_empty[buf], ((it >> 1) & 1) ^ 1);
      mbar_wait(&a_full[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t tacc = tmem_base + buf * C::kN;
        const uint64_t da_s = desc_advance(da0, (uint32_t)(stage * C::kBoxBytes));
#pragma unroll
        for (int kh = 0; kh < 5; ++kh) {
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) {
            // forward: output (h, w) reads box pixel (h + kh, w + kw); dgrad: (h + 4 - kh, w + 4 - kw)
            const int rh = kFwd ? kh : 4 - kh, rw = kFwd ? kw : 4 - kw;
            const uint64_t da_t = desc_advance(da_s, (uint32_t)((rh * 16 + rw) * 128));
            const uint64_t db_t = desc_advance(db0, (uint32_t)((kh * 5 + kw) * 4096));
#pragma unroll
            for (int ks = 0; ks < C::kKSteps; ++ks) {
              const uint32_t acc = (kh > 0 || kw > 0 || ks > 0) ? 1u : 0u;
              if (TF) umma_tf32(tacc, desc_advance(da_t, ks * 32), desc_advance(db_t, ks * kBStep), idesc, acc);
              else    umma_bf16(tacc, desc_advance(da_t, ks * 32), desc_advance(db_t, ks * kBStep), idesc, acc);
            }
          }
        }
        umma_commit(&a_empty[stage]);
        umma_commit(&acc_full[buf]);
      }
      __syncwarp();
      if (++stage == C::kStages) { stage = 0; phase ^= 1; }
    }
  } else {
    // epilogue: warp q owns accumulator rows 32q..32q+31 = image rows 4q..4q+3 of the tile's 8 columns
    const int q = warp & 3;
    const int hl = lane >> 3, wl = lane & 7;
    int it = 0;
    for (int tile = tile0; tile < num_tiles; tile += tstride, ++it) {
      const int buf = it & 1;
      const int n = tile >> 1, wt = tile & 1;
      const int h = 4 * q + hl, w = 8 * wt + wl;  // conv output pixel of this lane
      mbar_wait(&acc_full[buf], (it >> 1) & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + buf * C::kN + ((uint32_t)(q * 32) << 16);
      if (kFwd) {
        const int self = (hl & 1) * 2 + (wl & 1);
        const int ph = h >> 1, pw = w >> 1;
        const int cbase = half * 32;
#pragma unroll 1
        for (int j = 0; j < C::kN / 16; ++j) {
          float v[16];
          tmem_ld16(tacc + j * 16, v);
          uint32_t am[4] = {0, 0, 0, 0};
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float x = fmaxf(v[i] + __ldg(bias + cbase + j * 16 + i), 0.f);
            int xi = self;
            float y = __shfl_xor_sync(0xffffffffu, x, 1);
            int yi = self ^ 1;
            if (y > x || (y == x && yi < xi)) { x = y; xi = yi; }
            y = __shfl_xor_sync(0xffffffffu, x, 8);
            yi = __shfl_xor_sync(0xffffffffu, xi, 8);
            if (y > x || (y == x && yi < xi)) { x = y; xi = yi; }
            v[i] = x;
            am[i >> 2] |= (uint32_t)xi << ((i & 3) * 8);
          }
          if (j == self && w < 14 && h < 14 && n < B) {
            const long long o = ((((long long)n * 7 + ph) * 7) + pw) * 64 + cbase + j * 16;
            float* po = static_cast<float*>(pool_out_v) + o;
#pragma unroll
            for (int i = 0; i < 16; i += 4)
              *reinterpret_cast<float4*>(po + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
            *reinterpret_cast<uint4*>(amax + o) = make_uint4(am[0], am[1], am[2], am[3]);
          }
        }
      } else {
        const bool ok = (w < 14) && (h < 14) && (n < B);
        float* o = dx + (long long)half * dx_plane_stride + ((((long long)n * 14 + h) * 14) + w) * 32;
#pragma unroll 1
        for (int j = 0; j < 2; ++j) {
          float v[16];
          tmem_ld16(tacc + j * 16, v);
          if (ok) {
#pragma unroll
            for (int i = 0; i < 16; i += 4)
              *reinterpret_cast<float4*>(o + j * 16 + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[buf]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 2 * C::kN);
  }
}

// ------------------------------------------------------------------------------------------------
// filter gradient: dW[(tap,ci), co] = sum_px in[px + shift(tap), ci] * dy[px, co].
// GEMM view per tap group g (4 taps): M = 4*32 (tap_local, ci), N = 64 co, K = 128 pixels of the tile.
//   A: four [128 px][32 ci] tiles (one TMA box per tap) read MN-major, stacked along M through LBO
//   B: the [128 px][64 co] dy tile read MN-major
//   bf16: A rows are 64 B (SW64), B rows 128 B (SW128).  tf32: every row is 128 B of fp32 (32 channels; dy as two
//   channel halves), 128B swizzle with 32 B atoms, 8 pixels per MMA.
// grid = (pixel chunks, 7 tap groups): a CTA owns ONE 128x64 accumulator (64 TMEM columns) over its share of the
// pixel tiles.
//
// Deterministic reduction (no floating-point atomics): every CTA stores its partial tile into a scratch plane, the
// `chunks` CTAs of a tap group meet at a counter, and each then sums a fixed slice of the group's output over all
// planes IN PLANE ORDER and writes it to dW (which may be a staging slot on another GPU).  All CTAs of the grid fit
// on the chip at once (<= 147 CTAs, one per SM), so a CTA spinning on its group's counter only ever waits for CTAs
// that are running or will be scheduled as soon as an SM frees up; the wait is bounded and traps.
// ------------------------------------------------------------------------------------------------
template <int TF>
struct WgCfg {
  static constexpr int kInTile = TF ? 16384 : 8192;           // [128 px][32 ci]
  static constexpr int kDyTile = TF ? 32768 : 16384;          // [128 px][64 co]
  static constexpr int kStageBytes = kDyTile + 4 * kInTile;   // 48 KB / 96 KB
  static constexpr int kStages = TF ? 2 : 4;
  static constexpr int kSmem = kStages * kStageBytes + 1024 + 512;
};
constexpr int kWgCounterWords = 64;  // head of the scratch buffer: one arrival counter per tap group

template <int TF>
__global__ void __launch_bounds__(kThreads, 1)
conv2_wgrad_tc_kernel(const __grid_constant__ CUtensorMap tmIn, const __grid_constant__ CUtensorMap tmDy,
                      float* __restrict__ dw, float* __restrict__ scratch, int num_tiles) {
  using C = WgCfg<TF>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
  uint64_t* empty = full + C::kStages;
  uint64_t* acc_full = empty + C::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = blockIdx.y;
  const int ntap = (g == 6) ? 1 : 4;
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmIn);
    prefetch_tmap(&tmDy);
    for (int s = 0; s < C::kStages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(acc_full, 1);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) tmem_alloc(tmem_slot, 64);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_wait();
  grid_dep_launch();

  if (warp == 0) {
    int stage = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int nb = tile / 7, hp = tile % 7;
      uint8_t* st = smem + stage * C::kStageBytes;
      mbar_wait(&empty[stage], phase ^ 1);
      if (elect_one()) {
        mbar_arrive_expect_tx(&full[stage], C::kDyTile + ntap * C::kInTile);
        tma_load_4d(st, &tmDy, &full[stage], 0, 0, 2 * hp, nb * 4);
        if (TF) tma_load_4d(st + 16384, &tmDy, &full[stage], 32, 0, 2 * hp, nb * 4);  // channels 32..63
        for (int tl = 0; tl < ntap; ++tl) {
          const int tap = g * 4 + tl, kh = tap / 5, kw = tap % 5;
          tma_load_4d(st + C::kDyTile + tl * C::kInTile, &tmIn, &full[stage], 0, kw - 2, 2 * hp + kh - 2, nb * 4);
        }
      }
      __syncwarp();
      if (++stage == C::kStages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = TF ? make_idesc_tf32(128, 64, 1, 1) : make_idesc_bf16(128, 64, 1, 1);
    // descriptor templates (start address = stage 0): A = tap tiles stacked along M, B = the dy tile
    const uint64_t db0 = TF ? make_smem_desc(smem_u32(smem), 16384, 512, kSwz128B32)          // N chunks = co halves
                            : make_smem_desc(smem_u32(smem), 8192, 1024, kSwz128);
    const uint64_t da0 = TF ? make_smem_desc(smem_u32(smem) + C::kDyTile, C::kInTile, 512, kSwz128B32)  // M chunks = taps
                            : make_smem_desc(smem_u32(smem) + C::kDyTile, 8192, 512, kSwz64);
    int stage = 0, it = 0;
    uint32_t phase = 0;
    for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++it) {
      mbar_wait(&full[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t da_s = desc_advance(da0, (uint32_t)(stage * C::kStageBytes));
        const uint64_t db_s = desc_advance(db0, (uint32_t)(stage * C::kStageBytes));
        if (TF) {
#pragma unroll
          for (int ks = 0; ks < 16; ++ks)  // 8 pixels (1024 B of every 128 B-wide tile) per MMA
            umma_tf32(tmem_base, desc_advance(da_s, ks * 1024), desc_advance(db_s, ks * 1024), idesc,
                      (it > 0 || ks > 0) ? 1u : 0u);
        } else {
#pragma unroll
          for (int ks = 0; ks < 8; ++ks)   // 16 pixels: two 512 B atoms of A (64 B rows), two 1024 B atoms of B
            umma_bf16(tmem_base, desc_advance(da_s, ks * 1024), desc_advance(db_s, ks * 2048), idesc,
                      (it > 0 || ks > 0) ? 1u : 0u);
        }
        umma_commit(&empty[stage]);
      }
      __syncwarp();
      if (++stage == C::kStages) { stage = 0; phase ^= 1; }
    }
    if (elect_one()) umma_commit(acc_full);
    __syncwarp();
  } else {
    // ---- epilogue warps: partial plane -> group rendezvous -> fixed-order slice reduction ----
    const int q = warp & 3;
    const int et = (int)threadIdx.x - 64;  // 0..127
    const int chunks = gridDim.x;
    unsigned* counters = reinterpret_cast<unsigned*>(scratch);
    float* planes = scratch + kWgCounterWords + (size_t)g * chunks * 8192;  // [chunks][128][64] of this group
    float* mine = planes + (size_t)blockIdx.x * 8192;
    mbar_wait(acc_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int j = 0; j < 4; ++j) {
      float v[16];
      tmem_ld16(tmem_base + j * 16 + ((uint32_t)(q * 32) << 16), v);
      float* o = mine + (q * 32 + lane) * 64 + j * 16;
#pragma unroll
      for (int i = 0; i < 16; i += 4) __stcg(reinterpret_cast<float4*>(o + i), make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");  // the four epilogue warps: all partial stores are issued
    if (et == 0) {
      __threadfence();  // one fence for the CTA (cumulative over the stores ordered before the barrier)
      // monotonic counter: this launch's arrivals fill one block of `chunks` counts
      const unsigned old = atomicAdd(&counters[g], 1u);
      const unsigned target = (old / (unsigned)chunks + 1u) * (unsigned)chunks;
      unsigned long long t0 = globaltimer_ns();
      unsigned itn = 0;
      while ((int)(*reinterpret_cast<volatile unsigned*>(&counters[g]) - target) < 0) {
        if ((++itn & 0x3fff) == 0 && globaltimer_ns() - t0 > kSpinTimeoutNs) {
          printf("tfd: conv2 wgrad group %d rendezvous timed out\n", g);
          __trap();
        }
      }
      __threadfence();
    }
    asm volatile("bar.sync 1, 128;" ::: "memory");
    // this CTA finalises a slice of the group's rows: float4 outputs, all planes' loads in flight, added in plane order
    const int rows_valid = ntap * 32;                               // rows of this group that exist in dW
    const int per = (rows_valid + chunks - 1) / chunks;             // rows this CTA finalises
    const int r0 = blockIdx.x * per, r1 = min(rows_valid, r0 + per);
    const float4* planes4 = reinterpret_cast<const float4*>(planes);
    float4* dw4 = reinterpret_cast<float4*>(dw + (size_t)g * 8192);  // (g*4 + tap_local)*32*64 + ci*64 + co == g*8192 + idx
    for (int i4 = et + r0 * 16; i4 < r1 * 16; i4 += 128) {
      float4 sacc = make_float4(0.f, 0.f, 0.f, 0.f);
      int c = 0;
      for (; c + 8 <= chunks; c += 8) {
        float4 v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) v[u] = __ldcg(planes4 + (size_t)(c + u) * 2048 + i4);
#pragma unroll
        for (int u = 0; u < 8; ++u) { sacc.x += v[u].x; sacc.y += v[u].y; sacc.z += v[u].z; sacc.w += v[u].w; }
      }
      for (; c < chunks; ++c) {
        const float4 v = __ldcg(planes4 + (size_t)c * 2048 + i4);
        sacc.x += v.x; sacc.y += v.y; sacc.z += v.z; sacc.w += v.w;
      }
      dw4[i4] = sacc;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 64);
  }
}

int num_sms() { return tfd_num_sms(); }

}  // namespace

// TFD_CONV2_BOX: 1 (default) = the one-box kernel where the operand rows are 128 B (tf32 forward / dgrad, bf16 dgrad),
// 0 = the five-box kernel everywhere.
static int conv2_box_mode() {
  static const int v = [] { const char* e = getenv("TFD_CONV2_BOX"); return e ? atoi(e) : 1; }();
  return v;
}

static int conv2_grid(int tiles, int tf) {
  int grid = tf ? 2 * tiles : tiles;  // tf32: (tile, channel half) work items, even CTAs half 0 / odd CTAs half 1
  const int cap = tf ? (num_sms() & ~1) : num_sms();
  return grid < cap ? grid : cap;
}

extern "C" int tfd_conv2_fwd(int dtype, const void* in, const void* w, const float* b, void* pool_out,
                             unsigned char* argmax, int B, void* stream) {
  TFD_REQUIRE(dtype == TFD_DTYPE_BF16 || dtype == TFD_DTYPE_TF32, "conv2_fwd: unknown operand type %d", dtype);
  const int tf = dtype == TFD_DTYPE_TF32;
  CUtensorMap tmA, tmW;
  int rc = tf ? tfd_make_tmap_nhwc(&tmA, in, B, 14, 14, 32, 1, 12, 16, 32, 128, 4)
              : tfd_make_tmap_nhwc_bf16(&tmA, in, B, 14, 14, 32, 1, 12, 16, 64);
  if (rc) return rc;
  // filter [800][64]: bf16 whole rows; tf32 one 32-channel half per box, 32 B swizzle atoms (MN-major B operand)
  rc = tf ? tfd_make_tmap_2d(&tmW, w, 800, 64, 64, 200, 32, 129, 4) : tfd_make_tmap_2d_bf16(&tmW, w, 800, 64, 64, 200, 64, 128);
  if (rc) return rc;
  const int tiles = B * 2;
  const int grid = conv2_grid(tiles, tf);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tf && conv2_box_mode()) {
    using C = BoxCfg<true, 1>;
    rc = tfd_make_tmap_nhwc(&tmA, in, B, 14, 14, 32, 1, 20, 16, 32, 128, 4);
    if (rc) return rc;
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_box_kernel<true, 1>), C::kSmem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(conv2_box_kernel<true, 1>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW, b, pool_out,
                              argmax, (float*)nullptr, B, tiles, 0LL));
  } else if (tf) {
    using C = ConvCfg<true, 1>;
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_tc_kernel<true, 1>), C::kSmem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(conv2_tc_kernel<true, 1>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW, b, pool_out,
                              argmax, (float*)nullptr, B, tiles, 0LL));
  } else {
    using C = ConvCfg<true, 0>;
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_tc_kernel<true, 0>), C::kSmem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(conv2_tc_kernel<true, 0>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW, b, pool_out,
                              argmax, (float*)nullptr, B, tiles, 0LL));
  }
  return 0;
}

extern "C" int tfd_conv2_dgrad(int dtype, const void* dy, const void* w, float* dx, int B, void* stream) {
  TFD_REQUIRE(dtype == TFD_DTYPE_BF16 || dtype == TFD_DTYPE_TF32, "conv2_dgrad: unknown operand type %d", dtype);
  const int tf = dtype == TFD_DTYPE_TF32;
  CUtensorMap tmA, tmW;
  int rc = tf ? tfd_make_tmap_nhwc(&tmA, dy, B, 14, 14, 64, 1, 12, 16, 32, 128, 4)
              : tfd_make_tmap_nhwc_bf16(&tmA, dy, B, 14, 14, 64, 1, 12, 16, 128);
  if (rc) return rc;
  rc = tf ? tfd_make_tmap_2d(&tmW, w, 800, 64, 64, 200, 32, 128, 4) : tfd_make_tmap_2d_bf16(&tmW, w, 800, 64, 64, 200, 64, 128);
  if (rc) return rc;
  const int tiles = B * 2;
  const int grid = conv2_grid(tiles, tf);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (conv2_box_mode()) {
    rc = tf ? tfd_make_tmap_nhwc(&tmA, dy, B, 14, 14, 64, 1, 20, 16, 32, 128, 4)
            : tfd_make_tmap_nhwc(&tmA, dy, B, 14, 14, 64, 1, 20, 16, 64, 128, 2);
    if (rc) return rc;
    if (tf) {
      using C = BoxCfg<false, 1>;
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_box_kernel<false, 1>), C::kSmem)) return arc;
      TFD_CUDA_CHECK(tfd_launch(conv2_box_kernel<false, 1>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW,
                                (const float*)nullptr, (void*)nullptr, (unsigned char*)nullptr, dx, B, tiles,
                                (long long)B * 14 * 14 * 32));
    } else {
      using C = BoxCfg<false, 0>;
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_box_kernel<false, 0>), C::kSmem)) return arc;
      TFD_CUDA_CHECK(tfd_launch(conv2_box_kernel<false, 0>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW,
                                (const float*)nullptr, (void*)nullptr, (unsigned char*)nullptr, dx, B, tiles, 0LL));
    }
  } else if (tf) {
    using C = ConvCfg<false, 1>;
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_tc_kernel<false, 1>), C::kSmem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(conv2_tc_kernel<false, 1>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW,
                              (const float*)nullptr, (void*)nullptr, (unsigned char*)nullptr, dx, B, tiles,
                              (long long)B * 14 * 14 * 32));
  } else {
    using C = ConvCfg<false, 0>;
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_tc_kernel<false, 0>), C::kSmem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(conv2_tc_kernel<false, 0>, dim3(grid), dim3(kThreads), C::kSmem, st, tmA, tmW,
                              (const float*)nullptr, (void*)nullptr, (unsigned char*)nullptr, dx, B, tiles, 0LL));
  }
  return 0;
}

static int wgrad_chunks(int B) {
  const int tiles = ((B + 3) / 4) * 7;
  int chunks = num_sms() / 7;
  if (chunks > tiles) chunks = tiles;
  if (chunks < 1) chunks = 1;
  return chunks;
}

extern "C" long long tfd_conv2_wgrad_scratch_floats(int B) {
  return (long long)kWgCounterWords + 7LL * wgrad_chunks(B) * 8192;
}

extern "C" int tfd_conv2_wgrad(int dtype, const void* in, const void* dy, float* dw, float* scratch, int B, void* stream) {
  TFD_REQUIRE(dtype == TFD_DTYPE_BF16 || dtype == TFD_DTYPE_TF32, "conv2_wgrad: unknown operand type %d", dtype);
  TFD_REQUIRE(scratch != nullptr, "conv2_wgrad: scratch planes missing (tfd_conv2_wgrad_scratch_floats(B) floats, zeroed once)");
  const int tf = dtype == TFD_DTYPE_TF32;
  CUtensorMap tmIn, tmDy;
  int rc = tf ? tfd_make_tmap_nhwc(&tmIn, in, B, 14, 14, 32, 4, 2, 16, 32, 129, 4)
              : tfd_make_tmap_nhwc_bf16(&tmIn, in, B, 14, 14, 32, 4, 2, 16, 64);
  if (rc) return rc;
  rc = tf ? tfd_make_tmap_nhwc(&tmDy, dy, B, 14, 14, 64, 4, 2, 16, 32, 129, 4)
          : tfd_make_tmap_nhwc_bf16(&tmDy, dy, B, 14, 14, 64, 4, 2, 16, 128);
  if (rc) return rc;
  const int tiles = ((B + 3) / 4) * 7;
  const int chunks = wgrad_chunks(B);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (tf) {
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_wgrad_tc_kernel<1>), WgCfg<1>::kSmem)) return arc;
    // 194 KB alone would put the SM at the 196 KB carve-out; at the largest one a conv1-wgrad block (18.5 KB) fits next to it
    tfd_prefer_max_carveout(reinterpret_cast<const void*>(conv2_wgrad_tc_kernel<1>));
    TFD_CUDA_CHECK(tfd_launch(conv2_wgrad_tc_kernel<1>, dim3(chunks, 7), dim3(kThreads), WgCfg<1>::kSmem, st, tmIn, tmDy, dw,
                              scratch, tiles));
  } else {
    if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(conv2_wgrad_tc_kernel<0>), WgCfg<0>::kSmem)) return arc;
    // 194 KB alone would put the SM at the 196 KB carve-out; at the largest one a conv1-wgrad block (18.5 KB) fits next to it
    tfd_prefer_max_carveout(reinterpret_cast<const void*>(conv2_wgrad_tc_kernel<0>));
    TFD_CUDA_CHECK(tfd_launch(conv2_wgrad_tc_kernel<0>, dim3(chunks, 7), dim3(kThreads), WgCfg<0>::kSmem, st, tmIn, tmDy, dw,
                              scratch, tiles));
  }
  return 0;
}
