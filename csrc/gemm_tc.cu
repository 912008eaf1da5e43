// tcgen05 / TMEM / TMA GEMM for the dense layers: D[M,N] = A[M,K] * B[K,N], bf16 operands, fp32 accumulate.
//
// Replaces the tf.matmul call sites of the reference (/root/reference/mnist_python_m.py:121,127 forward and the
// autodiff-generated wgrad/dgrad matmuls of m:222).  One kernel covers forward, dgrad and wgrad because either
// operand may be K-major or MN-major: no transposed copies of activations or weights are ever materialised.
//
//   forward  y = x W        A = x [M,K] K-major          B = W [K,N] MN-major
//   dgrad    dx = dy W^T    A = dy [M,N'] K-major        B = W [K',N'] read as [N,K] K-major
//   wgrad    dW = x^T dy    A = x [K=batch, M] MN-major  B = dy [K=batch, N] MN-major
//
// Warp roles (192 threads): warp 0 = TMA producer, warp 1 = TMEM allocator + single-thread MMA issuer,
// warps 2..5 = epilogue (TMEM -> registers -> global).
//
// Fused parameter pull / gradient push: the weight matrix is row-sharded over up to 8
// parameter-server GPUs.  The kernel takes one tensor map (or output pointer) per shard segment and addresses
// the shards' peer-mapped memory directly:
//   TFD_SEG_B_K   forward: B's K rows are sharded; K-block kb is TMA-loaded from the shard that owns it, after
//                 the producer has seen that shard's parameter version flag (acquire, system scope)
//   TFD_SEG_B_N   dgrad:   B read as [N,K]; the N tile is TMA-loaded from its owner shard
//   TFD_SEG_OUT_M wgrad:   the output rows are sharded; the epilogue stores each tile straight into this
//                 worker's staging slot on the owner shard (no local gradient copy, no separate push of it)
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"
#include <stdlib.h>

namespace {

using namespace tfd;

constexpr int BM = 128;
constexpr int kBKbf16 = 64;  // K elements per pipeline stage for bf16 operands (32 for tf32): 128 B either way
constexpr int kThreads = 192;
constexpr int kMaxSeg = TFD_MAX_SEG;

struct TmapArray {
  CUtensorMap m[kMaxSeg];
};

struct SegInfo {
  int mode;                         // TFD_SEG_*
  int nseg;
  int bound[kMaxSeg + 1];           // segment row boundaries (rows of the sharded dimension)
  float* out_ptr[kMaxSeg];          // TFD_SEG_OUT_M
  const uint32_t* wait_flag[kMaxSeg];
  uint32_t wait_value;
  const uint32_t* wait_value_ptr;   // optional device word added to wait_value (the worker's step counter)
};

struct GemmParams {
  int M, N, K;          // logical problem
  int kblocks;          // ceil(K / BK)
  int splits;           // gridDim.z
  int stages;           // pipeline depth actually used (<= Cfg::kStages): short K loops take less shared memory, so
                        // several CTAs fit on an SM and one CTA's store drain overlaps another's loads
  int bar_off;          // byte offset of the barrier block behind max(stages * stage bytes, epilogue staging bytes)
  int mode;             // TFD_EPI_*
  float alpha;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  long long ld_out;
  const float* bias;            // [N] or null
  const __nv_bfloat16* mask;    // relu mask source (same shape as out), or null
  long long ld_mask;
  long long split_stride;       // STORE_F32 split-K: plane z at out_f32 + z*split_stride
};

// Element-type dependent tile constants.  A pipeline stage always holds BM x 128 B of A and BN x 128 B of B: 64 bf16 or
// 32 tf32 (fp32 in memory; the tensor core reads the upper 19 bits) along K per stage, four MMAs per stage either way
// (UMMA K = 32 B / element size).  An MN-major operand arrives as boxes of [BK k-rows][128 B]:
//   bf16: 64 elements wide, 128B swizzle (16 B chunks), descriptor LBO = box stride, SBO = 1024 (8 k-rows)
//   tf32: 32 elements wide, 128B swizzle with 32 B atoms (the only MN-major layout the tf32 MMA reads: 4 k-rows per
//         swizzle atom), descriptor layout type 1, SBO = 512, eight k-rows (= one MMA) every 1024 B
template <int TF>
struct El {
  static constexpr int kBK = TF ? 32 : 64;            // elements along K per stage
  static constexpr int kMnW = TF ? 32 : 64;           // elements along M/N per MN-major box (128 B)
  static constexpr int kMnBoxBytes = kBK * 128;       // 4 KB / 8 KB
  static constexpr int kMnKStep = TF ? 1024 : 2048;   // bytes between consecutive MMAs' k-rows in an MN-major box
  static constexpr uint32_t kMnSbo = TF ? 512 : 1024;
  static constexpr uint32_t kMnSwz = TF ? kSwz128B32 : kSwz128;
};

template <int BN>
struct Cfg {
  static constexpr int kStageBytes = BM * 128 + BN * 128;
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

__device__ __forceinline__ int find_seg(const SegInfo& sg, int row) {
  int s = 0;
#pragma unroll
  for (int i = 1; i < kMaxSeg; ++i)
    if (i < sg.nseg && row >= sg.bound[i]) s = i;
  return s;
}

template <int A_MN, int B_MN, int BN, int TF>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ TmapArray tmBs, GemmParams p,
               const __grid_constant__ SegInfo sg) {
  using C = Cfg<BN>;
  using E = El<TF>;
  constexpr int BK = E::kBK;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + p.bar_off);
  const int nstages = p.stages;
  uint64_t* empty_bar = full_bar + C::kStages;
  uint64_t* tmem_full_bar = empty_bar + C::kStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN;
  const int m0 = blockIdx.y * BM;
  const int kb_begin = (int)(((long long)blockIdx.z * p.kblocks) / p.splits);
  const int kb_end = (int)(((long long)(blockIdx.z + 1) * p.kblocks) / p.splits);

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmA);
    prefetch_tmap(&tmBs.m[0]);
    for (int s = 0; s < nstages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(tmem_full_bar, 1);
    fence_barrier_init();
    fence_proxy_async_smem();
  }
  if (warp == 1) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  grid_dep_wait();    // prologue above overlapped the previous kernel's tail; from here on we touch global memory
  grid_dep_launch();

  // Producer and MMA warps run their loops with all 32 lanes (warp-uniform control flow, addresses and descriptors:
  // they stay in uniform registers); only the TMA / MMA / commit instructions sit under elect_one().  Loops under
  // `if (lane == 0)` made the compiler shuttle five registers to the uniform file around every UTCHMMA.
  if (warp == 0) {
    // ===== TMA producer =====
    int stage = 0;
    uint32_t phase = 0;
    int gated = -1;  // last shard whose version flag this producer has already seen
    const uint32_t want = sg.wait_value + (sg.wait_value_ptr ? *sg.wait_value_ptr : 0u);
    // dgrad: the whole N tile comes from one shard
    int seg_n = 0, nrow0 = n0;
    if (sg.mode == TFD_SEG_B_N) {
      seg_n = find_seg(sg, n0);
      nrow0 = n0 - sg.bound[seg_n];
      if (sg.wait_flag[seg_n]) spin_until_ge(sg.wait_flag[seg_n], want, "gemm pull gate (N)");
    } else if (sg.mode != TFD_SEG_B_K && sg.wait_flag[0]) {
      spin_until_ge(sg.wait_flag[0], want, "gemm pull gate");
    }
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait(&empty_bar[stage], phase ^ 1);
      uint8_t* sa = smem + stage * C::kStageBytes;
      uint8_t* sb = sa + BM * 128;
      const int k0 = kb * BK;
      int s = 0, krow = k0;
      if (B_MN != 0 && sg.mode == TFD_SEG_B_K) {
        s = find_seg(sg, k0);
        krow = k0 - sg.bound[s];
        if (s != gated) {
          if (sg.wait_flag[s]) spin_until_ge(sg.wait_flag[s], want, "gemm pull gate (K)");
          gated = s;
        }
      }
      if (elect_one()) {
        mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
        if (A_MN == 0) {
          tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);            // box [128 rows(m), BK k]
        } else {
#pragma unroll
          for (int c = 0; c < BM / E::kMnW; ++c)                      // boxes [BK rows(k), 128 B of m]
            tma_load_2d(sa + c * E::kMnBoxBytes, &tmA, &full_bar[stage], m0 + c * E::kMnW, k0);
        }
        if (B_MN == 0) {
          tma_load_2d(sb, &tmBs.m[seg_n], &full_bar[stage], k0, nrow0);  // box [BN rows(n), BK k]
        } else {
#pragma unroll
          for (int c = 0; c < BN / E::kMnW; ++c)
            tma_load_2d(sb + c * E::kMnBoxBytes, &tmBs.m[s], &full_bar[stage], n0 + c * E::kMnW, krow);
        }
      }
      __syncwarp();
      if (++stage == nstages) { stage = 0; phase ^= 1; }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = TF ? make_idesc_tf32(BM, BN, A_MN, B_MN) : make_idesc_bf16(BM, BN, A_MN, B_MN);
    // descriptor templates at stage 0; per stage / per MMA only the start address moves.
    // K-major: one MMA's K = 32 B inside the 128 B swizzle row.  MN-major: its k-rows are E::kMnKStep bytes further
    // down the box (bf16: 16 rows = two 1024 B atoms; tf32: 8 rows = two 512 B atoms).
    const uint32_t s0 = smem_u32(smem);
    const uint64_t da0 = A_MN ? make_smem_desc(s0, E::kMnBoxBytes, E::kMnSbo, E::kMnSwz) : make_smem_desc(s0, 16, 1024, kSwz128);
    const uint64_t db0 = B_MN ? make_smem_desc(s0 + BM * 128, E::kMnBoxBytes, E::kMnSbo, E::kMnSwz)
                              : make_smem_desc(s0 + BM * 128, 16, 1024, kSwz128);
    constexpr uint32_t kAStep = A_MN ? E::kMnKStep : 32, kBStep = B_MN ? E::kMnKStep : 32;
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = kb_begin; kb < kb_end; ++kb) {
      mbar_wait(&full_bar[stage], phase);
      tc_fence_after();
      if (elect_one()) {
        const uint64_t da_s = desc_advance(da0, (uint32_t)(stage * C::kStageBytes));
        const uint64_t db_s = desc_advance(db0, (uint32_t)(stage * C::kStageBytes));
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint64_t da = desc_advance(da_s, ks * kAStep), db = desc_advance(db_s, ks * kBStep);
          if (TF) umma_tf32(tmem_base, da, db, idesc, (kb > kb_begin || ks > 0) ? 1u : 0u);
          else    umma_bf16(tmem_base, da, db, idesc, (kb > kb_begin || ks > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
        if (kb == kb_end - 1) umma_commit(tmem_full_bar);
      }
      __syncwarp();
      if (++stage == nstages) { stage = 0; phase ^= 1; }
    }
  } else {
    // ===== epilogue: warps 2..5 own TMEM lane quadrants (warp % 4) =====
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    float* out_f32 = p.out_f32;
    int out_row = row;
    if (sg.mode == TFD_SEG_OUT_M) {  // sharded output: this tile lives in one shard's staging slot
      const int s = find_seg(sg, m0);
      out_f32 = sg.out_ptr[s];
      out_row = row - sg.bound[s];
    }
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool row_ok = row < p.M;
    const bool vec_ok = ((p.ld_out & 3) == 0);
    if (p.mode == TFD_EPI_STORE_F32 && vec_ok && n0 + BN <= p.N) {
      // Coalesced fp32 store: TMEM -> registers -> shared (the pipeline stages are idle by now) -> global, so a
      // warp instruction writes whole 512 B row segments instead of 32 scattered 16 B pieces.  That matters most
      // when `out` is a staging slot on another GPU: NVLink wants long contiguous writes.
      constexpr int kPitch = BN + 4;  // floats; (BN/4 + 1) float4 per row is odd -> conflict-free both ways
      float* stg = reinterpret_cast<float*>(smem) + (size_t)q * 32 * kPitch;
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
#pragma unroll
        for (int i = 0; i < 16; i += 4) {
          float4 o4 = make_float4(v[i] * p.alpha, v[i + 1] * p.alpha, v[i + 2] * p.alpha, v[i + 3] * p.alpha);
          if (p.bias) {
            o4.x += __ldg(p.bias + n0 + c + i); o4.y += __ldg(p.bias + n0 + c + i + 1);
            o4.z += __ldg(p.bias + n0 + c + i + 2); o4.w += __ldg(p.bias + n0 + c + i + 3);
          }
          *reinterpret_cast<float4*>(stg + lane * kPitch + c + i) = o4;
        }
      }
      __syncwarp();
      float* obase = out_f32 + (long long)blockIdx.z * p.split_stride + n0;
      constexpr int kLanesPerRow = BN / 4;                       // float4 per row
      constexpr int kRowsPerInst = kLanesPerRow >= 32 ? 1 : 32 / kLanesPerRow;
      constexpr int kInstPerRow = kLanesPerRow >= 32 ? kLanesPerRow / 32 : 1;
#pragma unroll 4
      for (int r0 = 0; r0 < 32; r0 += kRowsPerInst) {
        const int r = r0 + (kLanesPerRow >= 32 ? 0 : lane / kLanesPerRow);
        const int grow = m0 + q * 32 + r;
#pragma unroll
        for (int k = 0; k < kInstPerRow; ++k) {
          const int col = ((kLanesPerRow >= 32 ? lane : lane % kLanesPerRow) + 32 * k) * 4;
          if (grow < p.M) {
            const float4 o4 = *reinterpret_cast<const float4*>(stg + r * kPitch + col);
            *reinterpret_cast<float4*>(obase + (long long)(out_row - lane + r) * p.ld_out + col) = o4;
          }
        }
      }
    } else if (p.mode == TFD_EPI_STORE_BF16 && (p.ld_out & 7) == 0 && n0 + BN <= p.N) {
      // Coalesced bf16 store (bf16 gradient push): same staging idea, rows of BN*2 bytes.
      constexpr int kPitchB = BN * 2 + 16;  // bytes; odd multiple of 16 -> conflict-free
      uint8_t* stg = smem + (size_t)q * 32 * kPitchB;
#pragma unroll 1
      for (int c = 0; c < BN; c += 16) {
        float v[16];
        tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
        if (p.bias) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = v[i] * p.alpha + __ldg(p.bias + n0 + c + i);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] *= p.alpha;
        }
        uint4* d = reinterpret_cast<uint4*>(stg + lane * kPitchB + c * 2);
        d[0] = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
        d[1] = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]),
                          pack_bf16x2(v[14], v[15]));
      }
      __syncwarp();
      __nv_bfloat16* ob = (sg.mode == TFD_SEG_OUT_M ? reinterpret_cast<__nv_bfloat16*>(out_f32) : p.out_bf16) + n0;
      constexpr int kLanesPerRow = BN / 8;  // 16-byte pieces per row
      constexpr int kRowsPerInst = 32 / kLanesPerRow;
#pragma unroll 4
      for (int r0 = 0; r0 < 32; r0 += kRowsPerInst) {
        const int r = r0 + lane / kLanesPerRow;
        const int piece = lane % kLanesPerRow;
        if (m0 + q * 32 + r < p.M) {
          const uint4 o4 = *reinterpret_cast<const uint4*>(stg + r * kPitchB + piece * 16);
          *reinterpret_cast<uint4*>(ob + (long long)(out_row - lane + r) * p.ld_out + piece * 8) = o4;
        }
      }
    } else
#pragma unroll 1
    for (int c = 0; c < BN; c += 16) {
      float v[16];
      tmem_ld16(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, v);
      const int col0 = n0 + c;
      if (!row_ok || col0 >= p.N) continue;
      const int ncols = min(16, p.N - col0);
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] *= p.alpha;
      if (p.bias) {
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i < ncols) v[i] += __ldg(p.bias + col0 + i);
      }
      if (p.mode == TFD_EPI_BIAS_RELU_BF16 || p.mode == TFD_EPI_BIAS_RELU_F32) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      if (p.mode == TFD_EPI_RELUMASK_F32) {
        const float* mrow = reinterpret_cast<const float*>(p.mask) + (long long)row * p.ld_mask + col0;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i < ncols && !(mrow[i] > 0.f)) v[i] = 0.f;
      }
      if (p.mode == TFD_EPI_RELUMASK_BF16) {
        const __nv_bfloat16* mrow = p.mask + (long long)row * p.ld_mask + col0;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i < ncols && !(__bfloat162float(mrow[i]) > 0.f)) v[i] = 0.f;
      }
      if (p.mode == TFD_EPI_STORE_F32 || p.mode == TFD_EPI_BIAS_RELU_F32 || p.mode == TFD_EPI_RELUMASK_F32) {
        float* o = out_f32 + (long long)blockIdx.z * p.split_stride + (long long)out_row * p.ld_out + col0;
        if (ncols == 16 && vec_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
          for (int i = 0; i < ncols; ++i) o[i] = v[i];
        }
      } else if (p.mode == TFD_EPI_ATOMIC_F32) {
        float* o = out_f32 + (long long)out_row * p.ld_out + col0;
        if (ncols == 16 && vec_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) red_add_v4_f32(o + i, v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
          for (int i = 0; i < ncols; ++i) red_add_f32(o + i, v[i]);
        }
      } else {  // bf16 outputs
        __nv_bfloat16* o = (sg.mode == TFD_SEG_OUT_M ? reinterpret_cast<__nv_bfloat16*>(out_f32) : p.out_bf16) +
                           (long long)out_row * p.ld_out + col0;
        if (ncols == 16 && (p.ld_out & 7) == 0) {
          uint4 w0 = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]),
                                pack_bf16x2(v[6], v[7]));
          uint4 w1 = make_uint4(pack_bf16x2(v[8], v[9]), pack_bf16x2(v[10], v[11]), pack_bf16x2(v[12], v[13]),
                                pack_bf16x2(v[14], v[15]));
          *reinterpret_cast<uint4*>(o) = w0;
          *reinterpret_cast<uint4*>(o + 8) = w1;
        } else {
          for (int i = 0; i < ncols; ++i) o[i] = __float2bfloat16(v[i]);
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
  }
}

template <int A_MN, int B_MN, int BN, int TF>
int launch(const CUtensorMap& tmA, const TmapArray& tmBs, const GemmParams& p, const SegInfo& sg, cudaStream_t stream) {
  using C = Cfg<BN>;
  auto kern = gemm_tc_kernel<A_MN, B_MN, BN, TF>;
  if (int rc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(kern), C::kSmemBytes)) return rc;
  dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, p.splits);
  GemmParams q = p;
  const int per_split = (p.kblocks + p.splits - 1) / p.splits;
  q.stages = per_split < C::kStages ? per_split : C::kStages;
  // short K loops (wgrad: K = batch): two stages are enough and keep the CTA at ~69 KB so three fit on an SM and one
  // CTA's store drain overlaps another's loads (tf32 has 4 K blocks for K = 128 where bf16 has 2)
  if (per_split <= 4 && q.stages > 2) q.stages = 2;
  static const int cap_kb = [] {  // tuning knob: cap the pipeline so two CTAs (of this or another kernel) share an SM
    const char* e = getenv("TFD_GEMM_SMEM_KB");
    return e ? atoi(e) : 0;
  }();
  if (cap_kb > 0) {
    int cap_stages = (cap_kb * 1024 - 1280) / C::kStageBytes;
    if (cap_stages < 2) cap_stages = 2;
    if (q.stages > cap_stages) q.stages = cap_stages;
  }
  size_t staging = 0;  // the coalesced store epilogues stage the tile in the (then idle) pipeline buffers
  if (p.mode == TFD_EPI_STORE_F32) staging = (size_t)BM * (BN + 4) * 4;
  if (p.mode == TFD_EPI_STORE_BF16) staging = (size_t)BM * (BN * 2 + 16);
  size_t data = (size_t)q.stages * C::kStageBytes;
  if (data < staging) data = staging;
  data = (data + 1023) & ~size_t(1023);
  q.bar_off = (int)data;
  const size_t smem = data + 1024 /*align slack*/ + 256 /*barriers*/;
  TFD_REQUIRE(smem <= (size_t)C::kSmemBytes, "gemm: shared memory %zu exceeds the configured maximum", smem);
  TFD_CUDA_CHECK(tfd_launch(kern, grid, dim3(kThreads), smem, stream, tmA, tmBs, q, sg));
  return 0;
}

}  // namespace

static int gemm_seg_impl(const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb, int b_mn, int M,
                         int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                         const float* bias, const void* relu_mask, long long ld_mask, long long split_stride,
                         int seg_mode, int nseg, const int* seg_rows, void* const* out_ptrs,
                         const unsigned* const* wait_flags, unsigned wait_value, const unsigned* wait_value_ptr,
                         int dtype, void* stream);

// Segmented form: the sharded operand is given as nseg pieces.  seg_rows[nseg+1] are the row boundaries of the
// sharded dimension (K for TFD_SEG_B_K, N for TFD_SEG_B_N, M for TFD_SEG_OUT_M; multiples of 128, first = 0).
// B_ptrs[s] points at shard s's rows (row 0 of the piece); out_ptrs[s] likewise for TFD_SEG_OUT_M.
extern "C" int tfd_gemm_bf16_seg(const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb,
                                 int b_mn, int M, int N, int K, int bn, int splits, int mode, float alpha, void* out,
                                 long long ld_out, const float* bias, const void* relu_mask, long long ld_mask,
                                 long long split_stride, int seg_mode, int nseg, const int* seg_rows,
                                 void* const* out_ptrs, const unsigned* const* wait_flags, unsigned wait_value,
                                 const unsigned* wait_value_ptr, void* stream) {
  return gemm_seg_impl(A, lda, a_mn, B_ptrs, ldb, b_mn, M, N, K, bn, splits, mode, alpha, out, ld_out, bias, relu_mask,
                       ld_mask, split_stride, seg_mode, nseg, seg_rows, out_ptrs, wait_flags, wait_value, wait_value_ptr,
                       TFD_DTYPE_BF16, stream);
}

// Same GEMM with the operand type as a parameter: TFD_DTYPE_BF16 (bf16 in memory) or TFD_DTYPE_TF32 (fp32 in memory,
// tcgen05 kind::tf32: 10-bit mantissa products, fp32 accumulation) -- the precision arm that may be set against the
// reference's fp32 graph (/root/reference/mnist_python_m.py:198-200).  Leading dimensions in elements of that type.
extern "C" int tfd_gemm_seg(int dtype, const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb,
                            int b_mn, int M, int N, int K, int bn, int splits, int mode, float alpha, void* out,
                            long long ld_out, const float* bias, const void* relu_mask, long long ld_mask,
                            long long split_stride, int seg_mode, int nseg, const int* seg_rows, void* const* out_ptrs,
                            const unsigned* const* wait_flags, unsigned wait_value, const unsigned* wait_value_ptr,
                            void* stream) {
  return gemm_seg_impl(A, lda, a_mn, B_ptrs, ldb, b_mn, M, N, K, bn, splits, mode, alpha, out, ld_out, bias, relu_mask,
                       ld_mask, split_stride, seg_mode, nseg, seg_rows, out_ptrs, wait_flags, wait_value, wait_value_ptr,
                       dtype, stream);
}

static int gemm_seg_impl(const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb, int b_mn, int M,
                         int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                         const float* bias, const void* relu_mask, long long ld_mask, long long split_stride,
                         int seg_mode, int nseg, const int* seg_rows, void* const* out_ptrs,
                         const unsigned* const* wait_flags, unsigned wait_value, const unsigned* wait_value_ptr,
                         int dtype, void* stream) {
  TFD_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
  TFD_REQUIRE(dtype == TFD_DTYPE_BF16 || dtype == TFD_DTYPE_TF32, "gemm: unknown operand type %d", dtype);
  const int tf = dtype == TFD_DTYPE_TF32;
  const int BK = tf ? 32 : 64, eb = tf ? 4 : 2, mnw = tf ? 32 : 64;
  const int swz_mn = tf ? 129 : 128;  // MN-major boxes: 128B swizzle, 32 B atoms for tf32
  TFD_REQUIRE(bn == 64 || bn == 128 || bn == 256, "gemm: bn must be 64/128/256, got %d", bn);
  TFD_REQUIRE(nseg >= 1 && nseg <= kMaxSeg, "gemm: %d segments (max %d)", nseg, kMaxSeg);
  TFD_REQUIRE(seg_mode == TFD_SEG_NONE || seg_rows != nullptr, "gemm: segmented call without boundaries");
  TFD_REQUIRE(seg_mode != TFD_SEG_B_K || b_mn == 1, "gemm: TFD_SEG_B_K needs B stored [K,N]");
  TFD_REQUIRE(seg_mode != TFD_SEG_B_N || b_mn == 0, "gemm: TFD_SEG_B_N needs B stored [N,K]");
  TFD_REQUIRE(!tf || mode == TFD_EPI_STORE_F32 || mode == TFD_EPI_ATOMIC_F32 || mode == TFD_EPI_BIAS_RELU_F32 ||
                  mode == TFD_EPI_RELUMASK_F32 || mode == TFD_EPI_STORE_BF16,
              "gemm: tf32 operands go with fp32 outputs or a plain bf16 store (bf16 gradient slots); epilogue %d", mode);
  TFD_REQUIRE(seg_mode != TFD_SEG_OUT_M || mode == TFD_EPI_STORE_F32 || mode == TFD_EPI_ATOMIC_F32 ||
                  mode == TFD_EPI_STORE_BF16,
              "gemm: TFD_SEG_OUT_M needs a plain store epilogue (fp32, fp32 atomic or bf16)");
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.kblocks = (K + BK - 1) / BK;
  p.splits = splits < 1 ? 1 : (splits > p.kblocks ? p.kblocks : splits);
  TFD_REQUIRE(p.splits == 1 || mode == TFD_EPI_ATOMIC_F32 || (mode == TFD_EPI_STORE_F32 && split_stride > 0),
              "gemm: split-K needs the atomic epilogue or fp32 partial planes (split_stride)");
  TFD_REQUIRE(p.splits == 1 || bias == nullptr, "gemm: bias with split-K would be added once per split");
  p.split_stride = split_stride;
  p.mode = mode; p.alpha = alpha;
  p.out_f32 = (mode == TFD_EPI_STORE_F32 || mode == TFD_EPI_ATOMIC_F32 || mode == TFD_EPI_BIAS_RELU_F32 ||
               mode == TFD_EPI_RELUMASK_F32) ? static_cast<float*>(out) : nullptr;
  p.out_bf16 = p.out_f32 ? nullptr : static_cast<__nv_bfloat16*>(out);
  p.ld_out = ld_out; p.bias = bias;
  p.mask = static_cast<const __nv_bfloat16*>(relu_mask); p.ld_mask = ld_mask;
  TFD_REQUIRE((mode != TFD_EPI_RELUMASK_BF16 && mode != TFD_EPI_RELUMASK_F32) || relu_mask,
              "gemm: relu-mask epilogue without a mask");

  SegInfo sg;
  memset(&sg, 0, sizeof(sg));
  sg.mode = seg_mode; sg.nseg = nseg; sg.wait_value = wait_value; sg.wait_value_ptr = wait_value_ptr;
  for (int s = 0; s < nseg; ++s) {
    sg.bound[s] = seg_rows ? seg_rows[s] : 0;
    sg.out_ptr[s] = out_ptrs ? static_cast<float*>(out_ptrs[s]) : nullptr;
    sg.wait_flag[s] = wait_flags ? wait_flags[s] : nullptr;
    if (seg_mode != TFD_SEG_NONE) {
      TFD_REQUIRE(sg.bound[s] % 128 == 0, "gemm: segment boundary %d not a multiple of 128", sg.bound[s]);
      TFD_REQUIRE(s == 0 || sg.bound[s] > sg.bound[s - 1], "gemm: segment boundaries must increase");
    }
  }
  sg.bound[nseg] = seg_rows ? seg_rows[nseg] : 0;
  TFD_REQUIRE(seg_mode != TFD_SEG_B_N || bn <= 128, "gemm: TFD_SEG_B_N needs bn <= 128 (an N tile must not straddle shards)");

  CUtensorMap tmA;
  TmapArray tmBs;
  int rc;
  if (a_mn == 0) rc = tfd_make_tmap_2d(&tmA, A, M, K, lda, BM, BK, 128, eb);
  else           rc = tfd_make_tmap_2d(&tmA, A, K, M, lda, BK, mnw, swz_mn, eb);
  if (rc) return rc;
  const bool seg_b = (seg_mode == TFD_SEG_B_K || seg_mode == TFD_SEG_B_N);
  for (int s = 0; s < (seg_b ? nseg : 1); ++s) {
    const int rows = seg_b ? (sg.bound[s + 1] - sg.bound[s]) : (b_mn == 0 ? N : K);
    TFD_REQUIRE(rows > 0, "gemm: empty segment %d", s);
    if (b_mn == 0) rc = tfd_make_tmap_2d(&tmBs.m[s], B_ptrs[s], rows, K, ldb, bn, BK, 128, eb);
    else           rc = tfd_make_tmap_2d(&tmBs.m[s], B_ptrs[s], rows, N, ldb, BK, mnw, swz_mn, eb);
    if (rc) return rc;
  }
  for (int s = (seg_b ? nseg : 1); s < kMaxSeg; ++s) tmBs.m[s] = tmBs.m[0];
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define TFD_DISPATCH(AM, BMj)                                                           \
  if (a_mn == AM && b_mn == BMj) {                                                      \
    if (tf) {                                                                           \
      if (bn == 64) return launch<AM, BMj, 64, 1>(tmA, tmBs, p, sg, st);                \
      if (bn == 128) return launch<AM, BMj, 128, 1>(tmA, tmBs, p, sg, st);              \
      return launch<AM, BMj, 256, 1>(tmA, tmBs, p, sg, st);                             \
    }                                                                                   \
    if (bn == 64) return launch<AM, BMj, 64, 0>(tmA, tmBs, p, sg, st);                  \
    if (bn == 128) return launch<AM, BMj, 128, 0>(tmA, tmBs, p, sg, st);                \
    return launch<AM, BMj, 256, 0>(tmA, tmBs, p, sg, st);                               \
  }
  TFD_DISPATCH(0, 0) TFD_DISPATCH(0, 1) TFD_DISPATCH(1, 0) TFD_DISPATCH(1, 1)
#undef TFD_DISPATCH
  TFD_REQUIRE(false, "gemm: bad major flags %d %d", a_mn, b_mn);
}

// a_mn / b_mn: operand major-ness as stored (see header).  A: a_mn=0 -> [M,K] ld=lda ; a_mn=1 -> [K,M].
// B: b_mn=0 -> [N,K] ; b_mn=1 -> [K,N].
extern "C" int tfd_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M,
                             int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                             const float* bias, const void* relu_mask, long long ld_mask, long long split_stride,
                             const unsigned* wait_flag, unsigned wait_value, void* stream) {
  const void* bp[1] = {B};
  const unsigned* wf[1] = {wait_flag};
  return tfd_gemm_bf16_seg(A, lda, a_mn, bp, ldb, b_mn, M, N, K, bn, splits, mode, alpha, out, ld_out, bias, relu_mask,
                           ld_mask, split_stride, TFD_SEG_NONE, 1, nullptr, nullptr, wf, wait_value, nullptr, stream);
}

extern "C" int tfd_gemm(int dtype, const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M,
                        int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                        const float* bias, const void* relu_mask, long long ld_mask, long long split_stride,
                        void* stream) {
  const void* bp[1] = {B};
  return tfd_gemm_seg(dtype, A, lda, a_mn, bp, ldb, b_mn, M, N, K, bn, splits, mode, alpha, out, ld_out, bias, relu_mask,
                      ld_mask, split_stride, TFD_SEG_NONE, 1, nullptr, nullptr, nullptr, 0, nullptr, stream);
}

// ------------------------------------------------------------------------------------------------
// Split-K GEMM for callers that cannot allocate.
//
// The split-K path above needs scratch: one fp32 partial plane per split, summed afterwards.  Normally the
// caller allocates those planes.  Some callers cannot: a host that has handed every byte of HBM to its own pool
// (or runs under a fixed memory budget) has nothing left to cudaMalloc from, and a latency-critical serving loop
// must not enter the CUDA allocator at all, because cudaMalloc/cudaFree synchronise the whole device under the
// other streams' feet.  For them the planes live in the file-scope array below: it is part of the library image,
// so every GPU that loads the library carries these 8 MiB from load time until the process exits, whether or not
// this entry point is ever called.
//
// The array's name is only a host-side stand-in for the device object -- handing it to a kernel would make the
// kernel use host (or unmapped) memory -- so the first call on each device asks the runtime for the real device
// address with cudaGetSymbolAddress, with that device current, and keeps it.  That first call therefore does
// one-time set-up work (address lookup, kernel attribute) and should be made before any stream capture begins.
// ------------------------------------------------------------------------------------------------
namespace {

constexpr long long kNoAllocScratchFloats = 2ll << 20;  // 2 Mi floats = 8 MiB: splits * M * N must fit
static __device__ __align__(256) float g_noalloc_planes[kNoAllocScratchFloats];

constexpr int kMaxDevices = 64;
float* g_noalloc_planes_dev[kMaxDevices] = {nullptr};  // device address of the array, per device ordinal

__global__ void __launch_bounds__(256)
sum_planes_kernel(const float* __restrict__ planes, int nplanes, long long plane_stride, int M, int N,
                  float* __restrict__ out, long long ld_out, const float* __restrict__ bias, int relu) {
  grid_dep_wait();
  const long long total = (long long)M * N;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / N), c = (int)(i % N);
    float s = bias ? bias[c] : 0.f;
    for (int z = 0; z < nplanes; ++z) s += planes[z * plane_stride + i];  // fixed order: deterministic
    out[(long long)r * ld_out + c] = relu ? fmaxf(s, 0.f) : s;
  }
}

}  // namespace

extern "C" int tfd_gemm_bf16_splitk_noalloc(const void* A, long long lda, int a_mn, const void* B, long long ldb,
                                            int b_mn, int M, int N, int K, int bn, int splits, float* out,
                                            long long ld_out, const float* bias, int relu, void* stream) {
  TFD_REQUIRE(M > 0 && N > 0 && K > 0, "splitk_noalloc: empty problem %dx%dx%d", M, N, K);
  TFD_REQUIRE(splits >= 1, "splitk_noalloc: splits must be >= 1, got %d", splits);
  const int kblocks = (K + kBKbf16 - 1) / kBKbf16;
  if (splits > kblocks) splits = kblocks;
  const long long need = (long long)splits * M * N;
  TFD_REQUIRE(need <= kNoAllocScratchFloats,
              "splitk_noalloc: splits*M*N = %lld floats exceeds the built-in scratch of %lld floats (8 MiB); "
              "use fewer splits or tfd_gemm_bf16 with caller-provided planes",
              need, kNoAllocScratchFloats);
  int dev = 0;
  TFD_CUDA_CHECK(cudaGetDevice(&dev));
  TFD_REQUIRE(dev >= 0 && dev < kMaxDevices, "splitk_noalloc: device ordinal %d out of range", dev);
  if (g_noalloc_planes_dev[dev] == nullptr) {
    // set-up, once per device, with that device current: the device address of the file-scope array
    float* p = nullptr;
    TFD_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_noalloc_planes));
    TFD_REQUIRE(p != nullptr, "splitk_noalloc: cudaGetSymbolAddress returned a null device address");
    g_noalloc_planes_dev[dev] = p;
  }
  float* planes = g_noalloc_planes_dev[dev];
  // Every plane element the reduction reads is written by this call's GEMM first, so nothing in the array has to
  // start at zero and the previous call's contents are simply overwritten.
  const long long plane_stride = (long long)M * N;
  int rc = tfd_gemm_bf16(A, lda, a_mn, B, ldb, b_mn, M, N, K, bn, splits, TFD_EPI_STORE_F32, 1.0f, planes, N, nullptr,
                         nullptr, 0, plane_stride, nullptr, 0, stream);
  if (rc) return rc;
  long long blocks = (plane_stride + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  sum_planes_kernel<<<(unsigned)blocks, 256, 0, static_cast<cudaStream_t>(stream)>>>(planes, splits, plane_stride, M, N, out,
                                                                                    ld_out, bias, relu);
  TFD_LAUNCH_CHECK();
  return 0;
}
