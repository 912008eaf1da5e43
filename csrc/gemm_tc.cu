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
// warps 2..5 = epilogue (TMEM -> registers -> global).  Operand pointers may be peer-mapped (NVLink) addresses:
// TMA pulls B tiles straight from the owning ps GPU (fused parameter pull), and the epilogue may store /
// red.add to a peer staging buffer (fused gradient push).
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"

namespace {

using namespace tfd;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int kThreads = 192;

struct GemmParams {
  int M, N, K;          // logical problem
  int kblocks;          // ceil(K / BK)
  int splits;           // gridDim.z
  int mode;             // TFD_EPI_*
  float alpha;
  float* out_f32;
  __nv_bfloat16* out_bf16;
  long long ld_out;
  const float* bias;            // [N] or null
  const __nv_bfloat16* mask;    // relu mask source (same shape as out), or null
  long long ld_mask;
  const uint32_t* wait_flag;    // optional: spin until *wait_flag >= wait_value before touching B (pull gate)
  uint32_t wait_value;
};

template <int BN>
struct Cfg {
  static constexpr int kStageBytes = BM * 128 + BN * 128;
  static constexpr int kStages = (BN == 256) ? 4 : (BN == 128 ? 6 : 8);
  static constexpr int kSmemBytes = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
};

template <int A_MN, int B_MN, int BN>
__global__ void __launch_bounds__(kThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, GemmParams p) {
  using C = Cfg<BN>;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + C::kStages * C::kStageBytes);
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
    prefetch_tmap(&tmB);
    for (int s = 0; s < C::kStages; ++s) {
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

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      if (p.wait_flag) spin_until_ge(p.wait_flag, p.wait_value, "gemm pull gate");
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&empty_bar[stage], phase ^ 1);
        uint8_t* sa = smem + stage * C::kStageBytes;
        uint8_t* sb = sa + BM * 128;
        mbar_arrive_expect_tx(&full_bar[stage], C::kStageBytes);
        const int k0 = kb * BK;
        if (A_MN == 0) {
          tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);            // box [128 rows(m), 64 k]
        } else {
          tma_load_2d(sa, &tmA, &full_bar[stage], m0, k0);            // box [64 rows(k), 64 m]
          tma_load_2d(sa + 8192, &tmA, &full_bar[stage], m0 + 64, k0);
        }
        if (B_MN == 0) {
          tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);            // box [BN rows(n), 64 k]
        } else {
#pragma unroll
          for (int c = 0; c < BN / 64; ++c) tma_load_2d(sb + c * 8192, &tmB, &full_bar[stage], n0 + c * 64, k0);
        }
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      constexpr uint32_t idesc = make_idesc_bf16(BM, BN, A_MN, B_MN);
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb_begin; kb < kb_end; ++kb) {
        mbar_wait(&full_bar[stage], phase);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + stage * C::kStageBytes);
        const uint32_t sb = sa + BM * 128;
#pragma unroll
        for (int ks = 0; ks < BK / 16; ++ks) {
          // K-major: 16 bf16 = 32 B inside the 128 B swizzle row.  MN-major: 16 k-rows = two 1024 B atoms.
          const uint64_t da = A_MN ? make_smem_desc(sa + ks * 2048, 8192, 1024, kSwz128)
                                   : make_smem_desc(sa + ks * 32, 16, 1024, kSwz128);
          const uint64_t db = B_MN ? make_smem_desc(sb + ks * 2048, 8192, 1024, kSwz128)
                                   : make_smem_desc(sb + ks * 32, 16, 1024, kSwz128);
          umma_bf16(tmem_base, da, db, idesc, (kb > kb_begin || ks > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[stage]);  // frees the smem slot once these MMAs have read it
        if (++stage == C::kStages) { stage = 0; phase ^= 1; }
      }
      umma_commit(tmem_full_bar);
    }
  } else {
    // ===== epilogue: warps 2..5 own TMEM lane quadrants (warp % 4) =====
    const int q = warp & 3;
    const int row = m0 + q * 32 + lane;
    mbar_wait(tmem_full_bar, 0);
    tc_fence_after();
    const bool row_ok = row < p.M;
    const bool vec_ok = ((p.ld_out & 3) == 0);
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
      if (p.mode == TFD_EPI_BIAS_RELU_BF16) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = fmaxf(v[i], 0.f);
      }
      if (p.mode == TFD_EPI_RELUMASK_BF16) {
        const __nv_bfloat16* mrow = p.mask + (long long)row * p.ld_mask + col0;
#pragma unroll
        for (int i = 0; i < 16; ++i)
          if (i < ncols && !(__bfloat162float(mrow[i]) > 0.f)) v[i] = 0.f;
      }
      if (p.mode == TFD_EPI_STORE_F32) {
        float* o = p.out_f32 + (long long)row * p.ld_out + col0;
        if (ncols == 16 && vec_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) *reinterpret_cast<float4*>(o + i) = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
          for (int i = 0; i < ncols; ++i) o[i] = v[i];
        }
      } else if (p.mode == TFD_EPI_ATOMIC_F32) {
        float* o = p.out_f32 + (long long)row * p.ld_out + col0;
        if (ncols == 16 && vec_ok) {
#pragma unroll
          for (int i = 0; i < 16; i += 4) red_add_v4_f32(o + i, v[i], v[i + 1], v[i + 2], v[i + 3]);
        } else {
          for (int i = 0; i < ncols; ++i) red_add_f32(o + i, v[i]);
        }
      } else {  // bf16 outputs
        __nv_bfloat16* o = p.out_bf16 + (long long)row * p.ld_out + col0;
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

template <int A_MN, int B_MN, int BN>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const GemmParams& p, cudaStream_t stream) {
  using C = Cfg<BN>;
  auto kern = gemm_tc_kernel<A_MN, B_MN, BN>;
  static bool attr_set = false;
  if (!attr_set) {
    TFD_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::kSmemBytes));
    attr_set = true;
  }
  dim3 grid((p.N + BN - 1) / BN, (p.M + BM - 1) / BM, p.splits);
  kern<<<grid, kThreads, C::kSmemBytes, stream>>>(tmA, tmB, p);
  TFD_LAUNCH_CHECK();
  return 0;
}

}  // namespace

// a_mn / b_mn: operand major-ness as stored (see header).  A: a_mn=0 -> [M,K] ld=lda ; a_mn=1 -> [K,M].
// B: b_mn=0 -> [N,K] ; b_mn=1 -> [K,N].
extern "C" int tfd_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M,
                             int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                             const float* bias, const void* relu_mask, long long ld_mask, const unsigned* wait_flag,
                             unsigned wait_value, void* stream) {
  TFD_REQUIRE(M > 0 && N > 0 && K > 0, "gemm: empty problem %dx%dx%d", M, N, K);
  TFD_REQUIRE(bn == 64 || bn == 128 || bn == 256, "gemm: bn must be 64/128/256, got %d", bn);
  GemmParams p;
  p.M = M; p.N = N; p.K = K;
  p.kblocks = (K + BK - 1) / BK;
  p.splits = splits < 1 ? 1 : (splits > p.kblocks ? p.kblocks : splits);
  TFD_REQUIRE(p.splits == 1 || mode == TFD_EPI_ATOMIC_F32, "gemm: split-K needs the atomic epilogue");
  p.mode = mode; p.alpha = alpha;
  p.out_f32 = (mode == TFD_EPI_STORE_F32 || mode == TFD_EPI_ATOMIC_F32) ? static_cast<float*>(out) : nullptr;
  p.out_bf16 = p.out_f32 ? nullptr : static_cast<__nv_bfloat16*>(out);
  p.ld_out = ld_out; p.bias = bias;
  p.mask = static_cast<const __nv_bfloat16*>(relu_mask); p.ld_mask = ld_mask;
  TFD_REQUIRE(mode != TFD_EPI_RELUMASK_BF16 || relu_mask, "gemm: relu-mask epilogue without a mask");
  p.wait_flag = wait_flag; p.wait_value = wait_value;

  CUtensorMap tmA, tmB;
  int rc;
  if (a_mn == 0) rc = tfd_make_tmap_2d_bf16(&tmA, A, M, K, lda, BM, BK, 128);
  else           rc = tfd_make_tmap_2d_bf16(&tmA, A, K, M, lda, BK, 64, 128);
  if (rc) return rc;
  if (b_mn == 0) rc = tfd_make_tmap_2d_bf16(&tmB, B, N, K, ldb, bn, BK, 128);
  else           rc = tfd_make_tmap_2d_bf16(&tmB, B, K, N, ldb, BK, 64, 128);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
#define TFD_DISPATCH(AM, BMj)                                              \
  if (a_mn == AM && b_mn == BMj) {                                         \
    if (bn == 64) return launch<AM, BMj, 64>(tmA, tmB, p, st);             \
    if (bn == 128) return launch<AM, BMj, 128>(tmA, tmB, p, st);           \
    return launch<AM, BMj, 256>(tmA, tmB, p, st);                          \
  }
  TFD_DISPATCH(0, 0) TFD_DISPATCH(0, 1) TFD_DISPATCH(1, 0) TFD_DISPATCH(1, 1)
#undef TFD_DISPATCH
  TFD_REQUIRE(false, "gemm: bad major flags %d %d", a_mn, b_mn);
}
