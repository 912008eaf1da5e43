// Parameter server over NVLink SHARP (NVLS multicast) -- EXPERIMENTAL: compiled and unit-checked on the host side,
// not yet run on hardware ("What the measurements say to build next", item 4).
//
// Every GPU maps one symmetric heap twice: at a unicast address (its own copy) and at a multicast address (all
// copies at once, through the switch).  That turns the reference's star topology (every byte through the ps,
// /root/reference/mnist_python_m.py:177) inside out without giving up its semantics (the owner shard alone holds
// the master weights and the Adam slots and applies the update):
//
//   push  (X2): a worker writes its gradient into ITS OWN copy of the heap -- plain local stores from the wgrad
//               epilogues -- and then adds 1 to the arrival counter of every shard with ONE multimem.red.
//   apply     : the owner shard waits for the arrivals and reads each gradient element with multimem.ld_reduce:
//               the switch fetches the element from every GPU and returns the sum, so the shard's NVLink ingress
//               is 1x the gradient instead of (workers)x.
//   pull  (X1): the owner stores the updated parameter (fp32, or bf16 for tensor-core operands) with multimem.st:
//               one store lands in every worker's mirror, egress 1x instead of (workers)x; workers then read
//               parameters from local memory only.  The version counter of every GPU is bumped the same way.
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"

namespace {

using namespace tfd;
typedef __nv_bfloat16 bf16;

__device__ __forceinline__ float4 mc_ld_reduce_add_f32x4(const float* mc) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc)
               : "memory");
  return r;
}
__device__ __forceinline__ void mc_st_f32x4(float* mc, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}
__device__ __forceinline__ void mc_st_bf16x4(bf16* mc, uint32_t lo, uint32_t hi) {
  asm volatile("multimem.st.relaxed.sys.global.v2.bf16x2 [%0], {%1, %2};" ::"l"(mc), "r"(lo), "r"(hi) : "memory");
}
__device__ __forceinline__ void mc_red_release_add_u32(unsigned* mc, unsigned v) {
  asm volatile("multimem.red.release.sys.global.add.u32 [%0], %1;" ::"l"(mc), "r"(v) : "memory");
}

constexpr int kMaxNvlsPieces = 16;
struct NvlsPieces {
  tfd_nvls_piece_t p[kMaxNvlsPieces];
  int n;
};

struct NvlsArgs {
  float *w, *m, *v;          // owner-private arenas (this shard's pieces)
  const float* mc_grad;      // multicast alias of the workers' gradient arenas (single-shard layout)
  float* mc_mirror_w;        // multicast alias of the fp32 parameter mirrors
  bf16* mc_mirror_shadow;    // multicast alias of the bf16 shadow mirrors
  float scale, lr, beta1, beta2, eps;
  unsigned* t_counter;       // [0] Adam step of this shard, [1] block counter (owner-private)
  const unsigned* arrive;    // this GPU's arrival counter (workers add 1 per step through the multicast alias)
  unsigned arrivals_per_step;
  unsigned* mc_version;      // multicast alias of the version counter: every shard adds 1 per applied step
  unsigned* global_step;     // owner-private; shard 0's is the global step
  int adam;                  // 0: publish the current parameters only (initialisation / restore)
};

// One element group (4 floats) of one piece: gradient from the switch, Adam in place, publish through the switch.
__device__ __forceinline__ void nvls_update4(const NvlsArgs& a, const tfd_nvls_piece_t& pc, long long e, float lr_t) {
  const long long o = pc.owner_off + e;
  float4 w4 = *reinterpret_cast<const float4*>(a.w + o);
  float wv[4] = {w4.x, w4.y, w4.z, w4.w};
  if (a.adam) {
    const float4 g4 = mc_ld_reduce_add_f32x4(a.mc_grad + pc.grad_off + e);
    const float4 m4 = *reinterpret_cast<const float4*>(a.m + o), v4 = *reinterpret_cast<const float4*>(a.v + o);
    float mv[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
    const float gv[4] = {g4.x * a.scale, g4.y * a.scale, g4.z * a.scale, g4.w * a.scale};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mv[j] = a.beta1 * mv[j] + (1.f - a.beta1) * gv[j];
      vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gv[j] * gv[j];
      wv[j] -= __fdividef(lr_t * mv[j], vv[j] * rsqrtf(fmaxf(vv[j], 1e-38f)) + a.eps);  // TF form (eps outside)
    }
    *reinterpret_cast<float4*>(a.w + o) = make_float4(wv[0], wv[1], wv[2], wv[3]);
    *reinterpret_cast<float4*>(a.m + o) = make_float4(mv[0], mv[1], mv[2], mv[3]);
    *reinterpret_cast<float4*>(a.v + o) = make_float4(vv[0], vv[1], vv[2], vv[3]);
  }
  if (pc.ld_shadow == 0) {
    mc_st_f32x4(a.mc_mirror_w + pc.mirror_off + e, make_float4(wv[0], wv[1], wv[2], wv[3]));
  } else {  // ld_shadow == cols (checked on the host): the bf16 rows are dense, 4 elements = one 8-byte store
    mc_st_bf16x4(a.mc_mirror_shadow + pc.shadow_off + e, pack_bf16x2(wv[0], wv[1]), pack_bf16x2(wv[2], wv[3]));
  }
}

__global__ void __launch_bounds__(256) nvls_apply_kernel(const NvlsArgs a, const __grid_constant__ NvlsPieces pt) {
  const unsigned t = a.t_counter[0] + 1u;
  if (a.adam) {
    if (threadIdx.x == 0) spin_until_ge(a.arrive, a.arrivals_per_step * t, "nvls apply: gradient arrivals");
    __syncthreads();
  }
  const float om_b1t = -expm1f((float)t * logf(a.beta1)), om_b2t = -expm1f((float)t * logf(a.beta2));
  const float lr_t = a.lr * sqrtf(om_b2t) / om_b1t;
  for (int q = 0; q < pt.n; ++q) {
    const tfd_nvls_piece_t pc = pt.p[q];
    const long long n4 = (pc.numel + 3) >> 2;  // pieces are padded to 64 elements in every arena
    for (long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x)
      nvls_update4(a, pc, i4 << 2, lr_t);
  }
  // publish: every block's multicast stores are fenced before it counts itself; the last block releases the version
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned done = atomicAdd(&a.t_counter[1], 1u);
    if (done == gridDim.x - 1) {
      a.t_counter[1] = 0;
      if (a.adam) {  // (a publish-only launch is followed by a host-side barrier instead of a version bump)
        a.t_counter[0] = t;
        if (a.global_step) *a.global_step = *a.global_step + 1;
        __threadfence_system();
        mc_red_release_add_u32(a.mc_version, 1u);
      }
    }
  }
}

// First kernel of a worker's step: wait (on LOCAL memory) until every shard has published the parameters of this
// step, which also means every shard has finished reading this worker's previous gradient.
__global__ void nvls_gate_kernel(const unsigned* version, const unsigned* step, unsigned per_step, unsigned base) {
  if (threadIdx.x == 0 && blockIdx.x == 0) spin_until_ge(version, base + per_step * step[0], "nvls gate: parameter version");
}

// Last kernel of a worker's step: the gradient arena is complete (stream order); tell every shard with one
// multicast reduction and advance the local step.
__global__ void nvls_signal_kernel(unsigned* mc_arrive, unsigned* step) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    __threadfence_system();
    mc_red_release_add_u32(mc_arrive, 1u);
    step[0] = step[0] + 1u;
  }
}

}  // namespace

#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int tfd_nvls_apply(float* w, float* m, float* v, const float* mc_grad, float* mc_mirror_w,
                              void* mc_mirror_shadow, const tfd_nvls_piece_t* pieces, int num_pieces, float scale,
                              float lr, float beta1, float beta2, float eps, unsigned* t_counter,
                              const unsigned* arrive, unsigned arrivals_per_step, unsigned* mc_version,
                              unsigned* global_step, int adam, void* stream) {
  TFD_REQUIRE(num_pieces >= 1 && num_pieces <= kMaxNvlsPieces, "nvls_apply: %d pieces (max %d)", num_pieces, kMaxNvlsPieces);
  TFD_REQUIRE(mc_mirror_w && mc_mirror_shadow && mc_version && t_counter, "nvls_apply: multicast pointers are required");
  TFD_REQUIRE(!adam || (mc_grad && arrive && m && v), "nvls_apply: the Adam form needs gradients, slots and arrivals");
  NvlsPieces pt;
  pt.n = num_pieces;
  long long total = 0;
  for (int i = 0; i < num_pieces; ++i) {
    pt.p[i] = pieces[i];
    TFD_REQUIRE(pieces[i].owner_off % 4 == 0 && pieces[i].grad_off % 4 == 0 && pieces[i].mirror_off % 4 == 0 &&
                    pieces[i].shadow_off % 4 == 0,
                "nvls_apply: piece %d offsets must be multiples of 4 elements", i);
    TFD_REQUIRE(pieces[i].ld_shadow == 0 || pieces[i].ld_shadow == pieces[i].cols,
                "nvls_apply: piece %d: bf16-published tensors need cols %% 8 == 0 (multimem.st moves >= 32 bits)", i);
    total += pieces[i].numel;
  }
  NvlsArgs a;
  a.w = w; a.m = m; a.v = v; a.mc_grad = mc_grad; a.mc_mirror_w = mc_mirror_w;
  a.mc_mirror_shadow = static_cast<bf16*>(mc_mirror_shadow);
  a.scale = scale; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.t_counter = t_counter; a.arrive = arrive; a.arrivals_per_step = arrivals_per_step; a.mc_version = mc_version;
  a.global_step = global_step; a.adam = adam;
  long long blocks = (total / 4 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 3) blocks = 148 * 3;
  TFD_CUDA_CHECK(tfd_launch(nvls_apply_kernel, dim3((unsigned)blocks), dim3(256), 0, ST(stream), a, pt));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_nvls_gate(const unsigned* version, const unsigned* step, unsigned per_step, unsigned base, void* stream) {
  TFD_REQUIRE(version && step, "nvls_gate: null pointer");
  nvls_gate_kernel<<<1, 32, 0, ST(stream)>>>(version, step, per_step, base);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_nvls_signal(unsigned* mc_arrive, unsigned* step, void* stream) {
  TFD_REQUIRE(mc_arrive && step, "nvls_signal: null pointer");
  nvls_signal_kernel<<<1, 32, 0, ST(stream)>>>(mc_arrive, step);
  TFD_LAUNCH_CHECK();
  return 0;
}
