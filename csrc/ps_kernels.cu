// Parameter-server kernels: gradient push into the owning shard's staging slots, the shard-side
// reduce + scale + Adam + bf16 publish, small-tensor pull, and the cudaIpc symmetric heap.
//
// Reference behaviour replaced: the implicit ps<->worker Send/Recv traffic of TF's gRPC runtime (SURVEY.md 2.4
// X1-X5) and tf.train.SyncReplicasOptimizer / AdamOptimizer applying on the ps
// (/root/reference/mnist_python_m.py:208-222).  Tokens (X4) become monotonic counters in peer-visible memory:
// `arrive` (workers -> shard, one increment per pushed gradient) and `version` (shard -> workers, = number of
// applied updates), written with release and read with acquire semantics at system scope.
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"

namespace {

using namespace tfd;
typedef __nv_bfloat16 bf16;

constexpr int kMaxPieces = 40;
struct PieceTable {
  tfd_piece_t p[kMaxPieces];
  int n;
};

struct AdamArgs {
  float* w; float* m; float* v; bf16* shadow;
  const float* const* slots;
  int num_slots;
  float scale;
  long long n;
  float lr, beta1, beta2, eps;
  unsigned* t_counter;          // [0] = applied updates so far, [1] = block-done scratch
  const unsigned* wait_counter; // arrivals (local memory, incremented by workers over NVLink)
  unsigned* version_flag;
  unsigned* global_step;
};

__global__ void __launch_bounds__(256) adam_apply_kernel(AdamArgs a, const __grid_constant__ PieceTable pt) {
  __shared__ float s_lr_t;
  __shared__ unsigned s_t;
  if (threadIdx.x == 0) {
    const unsigned t = a.t_counter[0] + 1;  // this update's 1-based index
    if (a.wait_counter) spin_until_ge(a.wait_counter, (unsigned)a.num_slots * t, "adam: gradient arrivals");
    s_t = t;
    const double b1t = pow((double)a.beta1, (double)t), b2t = pow((double)a.beta2, (double)t);
    s_lr_t = (float)((double)a.lr * sqrt(1.0 - b2t) / (1.0 - b1t));
  }
  __syncthreads();
  const float lr_t = s_lr_t;
  const long long n4 = a.n >> 2;
  for (long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < n4; i4 += (long long)gridDim.x * blockDim.x) {
    const long long i = i4 << 2;
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int r = 0; r < a.num_slots; ++r) {
      const float4 x = *reinterpret_cast<const float4*>(a.slots[r] + i);
      g.x += x.x; g.y += x.y; g.z += x.z; g.w += x.w;
    }
    float gv[4] = {g.x * a.scale, g.y * a.scale, g.z * a.scale, g.w * a.scale};
    float4 w4 = *reinterpret_cast<float4*>(a.w + i);
    float4 m4 = *reinterpret_cast<float4*>(a.m + i);
    float4 v4 = *reinterpret_cast<float4*>(a.v + i);
    float wv[4] = {w4.x, w4.y, w4.z, w4.w}, mv[4] = {m4.x, m4.y, m4.z, m4.w}, vv[4] = {v4.x, v4.y, v4.z, v4.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      mv[j] = a.beta1 * mv[j] + (1.f - a.beta1) * gv[j];
      vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gv[j] * gv[j];
      wv[j] -= lr_t * mv[j] / (sqrtf(vv[j]) + a.eps);  // TF form: epsilon outside the bias correction
    }
    *reinterpret_cast<float4*>(a.w + i) = make_float4(wv[0], wv[1], wv[2], wv[3]);
    *reinterpret_cast<float4*>(a.m + i) = make_float4(mv[0], mv[1], mv[2], mv[3]);
    *reinterpret_cast<float4*>(a.v + i) = make_float4(vv[0], vv[1], vv[2], vv[3]);
    if (a.shadow) {
      // pieces are 64-element aligned in the arena, so one float4 lies in one piece
      for (int q = 0; q < pt.n; ++q) {
        const tfd_piece_t& pc = pt.p[q];
        if (i >= pc.arena_off && i < pc.arena_off + pc.numel) {
          if (pc.ld_shadow > 0) {
            const long long e0 = i - pc.arena_off;
            if (pc.ld_shadow == pc.cols) {
              bf16* d = a.shadow + pc.shadow_off + e0;
              *reinterpret_cast<uint2*>(d) = make_uint2(pack_bf16x2(wv[0], wv[1]), pack_bf16x2(wv[2], wv[3]));
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const long long e = e0 + j;
                if (e < pc.numel)
                  a.shadow[pc.shadow_off + (e / pc.cols) * pc.ld_shadow + (e % pc.cols)] = __float2bfloat16(wv[j]);
              }
            }
          }
          break;
        }
      }
    }
  }
  // publish: last block bumps the counters and releases the new parameter version
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned done = atomicAdd(&a.t_counter[1], 1u);
    if (done == gridDim.x - 1) {
      a.t_counter[1] = 0;
      a.t_counter[0] = s_t;
      if (a.global_step) *a.global_step = *a.global_step + 1;
      __threadfence_system();
      if (a.version_flag) st_release_sys_u32(a.version_flag, s_t);
    }
  }
}

struct PushSeg {
  const float* src;
  float* dst;
  long long n;
};
constexpr int kMaxSegs = 32;
struct PushTable {
  PushSeg seg[kMaxSegs];
  unsigned* arrive[kMaxSegs];
  int nseg, narrive;
};

// grid (blocks_per_seg, nseg)
__global__ void __launch_bounds__(256) push_kernel(const __grid_constant__ PushTable t, float scale, unsigned* block_counter) {
  const PushSeg s = t.seg[blockIdx.y];
  const long long n4 = s.n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 x = *reinterpret_cast<const float4*>(s.src + (i << 2));
    x.x *= scale; x.y *= scale; x.z *= scale; x.w *= scale;
    *reinterpret_cast<float4*>(s.dst + (i << 2)) = x;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += (long long)gridDim.x * blockDim.x)
    s.dst[i] = s.src[i] * scale;
  __threadfence_system();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y;
    if (atomicAdd(block_counter, 1u) == total - 1) {
      *block_counter = 0;
      __threadfence_system();
      for (int k = 0; k < t.narrive; ++k) atom_add_release_sys_u32(t.arrive[k], 1u);
    }
  }
}

struct PullSeg {
  const float* src;
  float* dst;
  long long n;
  const unsigned* version;
};
struct PullTable {
  PullSeg seg[kMaxSegs];
  int nseg;
};

__global__ void __launch_bounds__(256) pull_kernel(const __grid_constant__ PullTable t, unsigned want) {
  const PullSeg s = t.seg[blockIdx.y];
  if (s.version) {
    if (threadIdx.x == 0) spin_until_ge(s.version, want, "pull: parameter version");
    __syncthreads();
  }
  const long long n4 = s.n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
    *reinterpret_cast<float4*>(s.dst + (i << 2)) = *reinterpret_cast<const float4*>(s.src + (i << 2));
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += (long long)gridDim.x * blockDim.x)
    s.dst[i] = s.src[i];
}

}  // namespace

#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int tfd_adam_apply(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                              float scale, long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1,
                              float beta2, float eps, unsigned* t_counter, const unsigned* wait_counter,
                              unsigned* version_flag, unsigned* global_step, void* stream) {
  TFD_REQUIRE(n % 4 == 0, "adam: arena length %lld not a multiple of 4", n);
  TFD_REQUIRE(num_pieces <= kMaxPieces, "adam: too many pieces (%d)", num_pieces);
  TFD_REQUIRE(num_slots >= 1, "adam: need at least one gradient slot");
  PieceTable pt;
  pt.n = num_pieces;
  for (int i = 0; i < num_pieces; ++i) pt.p[i] = pieces[i];
  AdamArgs a{w, m, v, static_cast<bf16*>(shadow), slots, num_slots, scale, n, lr, beta1, beta2, eps,
             t_counter, wait_counter, version_flag, global_step};
  long long blocks = (n / 4 + 255) / 256;
  if (blocks > 148 * 8) blocks = 148 * 8;
  adam_apply_kernel<<<(unsigned)blocks, 256, 0, ST(stream)>>>(a, pt);
  TFD_LAUNCH_CHECK();
  return 0;
}

/* Multi-segment push: segs[i] = {src, dst, n}; after every segment is written the last block adds 1 to each
 * arrive counter.  (tfd_push_grads is the single-segment convenience form.) */
extern "C" int tfd_push_multi(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                              unsigned* const* arrive, int narrive, float scale, unsigned* block_counter, void* stream) {
  TFD_REQUIRE(nseg >= 1 && nseg <= kMaxSegs && narrive <= kMaxSegs, "push: bad segment count %d/%d", nseg, narrive);
  PushTable t;
  t.nseg = nseg; t.narrive = narrive;
  long long maxn = 0;
  for (int i = 0; i < nseg; ++i) {
    t.seg[i] = PushSeg{static_cast<const float*>(srcs[i]), static_cast<float*>(dsts[i]), ns[i]};
    TFD_REQUIRE(((uintptr_t)srcs[i] & 15) == 0 && ((uintptr_t)dsts[i] & 15) == 0, "push: segment %d not 16B aligned", i);
    if (ns[i] > maxn) maxn = ns[i];
  }
  for (int i = 0; i < narrive; ++i) t.arrive[i] = arrive[i];
  long long bx = (maxn / 4 + 255) / 256;
  if (bx < 1) bx = 1;
  if (bx > 148) bx = 148;
  push_kernel<<<dim3((unsigned)bx, nseg), 256, 0, ST(stream)>>>(t, scale, block_counter);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_push_grads(const float* src, float* dst, long long n, float scale, unsigned* arrive_counter,
                              unsigned* block_counter, void* stream) {
  const void* s[1] = {src};
  void* d[1] = {dst};
  long long ns[1] = {n};
  unsigned* a[1] = {arrive_counter};
  return tfd_push_multi(s, d, ns, 1, a, arrive_counter ? 1 : 0, scale, block_counter, stream);
}

extern "C" int tfd_pull_multi(const void* const* srcs, void* const* dsts, const long long* ns,
                              const unsigned* const* versions, int nseg, unsigned want_version, void* stream) {
  TFD_REQUIRE(nseg >= 1 && nseg <= kMaxSegs, "pull: bad segment count %d", nseg);
  PullTable t;
  t.nseg = nseg;
  long long maxn = 0;
  for (int i = 0; i < nseg; ++i) {
    t.seg[i] = PullSeg{static_cast<const float*>(srcs[i]), static_cast<float*>(dsts[i]), ns[i], versions ? versions[i] : nullptr};
    if (ns[i] > maxn) maxn = ns[i];
  }
  long long bx = (maxn / 4 + 255) / 256;
  if (bx < 1) bx = 1;
  if (bx > 64) bx = 64;
  pull_kernel<<<dim3((unsigned)bx, nseg), 256, 0, ST(stream)>>>(t, want_version);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_pull_params(const float* src, float* dst, long long n, const unsigned* version_flag, unsigned version,
                               void* stream) {
  const void* s[1] = {src};
  void* d[1] = {dst};
  long long ns[1] = {n};
  const unsigned* v[1] = {version_flag};
  return tfd_pull_multi(s, d, ns, v, 1, version, stream);
}

// ---- symmetric heap over cudaIpc ----
extern "C" int tfd_ipc_alloc(size_t bytes, void** dev_ptr, void* handle_out64) {
  TFD_CUDA_CHECK(cudaMalloc(dev_ptr, bytes));
  TFD_CUDA_CHECK(cudaMemset(*dev_ptr, 0, bytes));
  cudaIpcMemHandle_t h;
  TFD_CUDA_CHECK(cudaIpcGetMemHandle(&h, *dev_ptr));
  static_assert(sizeof(h) == 64, "cudaIpcMemHandle_t is 64 bytes");
  memcpy(handle_out64, &h, 64);
  return 0;
}
extern "C" int tfd_ipc_open(const void* handle64, void** dev_ptr) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64, 64);
  TFD_CUDA_CHECK(cudaIpcOpenMemHandle(dev_ptr, h, cudaIpcMemLazyEnablePeerAccess));
  return 0;
}
extern "C" int tfd_ipc_close(void* dev_ptr) {
  TFD_CUDA_CHECK(cudaIpcCloseMemHandle(dev_ptr));
  return 0;
}
extern "C" int tfd_free(void* dev_ptr) {
  TFD_CUDA_CHECK(cudaFree(dev_ptr));
  return 0;
}
extern "C" int tfd_enable_peer_access(int device, int peer) {
  int can = 0;
  TFD_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, device, peer));
  TFD_REQUIRE(can, "device %d cannot access peer %d", device, peer);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) { cudaGetLastError(); return 0; }
  TFD_CUDA_CHECK(e);
  return 0;
}
