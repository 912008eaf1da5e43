// CUDA-core kernels: casts, Philox init, conv1 (Cin=1, tensor-core hostile), pooling backward, the fused
// output head, and slow reference kernels used only to validate the tcgen05 paths.
//
// Call sites replaced (all /root/reference/mnist_python_m.py): conv1+bias+relu+pool :109-111, dropout :124,
// output matmul :127, softmax cross entropy :205, accuracy :206-207, tf.random_normal :186-195.
#include "../include/tfd_b200.h"
#include "tfd_common.cuh"

namespace {

using namespace tfd;
typedef __nv_bfloat16 bf16;

// ------------------------------------------------------------------------------------------------
// Philox4x32-10
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 c, uint2 k) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const uint32_t hi0 = __umulhi(M0, c.x), lo0 = M0 * c.x;
    const uint32_t hi1 = __umulhi(M1, c.z), lo1 = M1 * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += W0;
    k.y += W1;
  }
  return c;
}
__device__ __forceinline__ float u01(uint32_t x) { return ((float)x + 0.5f) * 2.3283064365386963e-10f; }

// Activation storage type of a precision arm: bf16 (bf16 tensor-core operands) or float (tf32 arm).
__device__ __forceinline__ float to_f(bf16 v) { return __bfloat162float(v); }
__device__ __forceinline__ float to_f(float v) { return v; }
template <typename T> __device__ __forceinline__ T from_f(float v);
template <> __device__ __forceinline__ bf16 from_f<bf16>(float v) { return __float2bfloat16(v); }
template <> __device__ __forceinline__ float from_f<float>(float v) { return v; }

// "Last block" bookkeeping of the fixed-order reductions below: every block adds 1 to a monotonic counter after its
// partial results are visible (fence); the block that takes the launch's last ticket does the final sum.  The counter
// is never reset, so a kernel can sit in a CUDA graph and be replayed.
__device__ __forceinline__ bool last_block_arrives(unsigned* counter, unsigned nblocks) {
  __shared__ unsigned s_last;
  __syncthreads();  // every thread's partial results are ordered before thread 0's fence (cumulativity does the rest:
                    // ONE fence per block -- a fence in every thread made these kernels membar-bound)
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned last = (atomicAdd(counter, 1u) % nblocks == nblocks - 1u) ? 1u : 0u;
    if (last) __threadfence();
    s_last = last;
  }
  __syncthreads();
  return s_last != 0;
}

__global__ void random_normal_kernel(float* __restrict__ out, long long n, unsigned long long seed, float stddev,
                                     long long off4) {
  grid_dep_wait();
  grid_dep_launch();
  const long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x;  // 4 values per thread
  if (i4 * 4 >= n) return;
  const long long c4 = i4 + off4;  // counter = global index of the 4-element group (shard-independent stream)
  uint4 r = philox4x32_10(make_uint4((uint32_t)c4, (uint32_t)(c4 >> 32), 0x5eed, 0),
                          make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
  float z[4];
  const float r0 = sqrtf(-2.f * logf(u01(r.x))), r1 = sqrtf(-2.f * logf(u01(r.z)));
  sincosf(6.283185307179586f * u01(r.y), &z[1], &z[0]);
  sincosf(6.283185307179586f * u01(r.w), &z[3], &z[2]);
  z[0] *= r0; z[1] *= r0; z[2] *= r1; z[3] *= r1;
  for (int j = 0; j < 4; ++j)
    if (i4 * 4 + j < n) out[i4 * 4 + j] = z[j] * stddev;
}

__global__ void cast_bf16_kernel(const float* __restrict__ src, long long ld_src, bf16* __restrict__ dst,
                                 long long ld_dst, int rows, int cols) {
  grid_dep_wait();
  grid_dep_launch();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (long long)rows * ld_dst) return;
  const int r = (int)(i / ld_dst), c = (int)(i % ld_dst);
  dst[i] = __float2bfloat16(c < cols ? src[r * ld_src + c] : 0.f);
}

// ------------------------------------------------------------------------------------------------
// conv1 forward: 5x5 SAME, 1 -> 32, + bias + ReLU + 2x2 max pool.  grid (B, 2): half image per block.
// warp = pooled row, lane = output channel; the 6x6 input window is a shared-memory broadcast.
//
// Gate (ps cluster workers): conv1 is the first kernel of a step.  Its blocks first load their input patch, then one
// thread waits until every parameter shard's version word IN THIS GPU'S OWN MEMORY has reached the worker's step
// counter -- the owner shards store the word (and, before it, the updated small parameters into this worker's mirror)
// over NVLink at the end of their Adam kernel -- so the wait costs local L2 polls, not NVLink round trips, and no
// stand-alone pull kernel sits in front of the step.
// ------------------------------------------------------------------------------------------------
struct GateList {
  const unsigned* flag[TFD_MAX_SEG];
  int n;
  const unsigned* want_ptr;
  unsigned want_mul;  // a flag must reach want_mul * *want_ptr (the multicast transport counts one version per shard)
};

// Phase stamps of the most recent conv1 forward on this GPU (globaltimer ns): [0] block (0,0) starts, [1] its gate
// has passed, [2] it is done.  Read with tfd_debug_conv1_stamps.
__device__ unsigned long long g_conv1_stamps[4];

template <typename T>
__global__ void __launch_bounds__(224) conv1_fwd_kernel(const float* __restrict__ x, const float* w, const float* b,
                                                        T* __restrict__ out, unsigned char* __restrict__ amax,
                                                        const __grid_constant__ GateList gate) {
  grid_dep_wait();
  grid_dep_launch();
  __shared__ __align__(16) float patch[18][32];  // input rows 14*half-2 .. +17, cols -2 .. 29
  const int n = blockIdx.x, half = blockIdx.y;
  const bool stamp = n == 0 && half == 0 && threadIdx.x == 0;
  if (stamp) g_conv1_stamps[0] = globaltimer_ns();
  for (int i = threadIdx.x; i < 18 * 32; i += blockDim.x) {
    const int pr = i >> 5, pc = i & 31;
    const int r = 14 * half - 2 + pr, c = pc - 2;
    patch[pr][pc] = (r >= 0 && r < 28 && c >= 0 && c < 28) ? x[(long long)n * 784 + r * 28 + c] : 0.f;
  }
  const int c = threadIdx.x & 31, prow = threadIdx.x >> 5;  // pooled row inside the half: 0..6
  float wr[25];
  float bias;
  if (gate.n > 0) {
    if ((int)threadIdx.x < gate.n)
      spin_until_ge(gate.flag[threadIdx.x], gate.want_mul * *gate.want_ptr, "conv1: parameter version");
    __syncthreads();
    if (stamp) g_conv1_stamps[1] = globaltimer_ns();
#pragma unroll
    for (int t = 0; t < 25; ++t) wr[t] = __ldcg(w + t * 32 + c);  // the mirror was written by another GPU: read at L2
    bias = __ldcg(b + c);
  } else {
#pragma unroll
    for (int t = 0; t < 25; ++t) wr[t] = w[t * 32 + c];
    bias = b[c];
    __syncthreads();
  }
  const int ph = half * 7 + prow;
  for (int pw = 0; pw < 14; ++pw) {
    float win[6][6];  // 8-byte shared loads: 18 LDS.64 instead of 36 LDS.32 (the kernel is LDS-issue bound)
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      const float2* pr = reinterpret_cast<const float2*>(&patch[2 * prow + i][2 * pw]);
#pragma unroll
      for (int j = 0; j < 3; ++j) {
        const float2 t2 = pr[j];
        win[i][2 * j] = t2.x;
        win[i][2 * j + 1] = t2.y;
      }
    }
    float best = -1.f;
    int bi = 0;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy)
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        float acc = bias;
#pragma unroll
        for (int kh = 0; kh < 5; ++kh)
#pragma unroll
          for (int kw = 0; kw < 5; ++kw) acc = fmaf(win[dy + kh][dx + kw], wr[kh * 5 + kw], acc);
        acc = fmaxf(acc, 0.f);
        if (acc > best) { best = acc; bi = dy * 2 + dx; }
      }
    const long long o = (((long long)n * 14 + ph) * 14 + pw) * 32 + c;
    out[o] = from_f<T>(best);
    amax[o] = (unsigned char)bi;
  }
  if (stamp) g_conv1_stamps[2] = globaltimer_ns();
}


// conv1 filter gradient: pool/ReLU backward folded in (only the arg-max position of each 2x2 window carries
// gradient).  One block of 16 warps per image: warp = pixel group, lane = channel; a thread touches at most 13 of the
// 196 pooled pixels and all its global loads are in flight at once.  512 threads x <= 64 registers and 18 KB of shared
// memory leave room for a conv2 wgrad CTA (192 threads, ~194 KB) on the same SM: the two weight-gradient kernels of
// the backward tail overlap instead of queueing (a 1024-thread shape owned the whole register file).  dpool arrives as nsplit partial planes (the tf32 conv2 dgrad
// splits its K over two CTAs) that are added here in plane order.  Reduction without atomics: warps -> shared memory
// in four fixed passes -> one partial [26][32] per image in `scratch` -> the last block to finish sums the images'
// partials in a fixed order and writes dw / db (which may be staging slots on another GPU).
// scratch = 16 counter words + (B + 15) * 832 floats (tfd_conv1_wgrad_scratch_floats), zeroed once by the caller.
template <typename T>
__global__ void __launch_bounds__(512, 2) conv1_wgrad_kernel(const float* __restrict__ x, const float* __restrict__ dpool,
                                                          int nsplit, long long split_stride, const T* __restrict__ pool,
                                                          const unsigned char* __restrict__ amax, float* __restrict__ dw,
                                                          float* __restrict__ db, float* __restrict__ scratch, int B,
                                                          int gsz) {
  grid_dep_wait();
  grid_dep_launch();
  // row stride 34: a warp holds the 32 channels of ONE pooled pixel, whose argmax picks one of the four positions of
  // its 2x2 window -- with 34 words per row those four land in four different banks (offsets 0, 1, 34, 35), with 33
  // two of them shared a bank and every LDS of the tap loop was a two-way conflict
  __shared__ float img[32][34];
  __shared__ __align__(16) float red[4][26][32];  // 13 KB: small enough to sit next to a conv2 wgrad CTA (~194 KB)
  const int n = blockIdx.x;
  const int c = threadIdx.x & 31, grp = threadIdx.x >> 5;  // 16 pixel groups
  // the image first (two loads per thread), then this thread's pixels: grp, grp+16, ...  ALL loads are unconditional
  // and independent of each other (clamped index, second plane through a pointer that falls back to the first with
  // weight 0), so the compiler issues the 54 loads back to back; with the loads under `if (pix < 196)` / a dynamic
  // plane loop they went out one round trip at a time and the kernel spent 13 serial L2 latencies.
  float xv[2];
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int i = threadIdx.x + k * 512;
    const int r = (i >> 5) - 2, cc = (i & 31) - 2;
    const bool in = r >= 0 && r < 28 && cc >= 0 && cc < 28;
    xv[k] = __ldg(x + (long long)n * 784 + (in ? r * 28 + cc : 0));
    if (!in) xv[k] = 0.f;
  }
  constexpr int kPix = 13;  // ceil(196 / 16)
  const float* plane1 = nsplit > 1 ? dpool + split_stride : dpool;
  const float w1 = nsplit > 1 ? 1.f : 0.f;
  float g0[kPix], g1[kPix], pvq[kPix];
  int aq[kPix];
#pragma unroll
  for (int it = 0; it < kPix; ++it) {
    const int pix = grp + it * 16;
    const long long idx = ((long long)n * 196 + (pix < 196 ? pix : 0)) * 32 + c;
    g0[it] = __ldg(dpool + idx);
    g1[it] = __ldg(plane1 + idx);
    pvq[it] = to_f(pool[idx]);
    aq[it] = amax[idx];
  }
#pragma unroll
  for (int k = 0; k < 2; ++k) {  // 32 x 32 zero-padded image
    const int i = threadIdx.x + k * 512;
    img[i >> 5][i & 31] = xv[k];
  }
  float gq[kPix];
#pragma unroll
  for (int it = 0; it < kPix; ++it) {
    const int pix = grp + it * 16;
    const float g = g0[it] + w1 * g1[it];
    gq[it] = (pix < 196 && pvq[it] > 0.f) ? g : 0.f;
  }
  __syncthreads();
  float acc[25];
#pragma unroll
  for (int t = 0; t < 25; ++t) acc[t] = 0.f;
  float accb = 0.f;
#pragma unroll
  for (int it = 0; it < kPix; ++it) {
    const int pix = grp + it * 16;
    if (pix < 196) {
      const int ph = pix / 14, pw = pix % 14;
      const float g = gq[it];
      const int h = 2 * ph + (aq[it] >> 1), wq = 2 * pw + (aq[it] & 1);
      accb += g;
#pragma unroll
      for (int kh = 0; kh < 5; ++kh)
#pragma unroll
        for (int kw = 0; kw < 5; ++kw) acc[kh * 5 + kw] = fmaf(g, img[h + kh][wq + kw], acc[kh * 5 + kw]);
    }
  }
  // 16 warps -> 4 rows of `red` in four passes (pass k: warps 4k..4k+3), a fixed order
#pragma unroll 1
  for (int pass = 0; pass < 4; ++pass) {
    if ((grp >> 2) == pass) {
      const int row = grp & 3;
      if (pass == 0) {
#pragma unroll
        for (int t = 0; t < 25; ++t) red[row][t][c] = acc[t];
        red[row][25][c] = accb;
      } else {
#pragma unroll
        for (int t = 0; t < 25; ++t) red[row][t][c] += acc[t];
        red[row][25][c] += accb;
      }
    }
    __syncthreads();
  }
  // Two-level sum in a fixed order.  Level 1: the last block of every group of `gsz` images adds the group's partials
  // (all of a thread's loads in flight at once); level 2: the last group to finish adds the <= 15 group sums.  One
  // block summing all B partials needed four dependent rounds of loads (~4 us at the tail of the step).
  unsigned* counters = reinterpret_cast<unsigned*>(scratch);  // [0] launch-wide, [1 + g] group g
  float* parts = scratch + 16;                                  // [B][832]
  float4* gparts4 = reinterpret_cast<float4*>(parts + (size_t)gridDim.x * 832);  // [ngroups][208]
  const int ngroups = ((int)gridDim.x + gsz - 1) / gsz;
  const int grpi = (int)blockIdx.x / gsz, g_first = grpi * gsz;
  const int g_n = min(gsz, (int)gridDim.x - g_first);
  for (int i = threadIdx.x; i < 832; i += 512) {
    const int t = i >> 5, cc = i & 31;
    float s = 0.f;
#pragma unroll
    for (int g4 = 0; g4 < 4; ++g4) s += red[g4][t][cc];
    __stcg(parts + (size_t)blockIdx.x * 832 + i, s);
  }
  if (!last_block_arrives(counters + 1 + grpi, (unsigned)g_n)) return;
  float4* red4 = reinterpret_cast<float4*>(&red[0][0][0]);  // [2][208]
  const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
  if (threadIdx.x < 416) {  // 208 float4 outputs x 2 interleaved halves of the group
    const int o4 = threadIdx.x % 208, half = threadIdx.x / 208;
    const float4* p4 = reinterpret_cast<const float4*>(parts) + (size_t)g_first * 208 + o4;
    float4 s = zero4;
    for (int bk = half; bk < g_n; bk += 16) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = (bk + 2 * u < g_n) ? __ldcg(p4 + (size_t)(bk + 2 * u) * 208) : zero4;
#pragma unroll
      for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
    red4[half * 208 + o4] = s;
  }
  __syncthreads();
  if (threadIdx.x < 208) {
    const float4 a = red4[threadIdx.x], b = red4[208 + threadIdx.x];
    __stcg(gparts4 + (size_t)grpi * 208 + threadIdx.x, make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w));
  }
  if (!last_block_arrives(counters, (unsigned)ngroups)) return;
  if (threadIdx.x < 208) {
    float4 v[15];
#pragma unroll
    for (int g = 0; g < 15; ++g) v[g] = g < ngroups ? __ldcg(gparts4 + (size_t)g * 208 + threadIdx.x) : zero4;
    float4 s = zero4;
#pragma unroll
    for (int g = 0; g < 15; ++g) { s.x += v[g].x; s.y += v[g].y; s.z += v[g].z; s.w += v[g].w; }
    const int i = threadIdx.x * 4;
    if (i < 800) *reinterpret_cast<float4*>(dw + i) = s;
    else *reinterpret_cast<float4*>(db + (i - 800)) = s;
  }
}

// pool2 + ReLU backward: dpool fp32 [nsplit][B,7,7,64] (split-K partials of the fc1 dgrad, summed here)
// -> dy [B,14,14,64] (activation type T); db[64] = sum.  Block = 64 channels x 4 pixel lanes, 16 pooled pixels per
// block.  The bias gradient goes through per-block partials in `scratch` (16 counter words + nblocks * 64 floats,
// zeroed once) and a last-block sum in block order: no atomics.
template <typename T>
__global__ void __launch_bounds__(256) pool2_bwd_kernel(const float* __restrict__ dpool, int nsplit, long long split_stride,
                                                        const T* __restrict__ pool, const unsigned char* __restrict__ amax,
                                                        T* __restrict__ dy, float* __restrict__ db,
                                                        float* __restrict__ scratch, long long npix) {
  grid_dep_wait();
  grid_dep_launch();
  __shared__ float red[256];
  const int c = threadIdx.x & 63, pl = threadIdx.x >> 6;
  float gsum = 0.f;
  float g[4];
  float pv[4];
  int am[4];
  long long pp[4];
  // up to 4 partial planes, every load unconditional (clamped pixel, absent planes alias plane 0 with weight 0) so
  // that all 24 loads of a thread are in flight at once
  const float* pl_ptr[4];
  float pl_w[4];
#pragma unroll
  for (int sp = 0; sp < 4; ++sp) {
    pl_ptr[sp] = sp < nsplit ? dpool + sp * split_stride : dpool;
    pl_w[sp] = sp < nsplit ? 1.f : 0.f;
  }
  float gp[4][4];
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    pp[it] = (long long)blockIdx.x * 16 + it * 4 + pl;
    const long long idx = (pp[it] < npix ? pp[it] : 0) * 64 + c;
#pragma unroll
    for (int sp = 0; sp < 4; ++sp) gp[it][sp] = __ldg(pl_ptr[sp] + idx);
    pv[it] = to_f(pool[idx]);
    am[it] = amax[idx];
  }
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    g[it] = 0.f;
#pragma unroll
    for (int sp = 0; sp < 4; ++sp) g[it] += pl_w[sp] * gp[it][sp];  // plane order: deterministic
  }
#pragma unroll
  for (int it = 0; it < 4; ++it) {
    if (pp[it] >= npix) continue;
    const long long p = pp[it];
    const float gg = pv[it] > 0.f ? g[it] : 0.f;
    const int pw = (int)(p % 7), ph = (int)((p / 7) % 7);
    const long long n = p / 49;
    const T gb = from_f<T>(gg), z = from_f<T>(0.f);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const long long o = (((n * 14 + 2 * ph + (q >> 1)) * 14) + 2 * pw + (q & 1)) * 64 + c;
      dy[o] = (q == am[it]) ? gb : z;
    }
    gsum += gg;
  }
  red[threadIdx.x] = gsum;
  __syncthreads();
  unsigned* counter = reinterpret_cast<unsigned*>(scratch);
  float* parts = scratch + 16;
  if (threadIdx.x < 64)
    __stcg(parts + (size_t)blockIdx.x * 64 + threadIdx.x,
           red[threadIdx.x] + red[threadIdx.x + 64] + red[threadIdx.x + 128] + red[threadIdx.x + 192]);
  if (!last_block_arrives(counter, gridDim.x)) return;
  // 16 float4 channel groups x 16 interleaved groups of blocks: each thread sums every 16th block's partial in block
  // order with 8 loads in flight; the 16 group sums are then added in group order
  __shared__ float4 red4[16][16];
  {
    const int c4 = threadIdx.x & 15, bg = threadIdx.x >> 4;
    const float4* p4 = reinterpret_cast<const float4*>(parts) + c4;
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
    int bk = bg;
    for (; bk + 112 < (int)gridDim.x; bk += 128) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) v[u] = __ldcg(p4 + (size_t)(bk + 16 * u) * 16);
#pragma unroll
      for (int u = 0; u < 8; ++u) { s.x += v[u].x; s.y += v[u].y; s.z += v[u].z; s.w += v[u].w; }
    }
    for (; bk < (int)gridDim.x; bk += 16) {
      const float4 v = __ldcg(p4 + (size_t)bk * 16);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    red4[bg][c4] = s;
  }
  __syncthreads();
  if (threadIdx.x < 16) {
    float4 s = red4[0][threadIdx.x];
#pragma unroll
    for (int bg = 1; bg < 16; ++bg) {
      const float4 v = red4[bg][threadIdx.x];
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
    }
    *reinterpret_cast<float4*>(db + threadIdx.x * 4) = s;
  }
}


// ------------------------------------------------------------------------------------------------
// head: rows phase
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
head_rows_kernel(const float* __restrict__ pre, int pre_nsplit, long long pre_split_stride,
                 const T* __restrict__ hid_in, const float* __restrict__ hid_bias,
                 const float* __restrict__ out_w, const float* __restrict__ out_b, const float* __restrict__ labels,
                 int B, int H, long long ld_h, float keep_prob, unsigned long long seed, unsigned long long step_base,
                 const unsigned* __restrict__ step_ptr, int training, T* __restrict__ hid_out, T* __restrict__ dhid, float* __restrict__ dlogits,
                 float* __restrict__ loss_row, unsigned char* __restrict__ hit, unsigned char* __restrict__ keep_out) {
  grid_dep_wait();
  grid_dep_launch();
  extern __shared__ float sh[];  // H hidden values, then 8*10 partials, then 10 dlogits
  float* part = sh + H;
  float* dl = part + 80;
  const int r = blockIdx.x, tid = threadIdx.x;
  const unsigned long long step = step_base + (step_ptr ? (unsigned long long)*step_ptr : 0ull);
  const float inv_keep = keep_prob < 1.f ? 1.f / keep_prob : 1.f;
  for (int k = tid; k < H; k += 256) {
    float a;
    if (pre) {
      float pv[8];
#pragma unroll
      for (int sp = 0; sp < 8; ++sp) pv[sp] = sp < pre_nsplit ? pre[sp * pre_split_stride + (long long)r * H + k] : 0.f;
      a = hid_bias[k];
#pragma unroll
      for (int sp = 0; sp < 8; ++sp) a += pv[sp];
      for (int sp = 8; sp < pre_nsplit; ++sp) a += pre[sp * pre_split_stride + (long long)r * H + k];
      a = fmaxf(a, 0.f);
      if (keep_prob < 1.f) {
        const unsigned long long e = (unsigned long long)r * H + k;
        const uint4 rn = philox4x32_10(make_uint4((uint32_t)e, (uint32_t)(e >> 32), (uint32_t)step, (uint32_t)(step >> 32)),
                                       make_uint2((uint32_t)seed, (uint32_t)(seed >> 32) ^ 0xd509u));
        const bool keep = u01(rn.x) < keep_prob;
        a = keep ? a * inv_keep : 0.f;
        if (keep_out) keep_out[(long long)r * H + k] = keep;
      }
      // the value every consumer sees is the stored one (bf16-rounded on the bf16 arm)
      const T ab = from_f<T>(a);
      hid_out[(long long)r * ld_h + k] = ab;
      a = to_f(ab);
    } else {
      a = to_f(hid_in[(long long)r * ld_h + k]);
    }
    sh[k] = a;
  }
  __syncthreads();
  float p[10];
#pragma unroll
  for (int j = 0; j < 10; ++j) p[j] = 0.f;
  for (int k = tid; k < H; k += 256) {
    const float a = sh[k];
    const float* wr = out_w + (long long)k * 10;
#pragma unroll
    for (int j = 0; j < 10; ++j) p[j] = fmaf(a, __ldg(wr + j), p[j]);
  }
#pragma unroll
  for (int j = 0; j < 10; ++j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) p[j] += __shfl_xor_sync(0xffffffffu, p[j], o);
  }
  if ((tid & 31) == 0) {
#pragma unroll
    for (int j = 0; j < 10; ++j) part[(tid >> 5) * 10 + j] = p[j];
  }
  __syncthreads();
  if (tid == 0) {
    float lg[10], mx = -INFINITY;
    int am = 0, al = 0;
    float lmax = -INFINITY;
    for (int j = 0; j < 10; ++j) {
      float s = out_b[j];
      for (int wv = 0; wv < 8; ++wv) s += part[wv * 10 + j];
      lg[j] = s;
      if (s > mx) { mx = s; am = j; }
      const float y = labels[(long long)r * 10 + j];
      if (y > lmax) { lmax = y; al = j; }
    }
    float se = 0.f;
    for (int j = 0; j < 10; ++j) se += expf(lg[j] - mx);
    const float lse = mx + logf(se);
    float loss = 0.f, ysum = 0.f;
    for (int j = 0; j < 10; ++j) {
      const float y = labels[(long long)r * 10 + j];
      loss += y * (lse - lg[j]);
      ysum += y;
    }
    for (int j = 0; j < 10; ++j) {
      const float y = labels[(long long)r * 10 + j];
      // d/dlogit of mean_r sum_j -y_j log softmax_j  (dense labels: sum(y) need not be 1)
      const float d = (expf(lg[j] - lse) * ysum - y) / (float)B;
      dl[j] = d;
      dlogits[(long long)r * 10 + j] = d;
    }
    loss_row[r] = loss;
    hit[r] = (am == al);
  }
  if (!training) return;
  __syncthreads();
  float d[10];
#pragma unroll
  for (int j = 0; j < 10; ++j) d[j] = dl[j];
  for (int k = tid; k < H; k += 256) {
    const float* wr = out_w + (long long)k * 10;
    float g = 0.f;
#pragma unroll
    for (int j = 0; j < 10; ++j) g = fmaf(d[j], __ldg(wr + j), g);
    g = (sh[k] > 0.f) ? g * (pre ? inv_keep : 1.f) : 0.f;  // ReLU' and the dropout mask (kept <=> sh>0 or relu==0)
    dhid[(long long)r * ld_h + k] = from_f<T>(g);
  }
}

// head: columns phase.  32 columns per block: lane = column, warp = row group; the 8 row groups are reduced
// through shared memory in a fixed order (deterministic, no atomics).
template <typename T>
__global__ void __launch_bounds__(256)
head_cols_kernel(const T* __restrict__ hid, const T* __restrict__ dhid, long long ld_h,
                 const float* __restrict__ dlogits, const float* __restrict__ loss_row,
                 const unsigned char* __restrict__ hit, int B, int H, float* __restrict__ d_out_w,
                 float* __restrict__ d_hid_bias, float* __restrict__ d_out_b, float* __restrict__ scalars) {
  grid_dep_wait();
  grid_dep_launch();
  extern __shared__ float dls[];  // [B][10], then [8][11][32] partials
  float* part = dls + B * 10;
  const int tid = threadIdx.x, lane = tid & 31, wv = tid >> 5;
  for (int i = tid; i < B * 10; i += 256) dls[i] = dlogits[i];
  __syncthreads();
  const int k = blockIdx.x * 32 + lane;
  if (hid != nullptr) {
    float acc[10];
#pragma unroll
    for (int j = 0; j < 10; ++j) acc[j] = 0.f;
    float sb = 0.f;
    if (k < H) {
      for (int r = wv; r < B; r += 8) {
        const float hv = to_f(hid[(long long)r * ld_h + k]);
        sb += to_f(dhid[(long long)r * ld_h + k]);
#pragma unroll
        for (int j = 0; j < 10; ++j) acc[j] = fmaf(hv, dls[r * 10 + j], acc[j]);
      }
    }
#pragma unroll
    for (int j = 0; j < 10; ++j) part[(wv * 11 + j) * 32 + lane] = acc[j];
    part[(wv * 11 + 10) * 32 + lane] = sb;
    __syncthreads();
    for (int i = tid; i < 11 * 32; i += 256) {
      const int j = i >> 5, c = i & 31, kk = blockIdx.x * 32 + c;
      float sum = 0.f;
#pragma unroll
      for (int g = 0; g < 8; ++g) sum += part[(g * 11 + j) * 32 + c];
      if (kk < H) {
        if (j < 10) d_out_w[(long long)kk * 10 + j] = sum;
        else d_hid_bias[kk] = sum;
      }
    }
  }
  if (blockIdx.x == 0) {
    if (tid < 10 && d_out_b) {
      float s = 0.f;
      for (int r = 0; r < B; ++r) s += dls[r * 10 + tid];
      d_out_b[tid] = s;
    }
    if (wv == 1) {
      float s = 0.f;
      for (int r = lane; r < B; r += 32) s += loss_row[r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) scalars[0] = s / (float)B;
    }
    if (wv == 2) {
      float s = 0.f;
      for (int r = lane; r < B; r += 32) s += (float)hit[r];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) scalars[1] = s / (float)B;
    }
  }
}

template <typename T>
__global__ void colsum_kernel(const T* __restrict__ x, long long ld, int rows, int cols, float* __restrict__ out) {
  grid_dep_wait();
  grid_dep_launch();
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float s = 0.f;
  for (int r = 0; r < rows; ++r) s += to_f(x[(long long)r * ld + c]);
  out[c] = s;
}

// ------------------------------------------------------------------------------------------------
// reference kernels (validation only)
// ------------------------------------------------------------------------------------------------
__global__ void gemm_ref_kernel(const bf16* A, long long lda, const bf16* Bm, long long ldb, int M, int N, int K,
                                float* out, long long ld_out) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y;
  if (n >= N || m >= M) return;
  float s = 0.f;
  for (int k = 0; k < K; ++k) s = fmaf(__bfloat162float(A[m * lda + k]), __bfloat162float(Bm[k * ldb + n]), s);
  out[m * ld_out + n] = s;
}

__global__ void conv2_fwd_ref_kernel(const bf16* __restrict__ in, const bf16* __restrict__ w, const float* __restrict__ b,
                                     bf16* __restrict__ out, unsigned char* __restrict__ amax, long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int co = (int)(idx & 63);
  const long long p = idx >> 6;
  const int pw = (int)(p % 7), ph = (int)((p / 7) % 7);
  const long long n = p / 49;
  float best = -1.f;
  int bi = 0;
  for (int q = 0; q < 4; ++q) {
    const int h = 2 * ph + (q >> 1), wq = 2 * pw + (q & 1);
    float acc = b[co];
    for (int kh = 0; kh < 5; ++kh) {
      const int ih = h + kh - 2;
      if (ih < 0 || ih >= 14) continue;
      for (int kw = 0; kw < 5; ++kw) {
        const int iw = wq + kw - 2;
        if (iw < 0 || iw >= 14) continue;
        const bf16* ip = in + ((n * 14 + ih) * 14 + iw) * 32;
        const bf16* wp = w + ((kh * 5 + kw) * 32) * 64 + co;
        for (int ci = 0; ci < 32; ++ci) acc = fmaf(__bfloat162float(ip[ci]), __bfloat162float(wp[ci * 64]), acc);
      }
    }
    acc = fmaxf(acc, 0.f);
    if (acc > best) { best = acc; bi = q; }
  }
  out[idx] = __float2bfloat16(best);
  amax[idx] = (unsigned char)bi;
}

__global__ void conv2_dgrad_ref_kernel(const bf16* __restrict__ dy, const bf16* __restrict__ w, float* __restrict__ dx,
                                       long long total) {
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= total) return;
  const int ci = (int)(idx & 31);
  const long long p = idx >> 5;
  const int wq = (int)(p % 14), h = (int)((p / 14) % 14);
  const long long n = p / 196;
  float acc = 0.f;
  for (int kh = 0; kh < 5; ++kh) {
    const int oh = h + 2 - kh;
    if (oh < 0 || oh >= 14) continue;
    for (int kw = 0; kw < 5; ++kw) {
      const int ow = wq + 2 - kw;
      if (ow < 0 || ow >= 14) continue;
      const bf16* gp = dy + ((n * 14 + oh) * 14 + ow) * 64;
      const bf16* wp = w + ((kh * 5 + kw) * 32 + ci) * 64;
      for (int co = 0; co < 64; ++co) acc = fmaf(__bfloat162float(gp[co]), __bfloat162float(wp[co]), acc);
    }
  }
  dx[idx] = acc;
}

// grid (800*64/256, chunks): each block sums a chunk of images, then atomics
__global__ void conv2_wgrad_ref_kernel(const bf16* __restrict__ in, const bf16* __restrict__ dy, float* __restrict__ dw,
                                       int B, int imgs_per_chunk) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;  // (tap*32+ci)*64+co
  const int co = o & 63, ci = (o >> 6) & 31, tap = o >> 11;
  const int kh = tap / 5, kw = tap % 5;
  const int n0 = blockIdx.y * imgs_per_chunk, n1 = min(B, n0 + imgs_per_chunk);
  float acc = 0.f;
  for (int n = n0; n < n1; ++n)
    for (int h = 0; h < 14; ++h) {
      const int ih = h + kh - 2;
      if (ih < 0 || ih >= 14) continue;
      for (int wq = 0; wq < 14; ++wq) {
        const int iw = wq + kw - 2;
        if (iw < 0 || iw >= 14) continue;
        acc = fmaf(__bfloat162float(in[(((long long)n * 14 + ih) * 14 + iw) * 32 + ci]),
                   __bfloat162float(dy[(((long long)n * 14 + h) * 14 + wq) * 64 + co]), acc);
      }
    }
  atomicAdd(dw + o, acc);
}

struct ZeroTable {
  float* p[16];
  long long n[16];
  int cnt;
};
__global__ void zero_multi_kernel(const __grid_constant__ ZeroTable t) {
  grid_dep_wait();
  grid_dep_launch();
  float* p = t.p[blockIdx.y];
  const long long n = t.n[blockIdx.y], n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
    reinterpret_cast<float4*>(p)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    p[i] = 0.f;
}

}  // namespace

extern "C" int tfd_debug_conv1_stamps(unsigned long long* out4) {
  TFD_CUDA_CHECK(cudaMemcpyFromSymbol(out4, g_conv1_stamps, sizeof(g_conv1_stamps)));
  return 0;
}

#define ST(s) static_cast<cudaStream_t>(s)

extern "C" int tfd_random_normal(float* out, long long n, unsigned long long seed, float stddev, long long elem_offset,
                                 void* stream) {
  TFD_REQUIRE(elem_offset % 4 == 0, "random_normal: element offset %lld not a multiple of 4", elem_offset);
  const long long n4 = (n + 3) / 4;
  TFD_CUDA_CHECK(tfd_launch(random_normal_kernel, dim3((unsigned)((n4 + 255) / 256)), dim3(256), 0, ST(stream), out, n, seed, stddev, elem_offset / 4));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_cast_bf16(const float* src, long long ld_src, void* dst, long long ld_dst, int rows, int cols,
                             void* stream) {
  const long long total = (long long)rows * ld_dst;
  TFD_CUDA_CHECK(tfd_launch(cast_bf16_kernel, dim3((unsigned)((total + 255) / 256)), dim3(256), 0, ST(stream), src, ld_src, (bf16*)dst, ld_dst, rows, cols));
  TFD_LAUNCH_CHECK();
  return 0;
}

#define TFD_ACT_CHECK(name) \
  TFD_REQUIRE(dtype == TFD_DTYPE_BF16 || dtype == TFD_DTYPE_TF32, name ": unknown activation type %d", dtype)

extern "C" int tfd_conv1_fwd(int dtype, const float* x, const float* w, const float* b, void* pool_out,
                             unsigned char* argmax, int B, const unsigned* const* gate_flags, int ngate,
                             const unsigned* want_ptr, unsigned want_mul, void* stream) {
  TFD_ACT_CHECK("conv1_fwd");
  TFD_REQUIRE(ngate >= 0 && ngate <= TFD_MAX_SEG && (ngate == 0 || (gate_flags && want_ptr)),
              "conv1_fwd: bad gate list (%d flags)", ngate);
  GateList gl;
  memset(&gl, 0, sizeof(gl));
  gl.n = ngate;
  gl.want_ptr = want_ptr;
  gl.want_mul = want_mul ? want_mul : 1u;
  for (int i = 0; i < ngate; ++i) gl.flag[i] = gate_flags[i];
  if (dtype == TFD_DTYPE_TF32)
    TFD_CUDA_CHECK(tfd_launch(conv1_fwd_kernel<float>, dim3(B, 2), dim3(224), 0, ST(stream), x, w, b, (float*)pool_out, argmax, gl));
  else
    TFD_CUDA_CHECK(tfd_launch(conv1_fwd_kernel<bf16>, dim3(B, 2), dim3(224), 0, ST(stream), x, w, b, (bf16*)pool_out, argmax, gl));
  TFD_LAUNCH_CHECK();
  return 0;
}

// 16 counter words, one [26][32] partial per image, up to 15 group sums
extern "C" long long tfd_conv1_wgrad_scratch_floats(int B) { return 16 + (long long)B * 832 + 15 * 832; }
extern "C" long long tfd_pool2_bwd_scratch_floats(int B) { return 16 + (((long long)B * 49 + 15) / 16) * 64; }

extern "C" int tfd_conv1_wgrad(int dtype, const float* x, const float* dpool, int nsplit, long long split_stride,
                               const void* pool_out, const unsigned char* argmax, float* dw, float* db, float* scratch,
                               int B, void* stream) {
  TFD_ACT_CHECK("conv1_wgrad");
  TFD_REQUIRE(scratch != nullptr && nsplit >= 1 && nsplit <= 2, "conv1_wgrad: scratch missing or nsplit %d not 1 or 2", nsplit);
  const dim3 grid(B);
  const int gsz = B <= 240 ? 16 : (B + 14) / 15;  // images per first-level group of the final sum (at most 15 groups)
  // runs next to the conv2 wgrad CTAs (one per SM, ~194 KB of shared memory): ask for the same carve-out
  tfd_prefer_max_carveout(reinterpret_cast<const void*>(conv1_wgrad_kernel<float>));
  tfd_prefer_max_carveout(reinterpret_cast<const void*>(conv1_wgrad_kernel<bf16>));
  if (dtype == TFD_DTYPE_TF32)
    TFD_CUDA_CHECK(tfd_launch(conv1_wgrad_kernel<float>, grid, dim3(512), 0, ST(stream), x, dpool, nsplit, split_stride,
                              (const float*)pool_out, argmax, dw, db, scratch, B, gsz));
  else
    TFD_CUDA_CHECK(tfd_launch(conv1_wgrad_kernel<bf16>, grid, dim3(512), 0, ST(stream), x, dpool, nsplit, split_stride,
                              (const bf16*)pool_out, argmax, dw, db, scratch, B, gsz));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_pool2_bwd(int dtype, const float* dpool, int nsplit, long long split_stride, const void* pool_out,
                             const unsigned char* argmax, void* dy, float* db, float* scratch, int B, void* stream) {
  TFD_ACT_CHECK("pool2_bwd");
  TFD_REQUIRE(scratch != nullptr && nsplit <= 4, "pool2_bwd: scratch missing or more than 4 partial planes (%d)", nsplit);
  const long long npix = (long long)B * 49;
  const dim3 grid((unsigned)((npix + 15) / 16));
  if (dtype == TFD_DTYPE_TF32)
    TFD_CUDA_CHECK(tfd_launch(pool2_bwd_kernel<float>, grid, dim3(256), 0, ST(stream), dpool, nsplit < 1 ? 1 : nsplit,
                              split_stride, (const float*)pool_out, argmax, (float*)dy, db, scratch, npix));
  else
    TFD_CUDA_CHECK(tfd_launch(pool2_bwd_kernel<bf16>, grid, dim3(256), 0, ST(stream), dpool, nsplit < 1 ? 1 : nsplit,
                              split_stride, (const bf16*)pool_out, argmax, (bf16*)dy, db, scratch, npix));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_head_rows(int dtype, const float* pre_f32, int pre_nsplit, long long pre_split_stride, const void* hid_in,
                             const float* hid_bias, const float* out_w, const float* out_b, const float* labels, int B,
                             int H, long long ld_h, float keep_prob, unsigned long long seed, unsigned long long step,
                             const unsigned* step_ptr, int training, void* hid_out, void* dhid, float* dlogits,
                             float* loss_row, unsigned char* hit, unsigned char* keep_mask_out, void* stream) {
  TFD_ACT_CHECK("head_rows");
  TFD_REQUIRE((pre_f32 != nullptr) != (hid_in != nullptr), "head_rows: exactly one of pre_f32 / hid_in");
  TFD_REQUIRE(!pre_f32 || (hid_bias && hid_out), "head_rows: pre_f32 mode needs hid_bias and hid_out");
  TFD_REQUIRE(!training || dhid, "head_rows: training needs dhid");
  const size_t smem = (size_t)(H + 96) * sizeof(float);
  TFD_REQUIRE(smem <= 200 * 1024, "head_rows: H=%d too large", H);
  if (dtype == TFD_DTYPE_TF32) {
    if (smem > 48 * 1024)
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(head_rows_kernel<float>), (int)smem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(head_rows_kernel<float>, dim3(B), dim3(256), smem, ST(stream), pre_f32, pre_nsplit < 1 ? 1 : pre_nsplit,
                              pre_split_stride, (const float*)hid_in, hid_bias, out_w, out_b, labels, B, H, ld_h, keep_prob,
                              seed, step, step_ptr, training, (float*)hid_out, (float*)dhid, dlogits, loss_row, hit,
                              keep_mask_out));
  } else {
    if (smem > 48 * 1024)
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(head_rows_kernel<bf16>), (int)smem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(head_rows_kernel<bf16>, dim3(B), dim3(256), smem, ST(stream), pre_f32, pre_nsplit < 1 ? 1 : pre_nsplit,
                              pre_split_stride, (const bf16*)hid_in, hid_bias, out_w, out_b, labels, B, H, ld_h, keep_prob,
                              seed, step, step_ptr, training, (bf16*)hid_out, (bf16*)dhid, dlogits, loss_row, hit,
                              keep_mask_out));
  }
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_head_cols(int dtype, const void* hid, const void* dhid, long long ld_h, const float* dlogits,
                             const float* loss_row, const unsigned char* hit, int B, int H, float* d_out_w,
                             float* d_hid_bias, float* d_out_b, float* scalars, void* stream) {
  TFD_ACT_CHECK("head_cols");
  const size_t smem = ((size_t)B * 10 + 8 * 11 * 32) * sizeof(float);
  TFD_REQUIRE(smem <= 200 * 1024, "head_cols: batch %d too large for the dlogits tile", B);
  const int blocks = hid ? (H + 31) / 32 : 1;
  if (dtype == TFD_DTYPE_TF32) {
    if (smem > 48 * 1024)
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(head_cols_kernel<float>), (int)smem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(head_cols_kernel<float>, dim3(blocks), dim3(256), smem, ST(stream), (const float*)hid,
                              (const float*)dhid, ld_h, dlogits, loss_row, hit, B, H, d_out_w, d_hid_bias, d_out_b, scalars));
  } else {
    if (smem > 48 * 1024)
      if (int arc = tfd_ensure_dyn_smem(reinterpret_cast<const void*>(head_cols_kernel<bf16>), (int)smem)) return arc;
    TFD_CUDA_CHECK(tfd_launch(head_cols_kernel<bf16>, dim3(blocks), dim3(256), smem, ST(stream), (const bf16*)hid,
                              (const bf16*)dhid, ld_h, dlogits, loss_row, hit, B, H, d_out_w, d_hid_bias, d_out_b, scalars));
  }
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_colsum(int dtype, const void* x, long long ld, int rows, int cols, float* out, void* stream) {
  TFD_ACT_CHECK("colsum");
  if (dtype == TFD_DTYPE_TF32)
    TFD_CUDA_CHECK(tfd_launch(colsum_kernel<float>, dim3((cols + 255) / 256), dim3(256), 0, ST(stream), (const float*)x, ld, rows, cols, out));
  else
    TFD_CUDA_CHECK(tfd_launch(colsum_kernel<bf16>, dim3((cols + 255) / 256), dim3(256), 0, ST(stream), (const bf16*)x, ld, rows, cols, out));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_gemm_ref(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, float* out,
                            long long ld_out, void* stream) {
  gemm_ref_kernel<<<dim3((N + 127) / 128, M), 128, 0, ST(stream)>>>((const bf16*)A, lda, (const bf16*)B, ldb, M, N, K, out,
                                                                   ld_out);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_conv2_fwd_ref(const void* in, const void* w, const float* b, void* pool_out, unsigned char* argmax,
                                 int B, void* stream) {
  const long long total = (long long)B * 49 * 64;
  conv2_fwd_ref_kernel<<<(unsigned)((total + 127) / 128), 128, 0, ST(stream)>>>((const bf16*)in, (const bf16*)w, b,
                                                                              (bf16*)pool_out, argmax, total);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_conv2_dgrad_ref(const void* dy, const void* w, float* dx, int B, void* stream) {
  const long long total = (long long)B * 196 * 32;
  conv2_dgrad_ref_kernel<<<(unsigned)((total + 127) / 128), 128, 0, ST(stream)>>>((const bf16*)dy, (const bf16*)w, dx, total);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_conv2_wgrad_ref(const void* in, const void* dy, float* dw, int B, void* stream) {
  const int per = 4;
  conv2_wgrad_ref_kernel<<<dim3(800 * 64 / 256, (B + per - 1) / per), 256, 0, ST(stream)>>>((const bf16*)in, (const bf16*)dy,
                                                                                           dw, B, per);
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_zero_multi(float* const* ptrs, const long long* ns, int count, void* stream) {
  TFD_REQUIRE(count >= 1 && count <= 16, "zero_multi: bad count %d", count);
  ZeroTable t;
  t.cnt = count;
  long long maxn = 0;
  for (int i = 0; i < count; ++i) {
    TFD_REQUIRE(((uintptr_t)ptrs[i] & 15) == 0, "zero_multi: pointer %d not 16B aligned", i);
    t.p[i] = ptrs[i];
    t.n[i] = ns[i];
    if (ns[i] > maxn) maxn = ns[i];
  }
  long long bx = (maxn / 4 + 255) / 256;
  if (bx < 1) bx = 1;
  if (bx > 296) bx = 296;
  TFD_CUDA_CHECK(tfd_launch(zero_multi_kernel, dim3((unsigned)bx, count), dim3(256), 0, ST(stream), t));
  TFD_LAUNCH_CHECK();
  return 0;
}
