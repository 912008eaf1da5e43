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
#include <stdlib.h>

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
  const float* slots[TFD_MAX_SLOTS];  // by value: no dependent pointer load in front of the gradient loads
  int num_slots;
  float scale;
  long long n;
  float lr, beta1, beta2, eps;
  unsigned* t_counter;          // [0] = applied updates so far, [1] = block-done scratch
  const unsigned* wait_counter; // arrivals (local memory, incremented by workers over NVLink)
  unsigned* version_flag;
  unsigned* global_step;
  // async (apply-on-arrival): serve exactly one ticket of the shard's arrival ring
  unsigned* queue;              // ring of qlen words: worker id + 1 once that worker's slot is complete, else 0
  int qlen;
  const unsigned* shutdown;     // non-zero: leave without applying (the ps is being released)
  int slots_bf16;               // gradient slots hold bf16 (halves the push bytes); accumulation stays fp32
  const unsigned* mask_ptr;     // backup workers: bit w set = slot w contributes to this update (decided by decide_kernel)
  // ---- colocated strict-sync step: the rendezvous kernels folded into this one ----
  unsigned* signal_arrive[TFD_MAX_SEG];  // every shard's arrival counter (peer pointers): this rank's worker signals
  int n_signal;                          //   "my gradients are in your slot" from block 0 before anybody waits
  unsigned* local_step;                  // this rank's worker step word, advanced by the last block
};

// Publish table: what the owner shard stores into every worker's parameter mirror (peer memory) right after the
// update, so that a worker's next step reads its small parameters from LOCAL memory and waits on a LOCAL version word.
struct PubPiece {
  long long f32_off;   // byte offset of the variable's fp32 mirror inside a worker heap, or -1
  long long b16_off;   // byte offset of its bf16 mirror (bf16 arm, tensor-core operands), or -1
  int ld16;            // leading dimension of the bf16 mirror (elements)
  int pad;
};
struct PubTable {
  PubPiece pc[kMaxPieces];        // parallel to the PieceTable
  unsigned char* base[TFD_MAX_SLOTS];  // every worker's heap base (peer-mapped; the own rank's is local)
  unsigned* wver[TFD_MAX_SLOTS];       // every worker's LOCAL version word for this shard
  int nworkers;                        // 0: nothing is published (single GPU, legacy pull protocol)
};

// four gradient values at element index i of a staging slot (fp32 or bf16 storage)
__device__ __forceinline__ float4 load_grad4(const float* slot, long long i, int is_bf16) {
  if (!is_bf16) return __ldcs(reinterpret_cast<const float4*>(slot + i));
  const uint2 r = __ldcs(reinterpret_cast<const uint2*>(reinterpret_cast<const bf16*>(slot) + i));
  return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                     __uint_as_float(r.y & 0xffff0000u));
}

// Phase stamps of the most recent Adam kernel on this GPU (globaltimer ns; a handful of stores by one thread):
// [0] block 0 starts, [1] arrive signals issued, [2] all arrivals seen, [3] block 0 leaves the update loop,
// [4] the last block sees every block done, [5] version words released.  Read with tfd_debug_adam_stamps.
__device__ unsigned long long g_adam_stamps[8];

__global__ void __launch_bounds__(256) adam_apply_kernel(AdamArgs a, const __grid_constant__ PieceTable pt,
                                                         const __grid_constant__ PubTable pub) {
  grid_dep_wait();
  grid_dep_launch();
  // Colocated strict sync: this rank's worker half signals "my gradients are in your slots" to every shard from here
  // (the gradient kernels of this step ran earlier in the same stream and stored straight into the owners' slots), so
  // no stand-alone push kernel exists.  One thread per shard: the NVLink round trips of the releases overlap.
  const bool stamp = blockIdx.x == 0 && threadIdx.x == 0;
  if (stamp) g_adam_stamps[0] = globaltimer_ns();
  // (the gradient stores belong to kernels that have completed; the release of the reduction is the only fence)
  if (blockIdx.x == 0 && (int)threadIdx.x < a.n_signal) red_add_release_sys_u32(a.signal_arrive[threadIdx.x], 1u);
  if (stamp) g_adam_stamps[1] = globaltimer_ns();
  // Every thread derives the step index and lr_t itself (one broadcast load + a few SFU ops): no block barrier
  // and no serial prologue in front of the streaming loop.  Only one thread per block waits for the arrivals.
  const unsigned t = a.t_counter[0] + 1;  // this update's 1-based index
  if (a.wait_counter) {
    if (threadIdx.x == 0) spin_until_ge(a.wait_counter, (unsigned)a.num_slots * t, "adam: gradient arrivals");
    __syncthreads();
  }
  if (stamp) g_adam_stamps[2] = globaltimer_ns();
  // backup workers (R of N): only the slots named in the mask are summed, the average is over their count; an empty
  // mask means the ps is being released
  unsigned use_mask = 0xffffffffu;
  float scale = a.scale;
  if (a.mask_ptr) {
    use_mask = *a.mask_ptr;
    if (use_mask == 0) return;
    scale = 1.f / (float)__popc(use_mask);
  }
  const float* serve_slot = nullptr;
  if (a.queue) {
    // apply-on-arrival in ticket order: ticket (t-1) is the next one this shard serves.  Every block reads the
    // same ring entry; the last block clears it after all blocks are done.
    __shared__ unsigned s_who;
    if (threadIdx.x == 0) {
      const unsigned* q = a.queue + ((t - 1) % (unsigned)a.qlen);
      unsigned who = ld_acquire_sys_u32(q);
      unsigned long long t0 = globaltimer_ns();
      unsigned it = 0;
      while (who == 0 && !(a.shutdown && ld_acquire_sys_u32(a.shutdown) != 0)) {
        __nanosleep(128);
        who = ld_acquire_sys_u32(q);
        if ((++it & 0xff) == 0 && globaltimer_ns() - t0 > kSpinTimeoutNs) {
          printf("tfd: async apply waited 20 s for ticket %u\n", t - 1);
          __trap();
        }
      }
      s_who = who;
    }
    __syncthreads();
    if (s_who == 0) return;  // shutdown
    serve_slot = a.slots[s_who - 1];
  }
  // 1 - beta^t through expm1 keeps full relative accuracy for small t (where beta^t is close to 1)
  const float om_b1t = -expm1f((float)t * logf(a.beta1)), om_b2t = -expm1f((float)t * logf(a.beta2));
  const float lr_t = a.lr * sqrtf(om_b2t) / om_b1t;
  // Two float4 groups per thread per iteration, one block-stride apart (every warp instruction touches 512
  // contiguous bytes); all eight streaming loads are issued before the first use (memory-level parallelism).
  const long long n4 = a.n >> 2;
  const long long stride4 = (long long)gridDim.x * blockDim.x;
  int first = 0;
  if (a.mask_ptr) first = __ffs(use_mask) - 1;
  const float* g0 = serve_slot ? serve_slot : a.slots[first];
  const int nslots = serve_slot ? 1 : a.num_slots;
  int cur = 0;  // cached piece index for the shadow lookup
  int stored_remote = 0;  // this thread has written into another GPU's mirror
  for (long long i4 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i4 < n4; i4 += 2 * stride4) {
    const long long j0 = i4, j1 = i4 + stride4;
    long long idx[2] = {j0 << 2, j1 << 2};
    const bool ok0 = true, ok1 = j1 < n4;
    if (!ok1) idx[1] = idx[0];
    float4 w4[2], m4[2], v4[2], g4[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      w4[h] = __ldcs(reinterpret_cast<const float4*>(a.w + idx[h]));
      m4[h] = __ldcs(reinterpret_cast<const float4*>(a.m + idx[h]));
      v4[h] = __ldcs(reinterpret_cast<const float4*>(a.v + idx[h]));
      g4[h] = load_grad4(g0, idx[h], a.slots_bf16);
    }
    for (int r = first + 1; r < nslots; ++r) {
      if (!((use_mask >> r) & 1u)) continue;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const float4 x = load_grad4(a.slots[r], idx[h], a.slots_bf16);
        g4[h].x += x.x; g4[h].y += x.y; g4[h].z += x.z; g4[h].w += x.w;
      }
    }
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      if ((h == 1 && !ok1) || (h == 0 && !ok0)) continue;
      float wv[4] = {w4[h].x, w4[h].y, w4[h].z, w4[h].w}, mv[4] = {m4[h].x, m4[h].y, m4[h].z, m4[h].w};
      float vv[4] = {v4[h].x, v4[h].y, v4[h].z, v4[h].w};
      const float gv[4] = {g4[h].x * scale, g4[h].y * scale, g4[h].z * scale, g4[h].w * scale};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        mv[j] = a.beta1 * mv[j] + (1.f - a.beta1) * gv[j];
        vv[j] = a.beta2 * vv[j] + (1.f - a.beta2) * gv[j] * gv[j];
        // TF form: epsilon outside the bias correction.  MUFU-based sqrt and divide (2 ulp): the IEEE versions made
        // this streaming kernel issue-bound (158 instructions per element).
        wv[j] -= __fdividef(lr_t * mv[j], vv[j] * rsqrtf(fmaxf(vv[j], 1e-38f)) + a.eps);
      }
      const long long i = idx[h];
      __stcs(reinterpret_cast<float4*>(a.w + i), make_float4(wv[0], wv[1], wv[2], wv[3]));
      __stcs(reinterpret_cast<float4*>(a.m + i), make_float4(mv[0], mv[1], mv[2], mv[3]));
      __stcs(reinterpret_cast<float4*>(a.v + i), make_float4(vv[0], vv[1], vv[2], vv[3]));
      if (a.shadow || pub.nworkers > 0) {
        // pieces are 64-element aligned in the arena, so one float4 lies in one piece; a thread's indices grow
        // monotonically, so the piece found last time is almost always still the right one
        if (!(i >= pt.p[cur].arena_off && i < pt.p[cur].arena_off + pt.p[cur].numel)) {
          for (int q = 0; q < pt.n; ++q)
            if (i >= pt.p[q].arena_off && i < pt.p[q].arena_off + pt.p[q].numel) { cur = q; break; }
        }
        const tfd_piece_t& pc = pt.p[cur];
        if (i >= pc.arena_off && i < pc.arena_off + pc.numel) {
          const long long e0 = i - pc.arena_off;
          const uint2 packed = make_uint2(pack_bf16x2(wv[0], wv[1]), pack_bf16x2(wv[2], wv[3]));
          if (a.shadow && pc.ld_shadow > 0) {
            if (pc.ld_shadow == pc.cols) {
              *reinterpret_cast<uint2*>(a.shadow + pc.shadow_off + e0) = packed;
            } else {
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const long long e = e0 + j;
                if (e < pc.numel)
                  a.shadow[pc.shadow_off + (e / pc.cols) * pc.ld_shadow + (e % pc.cols)] = __float2bfloat16(wv[j]);
              }
            }
          }
          // every worker's mirror of a small variable gets the new values now (peer stores; mirrors are padded to
          // whole float4 groups, so the tail of a 10-element bias needs no special case)
          if (pub.nworkers > 0) {
            const PubPiece& pp = pub.pc[cur];
            stored_remote |= (pp.f32_off >= 0) | (pp.b16_off >= 0);
            if (pp.f32_off >= 0) {
              const float4 o4 = make_float4(wv[0], wv[1], wv[2], wv[3]);
              for (int wk = 0; wk < pub.nworkers; ++wk)
                *reinterpret_cast<float4*>(pub.base[wk] + pp.f32_off + 4 * e0) = o4;
            }
            if (pp.b16_off >= 0) {
              if (pp.ld16 == pc.cols) {
                for (int wk = 0; wk < pub.nworkers; ++wk)
                  *reinterpret_cast<uint2*>(pub.base[wk] + pp.b16_off + 2 * e0) = packed;
              } else {
                for (int wk = 0; wk < pub.nworkers; ++wk) {
                  bf16* d = reinterpret_cast<bf16*>(pub.base[wk] + pp.b16_off);
#pragma unroll
                  for (int j = 0; j < 4; ++j) {
                    const long long e = e0 + j;
                    if (e < pc.numel) d[(e / pc.cols) * pp.ld16 + (e % pc.cols)] = __float2bfloat16(wv[j]);
                  }
                }
              }
            }
          }
        }
      }
    }
  }
  // publish: the last block bumps the counters and releases the new parameter version.  One fence per block (the
  // barrier orders every thread's stores before it): system scope only in a block that stored into another GPU's
  // mirror -- it has to see those stores acknowledged over NVLink -- and GPU scope otherwise (arena, moments and
  // shadow live in this GPU's L2, which is also where a peer's loads and copies read them).
  const int any_remote = __syncthreads_or(stored_remote);
  __shared__ unsigned s_last;
  if (stamp) g_adam_stamps[3] = globaltimer_ns();
  if (threadIdx.x == 0) {
    if (any_remote) __threadfence_system(); else __threadfence();
    const unsigned done = atomicAdd(&a.t_counter[1], 1u);
    s_last = (done == gridDim.x - 1) ? 1u : 0u;
    if (s_last) {
      __threadfence();  // acquire side of the counter: the other blocks' fenced stores are ordered before what follows
      g_adam_stamps[4] = globaltimer_ns();
      a.t_counter[1] = 0;
      a.t_counter[0] = t;
      if (a.queue) a.queue[(t - 1) % (unsigned)a.qlen] = 0;
      if (a.global_step) *a.global_step = *a.global_step + 1;
      if (a.local_step) *a.local_step = *a.local_step + 1;
    }
  }
  __syncthreads();
  // Releases by separate threads of the last block, so their fences and NVLink round trips overlap: thread k < nworkers
  // stores worker k's version word, thread nworkers the shard's own.  Each is a release at system scope issued after
  // the barrier behind thread 0's bookkeeping and its observation that every block had fenced and counted itself.
  if (s_last) {
    if ((int)threadIdx.x < pub.nworkers) st_release_sys_u32(pub.wver[threadIdx.x], t);
    if ((int)threadIdx.x == pub.nworkers && a.version_flag) {
      // a shard nobody else waits on (single process: no arrivals, queue or mask) needs no system-scope release --
      // the next kernel in the stream is its only reader, and the release cost ~2 us at the very end of the step
      const bool solo = !a.wait_counter && !a.queue && !a.mask_ptr && pub.nworkers == 0;
      if (solo) *reinterpret_cast<volatile unsigned*>(a.version_flag) = t;
      else st_release_sys_u32(a.version_flag, t);
    }
    if ((int)threadIdx.x == pub.nworkers) g_adam_stamps[5] = globaltimer_ns();
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
  unsigned* arrive[kMaxSegs];   // sync: per-shard arrival counter.  async: per-shard ticket dispenser
  unsigned* queue[kMaxSegs];    // async: per-shard arrival ring
  int nseg, narrive;
  int qlen;                     // > 0 selects the async (ticket) protocol
  unsigned worker;              // this worker's index
  unsigned* my_ticket;          // async: local array [narrive]; receives ticket+1 = the version that serves it
  int dst_bf16;                 // destination slots hold bf16
  const unsigned* tag_ptr;      // backup workers: stamp every shard with *tag_ptr instead of counting arrivals
  unsigned* tag_cur;            //   ... and remember it (what the next pull waits for)
};

// grid (blocks_per_seg, nseg)
__global__ void __launch_bounds__(1024) push_kernel(const __grid_constant__ PushTable t, float scale, unsigned* block_counter,
                                                   unsigned* local_step) {
  grid_dep_wait();
  grid_dep_launch();
  const PushSeg s = t.seg[blockIdx.y];
  const long long n4 = s.n >> 2;
  bf16* dst16 = reinterpret_cast<bf16*>(s.dst);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 x = *reinterpret_cast<const float4*>(s.src + (i << 2));
    x.x *= scale; x.y *= scale; x.z *= scale; x.w *= scale;
    if (t.dst_bf16) *reinterpret_cast<uint2*>(dst16 + (i << 2)) = make_uint2(pack_bf16x2(x.x, x.y), pack_bf16x2(x.z, x.w));
    else *reinterpret_cast<float4*>(s.dst + (i << 2)) = x;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += (long long)gridDim.x * blockDim.x) {
    if (t.dst_bf16) dst16[i] = __float2bfloat16(s.src[i] * scale);
    else s.dst[i] = s.src[i] * scale;
  }
  __threadfence_system();
  __syncthreads();
  // The last block to finish signals every shard.  One thread per shard, so the NVLink round trips of the
  // release operations overlap instead of queueing behind each other (8 shards: ~2 us instead of ~20 us).
  __shared__ unsigned s_last;
  if (threadIdx.x == 0) {
    const unsigned total = gridDim.x * gridDim.y;
    s_last = (atomicAdd(block_counter, 1u) == total - 1) ? 1u : 0u;
    if (s_last) *block_counter = 0;  // (every block fenced at system scope before its increment; the release
                                      //  operations below are cumulative over what this thread has observed)
  }
  __syncthreads();
  if (!s_last) return;
  const int k = threadIdx.x;
  if (k < t.narrive) {
    if (t.qlen > 0) {
      // apply-on-arrival: take a ticket on the shard, then publish "worker w's slot is ready" in that ring cell
      const unsigned tk = atom_add_release_sys_u32(t.arrive[k], 1u);
      st_release_sys_u32(t.queue[k] + (tk % (unsigned)t.qlen), t.worker + 1u);
      t.my_ticket[k] = tk + 1u;
    } else if (t.tag_ptr) {
      st_release_sys_u32(t.arrive[k], *t.tag_ptr);  // "my slot holds a gradient for update <tag>"
    } else {
      red_add_release_sys_u32(t.arrive[k], 1u);
    }
  }
  if (k == 0 && t.tag_ptr) *t.tag_cur = *t.tag_ptr;
  if (k == 0 && local_step) *local_step = *local_step + 1;
}

struct PullSeg {
  const float* src;
  float* dst;
  long long n;
  const unsigned* version;
  const unsigned* want_seg;     // optional: wait for *version >= *want_seg instead of the common value
};
struct PullTable {
  PullSeg seg[kMaxSegs];
  int nseg;
  int tag_seg;                  // backup workers: the block of this segment also derives the next push's tag
  unsigned* tag_out;            //   = (version it observed) + 1
};

__global__ void __launch_bounds__(256) pull_kernel(const __grid_constant__ PullTable t, unsigned want_base,
                                                   const unsigned* want_ptr) {
  grid_dep_wait();
  grid_dep_launch();
  const PullSeg s = t.seg[blockIdx.y];
  const unsigned want = want_base + (want_ptr ? *want_ptr : 0u);
  if (s.version) {
    if (threadIdx.x == 0) {
      spin_until_ge(s.version, s.want_seg ? *s.want_seg : want, "pull: parameter version");
      if (t.tag_out && (int)blockIdx.y == t.tag_seg && blockIdx.x == 0) *t.tag_out = ld_acquire_sys_u32(s.version) + 1u;
    }
    __syncthreads();
  }
  // peer reads are latency-bound (~2 us per round trip): four independent 16-byte loads in flight per thread
  const long long n4 = s.n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  const float4* src4 = reinterpret_cast<const float4*>(s.src);
  float4* dst4 = reinterpret_cast<float4*>(s.dst);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += 4 * stride) {
    float4 v[4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (i + k * stride < n4) v[k] = __ldcs(src4 + i + k * stride);
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (i + k * stride < n4) dst4[i + k * stride] = v[k];
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < s.n; i += stride) s.dst[i] = s.src[i];
}

}  // namespace

extern "C" int tfd_debug_adam_stamps(unsigned long long* out8) {
  TFD_CUDA_CHECK(cudaMemcpyFromSymbol(out8, g_adam_stamps, sizeof(g_adam_stamps)));
  return 0;
}

#define ST(s) static_cast<cudaStream_t>(s)

// storage type of the gradient slots for the adam/push entry points (set once per shard by the runtime)
static int g_next_slots_bf16 = 0;
extern "C" int tfd_set_slot_dtype(int is_bf16) {
  g_next_slots_bf16 = is_bf16 ? 1 : 0;
  return 0;
}

static int adam_launch(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                       float scale, long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1,
                       float beta2, float eps, unsigned* t_counter, const unsigned* wait_counter,
                       unsigned* version_flag, unsigned* global_step, unsigned* queue, int qlen,
                       const unsigned* shutdown, int slots_bf16, const unsigned* mask_ptr, const tfd_pub_t* pubspec,
                       void* stream) {
  TFD_REQUIRE(n % 4 == 0, "adam: arena length %lld must be a multiple of 4", n);
  TFD_REQUIRE(num_pieces <= kMaxPieces, "adam: too many pieces (%d)", num_pieces);
  TFD_REQUIRE(num_slots >= 1, "adam: need at least one gradient slot");
  PieceTable pt;
  pt.n = num_pieces;
  for (int i = 0; i < num_pieces; ++i) pt.p[i] = pieces[i];
  TFD_REQUIRE(num_slots <= TFD_MAX_SLOTS, "adam: at most %d gradient slots", TFD_MAX_SLOTS);
  AdamArgs a;
  memset(&a, 0, sizeof(a));
  a.w = w; a.m = m; a.v = v; a.shadow = static_cast<bf16*>(shadow);
  for (int i = 0; i < TFD_MAX_SLOTS; ++i) a.slots[i] = i < num_slots ? slots[i] : nullptr;
  a.num_slots = num_slots; a.scale = scale; a.n = n; a.lr = lr; a.beta1 = beta1; a.beta2 = beta2; a.eps = eps;
  a.t_counter = t_counter; a.wait_counter = wait_counter; a.version_flag = version_flag; a.global_step = global_step;
  a.queue = queue; a.qlen = qlen; a.shutdown = shutdown; a.slots_bf16 = slots_bf16;
  a.mask_ptr = mask_ptr;
  PubTable pub;
  memset(&pub, 0, sizeof(pub));
  for (int i = 0; i < kMaxPieces; ++i) pub.pc[i].f32_off = pub.pc[i].b16_off = -1;
  if (pubspec) {
    TFD_REQUIRE(pubspec->nworkers >= 0 && pubspec->nworkers <= TFD_MAX_SLOTS && pubspec->n_signal >= 0 &&
                    pubspec->n_signal <= TFD_MAX_SEG,
                "adam: publish spec with %d workers / %d signals", pubspec->nworkers, pubspec->n_signal);
    pub.nworkers = pubspec->nworkers;
    for (int i = 0; i < pubspec->nworkers; ++i) {
      pub.base[i] = static_cast<unsigned char*>(pubspec->worker_base[i]);
      pub.wver[i] = pubspec->worker_version[i];
    }
    for (int i = 0; i < num_pieces && pubspec->pieces; ++i) {
      pub.pc[i].f32_off = pubspec->pieces[i].f32_off;
      pub.pc[i].b16_off = pubspec->pieces[i].b16_off;
      pub.pc[i].ld16 = pubspec->pieces[i].ld16;
    }
    a.n_signal = pubspec->n_signal;
    for (int i = 0; i < pubspec->n_signal; ++i) a.signal_arrive[i] = pubspec->signal_arrive[i];
    a.local_step = pubspec->local_step;
  }
  long long blocks = (n / 8 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 3) blocks = 148 * 3;  // one resident wave (80 registers -> 3 blocks of 256 per SM)
  TFD_CUDA_CHECK(tfd_launch(adam_apply_kernel, dim3((unsigned)blocks), dim3(256), 0, ST(stream), a, pt, pub));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_adam_apply(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                              float scale, long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1,
                              float beta2, float eps, unsigned* t_counter, const unsigned* wait_counter,
                              unsigned* version_flag, unsigned* global_step, void* stream) {
  return adam_launch(w, m, v, shadow, slots, num_slots, scale, n, pieces, num_pieces, lr, beta1, beta2, eps, t_counter,
                     wait_counter, version_flag, global_step, nullptr, 0, nullptr, g_next_slots_bf16, nullptr, nullptr,
                     stream);
}

// tfd_adam_apply for a strict-sync ps/worker cluster: on top of the update, the kernel (a) sends this rank's own
// "gradients delivered" signals before it waits (colocated ranks), (b) stores the new values of the small variables
// into every worker's mirror and (c) releases every worker's LOCAL version word -- see tfd_pub_t in the header.
extern "C" int tfd_adam_apply_pub(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                                  float scale, long long n, const tfd_piece_t* pieces, int num_pieces, float lr,
                                  float beta1, float beta2, float eps, unsigned* t_counter, const unsigned* wait_counter,
                                  unsigned* version_flag, unsigned* global_step, const tfd_pub_t* pub, void* stream) {
  return adam_launch(w, m, v, shadow, slots, num_slots, scale, n, pieces, num_pieces, lr, beta1, beta2, eps, t_counter,
                     wait_counter, version_flag, global_step, nullptr, 0, nullptr, g_next_slots_bf16, nullptr, pub, stream);
}

// Worker-side signal of the same protocol when the worker GPU hosts no shard (dedicated ps GPUs): one tiny block adds 1
// (release, system scope) to every shard's arrival counter and advances the worker's step word.
__global__ void signal_kernel(const __grid_constant__ PushTable t, unsigned* local_step) {
  grid_dep_wait();
  if ((int)threadIdx.x < t.narrive) {
    __threadfence_system();
    red_add_release_sys_u32(t.arrive[threadIdx.x], 1u);
  }
  if (threadIdx.x == 0 && local_step) *local_step = *local_step + 1;
}

extern "C" int tfd_signal_arrive(unsigned* const* arrive, int narrive, unsigned* local_step, void* stream) {
  TFD_REQUIRE(narrive >= 1 && narrive <= kMaxSegs, "signal: bad shard count %d", narrive);
  PushTable t;
  memset(&t, 0, sizeof(t));
  t.narrive = narrive;
  for (int i = 0; i < narrive; ++i) t.arrive[i] = arrive[i];
  TFD_CUDA_CHECK(tfd_launch(signal_kernel, dim3(1), dim3(32), 0, ST(stream), t, local_step));
  TFD_LAUNCH_CHECK();
  return 0;
}

// Stand-alone wait on a list of (local) version words: the MLP's step has no conv1 to fold the wait into.
__global__ void gate_kernel(const __grid_constant__ PullTable t, const unsigned* want_ptr) {
  grid_dep_wait();
  if ((int)threadIdx.x < t.nseg) spin_until_ge(t.seg[threadIdx.x].version, *want_ptr, "gate: parameter version");
}

extern "C" int tfd_gate_flags(const unsigned* const* flags, int n, const unsigned* want_ptr, void* stream) {
  TFD_REQUIRE(n >= 1 && n <= kMaxSegs && want_ptr, "gate: bad flag count %d", n);
  PullTable t;
  memset(&t, 0, sizeof(t));
  t.nseg = n;
  for (int i = 0; i < n; ++i) t.seg[i].version = flags[i];
  TFD_CUDA_CHECK(tfd_launch(gate_kernel, dim3(1), dim3(32), 0, ST(stream), t, want_ptr));
  TFD_LAUNCH_CHECK();
  return 0;
}

// Backup workers (replicas_to_aggregate R < workers N; m:62-65, SyncReplicasOptimizer's stale-gradient drop): every
// worker stamps its slot with the update index it computed for (tag = version it pulled + 1).  One thread waits
// until at least R stamps equal the next update index t and publishes that set as a bit mask; the Adam launch that
// follows sums exactly those slots.  Stamps older than t are stale and simply never match.
__global__ void decide_kernel(const unsigned* stamps, int nworkers, int need, const unsigned* t_counter,
                              unsigned* mask_out, const unsigned* shutdown) {
  grid_dep_wait();
  if (threadIdx.x != 0 || blockIdx.x != 0) return;
  const unsigned t = t_counter[0] + 1;
  unsigned long long t0 = globaltimer_ns();
  unsigned it = 0;
  for (;;) {
    unsigned mask = 0;
    for (int w = 0; w < nworkers; ++w)
      if (ld_acquire_sys_u32(stamps + w) == t) mask |= 1u << w;
    if (__popc(mask) >= need) { *mask_out = mask; return; }
    if (shutdown && ld_acquire_sys_u32(shutdown) != 0) { *mask_out = 0; return; }
    __nanosleep(128);
    if ((++it & 0xff) == 0 && globaltimer_ns() - t0 > kSpinTimeoutNs) {
      printf("tfd: waited 20 s for %d gradients of update %u (have mask %x)\n", need, t, mask);
      __trap();
    }
  }
}

extern "C" int tfd_adam_apply_backup(float* w, float* m, float* v, void* shadow, const float* const* slots,
                                     int num_slots, int replicas_to_aggregate, long long n, const tfd_piece_t* pieces,
                                     int num_pieces, float lr, float beta1, float beta2, float eps, unsigned* t_counter,
                                     const unsigned* stamps, unsigned* mask_word, const unsigned* shutdown,
                                     unsigned* version_flag, unsigned* global_step, void* stream) {
  TFD_REQUIRE(replicas_to_aggregate >= 1 && replicas_to_aggregate <= num_slots && num_slots <= 32,
              "adam_backup: need 1 <= R (%d) <= workers (%d) <= 32", replicas_to_aggregate, num_slots);
  decide_kernel<<<1, 32, 0, ST(stream)>>>(stamps, num_slots, replicas_to_aggregate, t_counter, mask_word, shutdown);
  TFD_LAUNCH_CHECK();
  return adam_launch(w, m, v, shadow, slots, num_slots, 1.0f, n, pieces, num_pieces, lr, beta1, beta2, eps, t_counter,
                     nullptr, version_flag, global_step, nullptr, 0, nullptr, g_next_slots_bf16, mask_word, nullptr, stream);
}

extern "C" int tfd_adam_serve(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                              long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1, float beta2,
                              float eps, unsigned* t_counter, unsigned* queue, int qlen, const unsigned* shutdown,
                              unsigned* version_flag, unsigned* global_step, void* stream) {
  TFD_REQUIRE(queue != nullptr && qlen >= num_slots, "adam_serve: ring of %d cells cannot hold %d workers", qlen, num_slots);
  return adam_launch(w, m, v, shadow, slots, num_slots, 1.0f, n, pieces, num_pieces, lr, beta1, beta2, eps, t_counter,
                     nullptr, version_flag, global_step, queue, qlen, shutdown, g_next_slots_bf16, nullptr, nullptr, stream);
}

/* Multi-segment push: segs[i] = {src, dst, n}; after every segment is written the last block adds 1 to each
 * arrive counter.  (tfd_push_grads is the single-segment convenience form.) */
static int push_launch(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                       unsigned* const* arrive, int narrive, float scale, unsigned* block_counter,
                       unsigned* local_step, unsigned* const* queues, int qlen, unsigned worker, unsigned* my_ticket,
                       const unsigned* tag_ptr, unsigned* tag_cur, void* stream) {
  TFD_REQUIRE(nseg >= 1 && nseg <= kMaxSegs && narrive <= kMaxSegs, "push: bad segment count %d/%d", nseg, narrive);
  PushTable t;
  memset(&t, 0, sizeof(t));
  t.nseg = nseg; t.narrive = narrive;
  t.qlen = qlen; t.worker = worker; t.my_ticket = my_ticket; t.dst_bf16 = g_next_slots_bf16;
  t.tag_ptr = tag_ptr; t.tag_cur = tag_cur;
  for (int i = 0; i < narrive && queues; ++i) t.queue[i] = queues[i];
  long long maxn = 0;
  for (int i = 0; i < nseg; ++i) {
    t.seg[i] = PushSeg{static_cast<const float*>(srcs[i]), static_cast<float*>(dsts[i]), ns[i]};
    TFD_REQUIRE(((uintptr_t)srcs[i] & 15) == 0 && ((uintptr_t)dsts[i] & 7) == 0, "push: segment %d misaligned", i);
    if (ns[i] > maxn) maxn = ns[i];
  }
  for (int i = 0; i < narrive; ++i) t.arrive[i] = arrive[i];
  // Few, fat blocks: every block ends with a system-scope fence (an NVLink round trip when the slot is on a peer),
  // so a handful of 1024-thread blocks per segment beats hundreds of small ones for these sub-MB payloads.
  long long bx = (maxn / 4 + 4095) / 4096;
  if (bx < 1) bx = 1;
  if (bx > 32) bx = 32;
  TFD_CUDA_CHECK(tfd_launch(push_kernel, dim3((unsigned)bx, nseg), dim3(1024), 0, ST(stream), t, scale, block_counter, local_step));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_push_multi(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                              unsigned* const* arrive, int narrive, float scale, unsigned* block_counter,
                              unsigned* local_step, void* stream) {
  return push_launch(srcs, dsts, ns, nseg, arrive, narrive, scale, block_counter, local_step, nullptr, 0, 0, nullptr,
                     nullptr, nullptr, stream);
}

extern "C" int tfd_push_multi_async(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                                    unsigned* const* tickets, unsigned* const* queues, int nshards, int qlen,
                                    unsigned worker, unsigned* my_ticket, unsigned* block_counter, unsigned* local_step,
                                    void* stream) {
  TFD_REQUIRE(qlen > 0 && my_ticket != nullptr && queues != nullptr, "push_async: ring / ticket array missing");
  return push_launch(srcs, dsts, ns, nseg, tickets, nshards, 1.0f, block_counter, local_step, queues, qlen, worker,
                     my_ticket, nullptr, nullptr, stream);
}

extern "C" int tfd_push_multi_stamp(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                                    unsigned* const* stamps, int nshards, const unsigned* tag_ptr, unsigned* tag_cur,
                                    unsigned* block_counter, unsigned* local_step, void* stream) {
  TFD_REQUIRE(tag_ptr && tag_cur, "push_stamp: tag words missing");
  return push_launch(srcs, dsts, ns, nseg, stamps, nshards, 1.0f, block_counter, local_step, nullptr, 0, 0, nullptr,
                     tag_ptr, tag_cur, stream);
}

extern "C" int tfd_push_grads(const float* src, float* dst, long long n, float scale, unsigned* arrive_counter,
                              unsigned* block_counter, void* stream) {
  const void* s[1] = {src};
  void* d[1] = {dst};
  long long ns[1] = {n};
  unsigned* a[1] = {arrive_counter};
  return tfd_push_multi(s, d, ns, 1, a, arrive_counter ? 1 : 0, scale, block_counter, nullptr, stream);
}

static int pull_launch(const void* const* srcs, void* const* dsts, const long long* ns,
                       const unsigned* const* versions, const unsigned* const* want_segs, int nseg,
                       unsigned want_version, const unsigned* want_ptr, int tag_seg, unsigned* tag_out, void* stream) {
  TFD_REQUIRE(nseg >= 1 && nseg <= kMaxSegs, "pull: bad segment count %d", nseg);
  PullTable t;
  memset(&t, 0, sizeof(t));
  t.nseg = nseg;
  t.tag_seg = tag_seg; t.tag_out = tag_out;
  long long maxn = 0;
  for (int i = 0; i < nseg; ++i) {
    t.seg[i] = PullSeg{static_cast<const float*>(srcs[i]), static_cast<float*>(dsts[i]), ns[i], versions ? versions[i] : nullptr,
                        want_segs ? want_segs[i] : nullptr};
    if (ns[i] > maxn) maxn = ns[i];
  }
  long long bx = (maxn / 4 + 1023) / 1024;  // ~4 float4 per thread
  if (bx < 1) bx = 1;
  if (bx > 148) bx = 148;
  TFD_CUDA_CHECK(tfd_launch(pull_kernel, dim3((unsigned)bx, nseg), dim3(256), 0, ST(stream), t, want_version, want_ptr));
  TFD_LAUNCH_CHECK();
  return 0;
}

extern "C" int tfd_pull_multi(const void* const* srcs, void* const* dsts, const long long* ns,
                              const unsigned* const* versions, int nseg, unsigned want_version,
                              const unsigned* want_ptr, void* stream) {
  return pull_launch(srcs, dsts, ns, versions, nullptr, nseg, want_version, want_ptr, -1, nullptr, stream);
}

extern "C" int tfd_pull_multi_async(const void* const* srcs, void* const* dsts, const long long* ns,
                                    const unsigned* const* versions, const unsigned* const* want_segs, int nseg,
                                    void* stream) {
  return pull_launch(srcs, dsts, ns, versions, want_segs, nseg, 0, nullptr, -1, nullptr, stream);
}

extern "C" int tfd_pull_multi_tag(const void* const* srcs, void* const* dsts, const long long* ns,
                                  const unsigned* const* versions, const unsigned* const* want_segs, int nseg,
                                  int tag_seg, unsigned* tag_out, void* stream) {
  TFD_REQUIRE(tag_out && tag_seg >= 0 && tag_seg < nseg, "pull_tag: bad tag segment %d", tag_seg);
  return pull_launch(srcs, dsts, ns, versions, want_segs, nseg, 0, nullptr, tag_seg, tag_out, stream);
}

extern "C" int tfd_pull_params(const float* src, float* dst, long long n, const unsigned* version_flag, unsigned version,
                               void* stream) {
  const void* s[1] = {src};
  void* d[1] = {dst};
  long long ns[1] = {n};
  const unsigned* v[1] = {version_flag};
  return tfd_pull_multi(s, d, ns, v, 1, version, nullptr, stream);
}

// Plain asynchronous copy (cudaMemcpyAsync, kind default: the driver sees a peer-mapped source and uses the copy
// engines over NVLink).  The worker's prefetch of a big matrix from its owner shards goes through here: no SM is
// involved, so the transfer really overlaps the conv layers that own every SM while it runs, and it is a memcpy node
// in the captured step graph.  One call per copy.
extern "C" int tfd_memcpy_async(void* dst, const void* src, size_t bytes, void* stream) {
  TFD_CUDA_CHECK(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault, ST(stream)));
  return 0;
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
