/* C ABI of libtfd_b200.so -- the native sm_100a runtime of tensorflow_distributed_b200.
 *
 * Every entry point returns 0 on success; otherwise tfd_last_error() describes the failure.
 * Pointers are device pointers unless stated; `stream` is a cudaStream_t (NULL = default stream).
 * Layouts follow the reference: NHWC activations, HWIO filters, [in,out] dense weights
 * (/root/reference/mnist_python_m.py:93-128, :185-196).
 */
#ifndef TFD_B200_H_
#define TFD_B200_H_

#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* tfd_last_error(void);
int tfd_version(void);
/* Programmatic dependent launch for the step kernels (default off: measured no gain; also env TFD_PDL=1). */
int tfd_set_pdl(int on);
int tfd_device_info(int device, int* sm_count, int* cc_major, int* cc_minor, size_t* smem_optin);

/* ---- GEMM epilogues (tfd_gemm_bf16 `mode`) ---- */
#define TFD_EPI_STORE_F32 0      /* out fp32 = alpha*acc (+bias)                                    */
#define TFD_EPI_ATOMIC_F32 1     /* out fp32 += alpha*acc via red.global.add (split-K, push-reduce) */
#define TFD_EPI_BIAS_RELU_BF16 2 /* out bf16 = relu(alpha*acc + bias)                               */
#define TFD_EPI_STORE_BF16 3     /* out bf16 = alpha*acc (+bias)                                    */
#define TFD_EPI_RELUMASK_BF16 4  /* out bf16 = mask>0 ? alpha*acc : 0   (dgrad through a ReLU)      */
#define TFD_EPI_BIAS_RELU_F32 6  /* out fp32 = relu(alpha*acc + bias)              (tf32 arm)       */
#define TFD_EPI_RELUMASK_F32 7   /* out fp32 = mask>0 ? alpha*acc : 0, mask fp32   (tf32 arm)       */

/* ---- operand types of the tensor-core kernels (the `dtype` / `precision` arguments) ----
 * TFD_DTYPE_BF16: bf16 operands in memory (fp32 master weights keep a bf16 shadow), tcgen05 kind::f16.
 * TFD_DTYPE_TF32: fp32 operands in memory, tcgen05 kind::tf32 (10-bit mantissa products, fp32 accumulation): no
 *                 shadow, activations stay fp32 -- the arm comparable to the reference's all-fp32 graph
 *                 (/root/reference/mnist_python_m.py:198-200). */
#define TFD_DTYPE_BF16 0
#define TFD_DTYPE_TF32 1

/* tcgen05/TMEM/TMA GEMM, bf16 operands, fp32 accumulation: D[M,N] = A[M,K] * B[K,N].
 * a_mn=0: A stored [M,K] (K contiguous); a_mn=1: A stored [K,M].  b_mn=0: B stored [N,K]; b_mn=1: B stored [K,N].
 * Leading dimensions in elements, multiples of 8.  A/B may be peer-mapped addresses (fused parameter pull);
 * `out` may be a peer-mapped staging buffer (fused gradient push).  wait_flag (optional): the TMA producer
 * spins until *wait_flag >= wait_value before loading (parameter-version gate).  Split-K: with the atomic
 * epilogue every split red.adds into `out`; with TFD_EPI_STORE_F32 split z stores its partial plane at
 * out + z*split_stride (the consumer sums the planes: no atomics, deterministic).
 * Replaces tf.matmul at mnist_python_m.py:121,127 and the matmul gradients generated at m:222. */
int tfd_gemm_bf16(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M, int N, int K,
                  int bn, int splits, int mode, float alpha, void* out, long long ld_out, const float* bias,
                  const void* relu_mask, long long ld_mask, long long split_stride, const unsigned* wait_flag,
                  unsigned wait_value, void* stream);

/* Sharded form (fused parameter pull / gradient push over NVLink peer memory).  The sharded dimension is split
 * at seg_rows[0..nseg] (multiples of 128, seg_rows[0] = 0): TFD_SEG_B_K shards B's K rows (forward),
 * TFD_SEG_B_N shards B read as [N,K] (dgrad), TFD_SEG_OUT_M shards the output rows (wgrad: every tile is stored
 * straight into out_ptrs[shard], i.e. this worker's staging slot on the owning ps GPU).  B_ptrs[s] / out_ptrs[s]
 * point at row 0 of shard s's piece (peer-mapped addresses are fine); wait_flags[s] (optional) is shard s's
 * parameter-version word: the TMA producer acquires *wait_flags[s] >= wait_value + *wait_value_ptr (pointer
 * optional) before its first load from s.
 * No counterpart in the reference beyond the implicit RecvTensor traffic. */
int tfd_gemm_bf16_seg(const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb, int b_mn, int M,
                      int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                      const float* bias, const void* relu_mask, long long ld_mask, long long split_stride, int seg_mode,
                      int nseg, const int* seg_rows, void* const* out_ptrs, const unsigned* const* wait_flags,
                      unsigned wait_value, const unsigned* wait_value_ptr, void* stream);

/* tfd_gemm_bf16 / tfd_gemm_bf16_seg with the operand type as the first argument (TFD_DTYPE_*); leading dimensions
 * are in elements of that type; tf32 operands go with the fp32-output epilogues (0, 1, 6, 7). */
int tfd_gemm(int dtype, const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M, int N,
             int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out, const float* bias,
             const void* relu_mask, long long ld_mask, long long split_stride, void* stream);
int tfd_gemm_seg(int dtype, const void* A, long long lda, int a_mn, const void* const* B_ptrs, long long ldb, int b_mn,
                 int M, int N, int K, int bn, int splits, int mode, float alpha, void* out, long long ld_out,
                 const float* bias, const void* relu_mask, long long ld_mask, long long split_stride, int seg_mode,
                 int nseg, const int* seg_rows, void* const* out_ptrs, const unsigned* const* wait_flags,
                 unsigned wait_value, const unsigned* wait_value_ptr, void* stream);

/* Split-K GEMM that allocates nothing: out[M,N] fp32 = A*B (+bias, optional ReLU), same operand conventions as
 * tfd_gemm_bf16.  No counterpart in the reference (it has no native code and no allocator contract at all).
 * For callers that cannot call cudaMalloc -- hosts whose own pool already owns all device memory or that run
 * under a fixed memory budget, and latency-critical loops that must not enter the CUDA allocator because
 * cudaMalloc/cudaFree synchronise the device.  The split-K partial planes live in an 8 MiB static __device__
 * array inside the library (every GPU that loads the library holds it until the process exits), so
 * splits*M*N must be <= 2*1024*1024 floats; larger requests are rejected with an error message, nothing is run.
 * The first call on each device looks the array's device address up (cudaGetSymbolAddress) and sets a kernel
 * attribute: make that first call before any stream capture.  All calls share the one array and it keeps the
 * last call's partial sums: calls that overlap on the GPU (different streams, or two host threads) race on it --
 * serialise them.  Nothing in the array has to start at zero (each call writes every element it later reads). */
int tfd_gemm_bf16_splitk_noalloc(const void* A, long long lda, int a_mn, const void* B, long long ldb, int b_mn, int M,
                                 int N, int K, int bn, int splits, float* out, long long ld_out, const float* bias,
                                 int relu, void* stream);

/* CUDA-core fp32 reference GEMM (validation only): D[M,N] = A[M,K](bf16,row) * B[K,N](bf16,row). */
int tfd_gemm_ref(const void* A, long long lda, const void* B, long long ldb, int M, int N, int K, float* out,
                 long long ld_out, void* stream);

/* ---- elementwise / CUDA-core kernels ---- */
/* fp32 [rows,cols] -> bf16 [rows,ld_dst] (pad columns zeroed). */
int tfd_cast_bf16(const float* src, long long ld_src, void* dst, long long ld_dst, int rows, int cols, void* stream);
/* tf.random_normal (m:186-195): Philox4x32-10 + Box-Muller, mean 0.  out[i] is value (elem_offset + i) of the
 * stream `seed`, so a ps shard that owns rows [r0,r1) of a variable generates exactly its slice. */
int tfd_random_normal(float* out, long long n, unsigned long long seed, float stddev, long long elem_offset,
                      void* stream);

/* ---- the CNN's layer kernels.  `dtype` selects the precision arm: TFD_DTYPE_BF16 -> activations are bf16 and the
 * tensor-core weights are bf16 shadows; TFD_DTYPE_TF32 -> activations and weights are fp32 (read as tf32). ---- */

/* conv1 5x5 SAME 1->32 + bias + ReLU + 2x2 max-pool (m:109-111).  x fp32 [B,784]; w fp32 [25,32] (HWIO);
 * out [B,14,14,32] (activation type); argmax u8 [B,14,14,32] (0..3 = position inside the 2x2 window).
 * gate_flags (optional, ngate <= TFD_MAX_SEG): every block first waits until each of these LOCAL version words has
 * reached want_mul * *want_ptr -- the parameter-server worker's wait for the owner shards' updates, folded into
 * the step's first kernel (the owners store the words into this GPU's memory; nothing is polled over NVLink). */
int tfd_conv1_fwd(int dtype, const float* x, const float* w, const float* b, void* pool_out, unsigned char* argmax, int B,
                  const unsigned* const* gate_flags, int ngate, const unsigned* want_ptr, unsigned want_mul, void* stream);
/* conv1 filter/bias gradient with pool+ReLU backward folded in.  dpool fp32: nsplit partial planes of
 * d(loss)/d(pool1) [B,14,14,32], split_stride elements apart, added here in plane order.  dw fp32 [25,32] and db fp32
 * [32] are WRITTEN (fixed-order reduction through `scratch`, no atomics; they may be peer-mapped staging slots).
 * scratch: tfd_conv1_wgrad_scratch_floats(B) floats, zeroed once by the caller and then owned by this kernel. */
long long tfd_conv1_wgrad_scratch_floats(int B);
int tfd_conv1_wgrad(int dtype, const float* x, const float* dpool, int nsplit, long long split_stride, const void* pool_out,
                    const unsigned char* argmax, float* dw, float* db, float* scratch, int B, void* stream);

/* conv2 5x5 SAME 32->64 + bias + ReLU + 2x2 max-pool (m:114-116) on tcgen05: implicit GEMM, im2col by TMA
 * 4-D boxes with out-of-bounds zero fill, filter resident in shared memory (tf32: one 32-channel half per CTA).
 * in [B,14,14,32]; w [800,64] (HWIO flattened; bf16 shadow or the fp32 master; may be a peer pointer);
 * out [B,7,7,64]; argmax u8 [B,7,7,64]. */
int tfd_conv2_fwd(int dtype, const void* in, const void* w, const float* b, void* pool_out, unsigned char* argmax, int B,
                  void* stream);
/* d(loss)/d(pool1) fp32 = conv2 data gradient of dy [B,14,14,64] (tcgen05 implicit GEMM).  bf16: dx is ONE plane
 * [B,14,14,32]; tf32: TWO partial planes (one per half of dy's channels), B*14*14*32 floats apart, which
 * tfd_conv1_wgrad adds (nsplit = 2). */
int tfd_conv2_dgrad(int dtype, const void* dy, const void* w, float* dx, int B, void* stream);
/* dW fp32 [800,64] = im2col(in)^T dy (tcgen05, split over pixel tiles; per-CTA partial planes in `scratch`, summed in
 * plane order: deterministic, no atomics; dW is WRITTEN and may be a peer-mapped staging slot).
 * scratch: tfd_conv2_wgrad_scratch_floats(B) floats, zeroed once by the caller. */
long long tfd_conv2_wgrad_scratch_floats(int B);
int tfd_conv2_wgrad(int dtype, const void* in, const void* dy, float* dw, float* scratch, int B, void* stream);
/* CUDA-core references of the three conv2 kernels (bf16 arm, validation only; wgrad ACCUMULATES into a zeroed dw). */
int tfd_conv2_fwd_ref(const void* in, const void* w, const float* b, void* pool_out, unsigned char* argmax, int B,
                      void* stream);
int tfd_conv2_dgrad_ref(const void* dy, const void* w, float* dx, int B, void* stream);
int tfd_conv2_wgrad_ref(const void* in, const void* dy, float* dw, int B, void* stream);

/* max-pool + ReLU backward of the conv2 block: dpool fp32 [nsplit][B,7,7,64] (split-K partials, summed here;
 * split_stride in elements) -> dy [B,14,14,64] (dense, zeros written); db fp32 [64] WRITTEN (fixed-order reduction
 * through `scratch` = tfd_pool2_bwd_scratch_floats(B) floats, zeroed once). */
long long tfd_pool2_bwd_scratch_floats(int B);
int tfd_pool2_bwd(int dtype, const float* dpool, int nsplit, long long split_stride, const void* pool_out,
                  const unsigned char* argmax, void* dy, float* db, float* scratch, int B, void* stream);

/* Head, row phase (one block per example): optional fc bias + ReLU + inverted dropout (m:121-124), output
 * layer (m:127), softmax cross entropy with dense labels (m:205), dlogits=(softmax-y)/B, accuracy hit
 * (m:206-207) and the gradient flowing back into the hidden layer.
 *   pre_f32 != NULL: hidden = dropout(relu(sum_s pre_f32[s] + hid_bias))   (CNN fc1; pre_f32 holds pre_nsplit
 *                    split-K partial planes, pre_split_stride elements apart: the GEMM never needs atomics)
 *   else           : hidden = hid_in (already bias+ReLU'd by the GEMM epilogue; MLP)
 * Writes hid_out [B,ld_h] (post-dropout, for the output-layer wgrad), dhid [B,ld_h] (training only) in the
 * activation type, dlogits fp32 [B,10], loss_row fp32 [B], hit u8 [B]. keep_prob>=1 disables dropout.  The dropout
 * stream is Philox(seed; step + *step_ptr, element): step_ptr (optional, device) lets a captured CUDA graph advance it. */
int tfd_head_rows(int dtype, const float* pre_f32, int pre_nsplit, long long pre_split_stride, const void* hid_in,
                  const float* hid_bias, const float* out_w, const float* out_b, const float* labels, int B, int H,
                  long long ld_h, float keep_prob, unsigned long long seed, unsigned long long step,
                  const unsigned* step_ptr, int training, void* hid_out, void* dhid, float* dlogits, float* loss_row,
                  unsigned char* hit, unsigned char* keep_mask_out, void* stream);
/* Head, column phase: d_out_w[H,10] = hid^T dlogits, d_hid_bias[H] = colsum(dhid), d_out_b[10], loss = mean,
 * acc = mean(hit).  Deterministic (no atomics).  scalars: fp32[2] = {loss, accuracy}. */
int tfd_head_cols(int dtype, const void* hid, const void* dhid, long long ld_h, const float* dlogits,
                  const float* loss_row, const unsigned char* hit, int B, int H, float* d_out_w, float* d_hid_bias,
                  float* d_out_b, float* scalars, void* stream);
/* zero up to 16 fp32 buffers in one launch. */
int tfd_zero_multi(float* const* ptrs, const long long* ns, int count, void* stream);
/* column sums of an activation matrix (MLP hidden-layer bias gradients): out[N] = sum_r x[r, :]. */
int tfd_colsum(int dtype, const void* x, long long ld, int rows, int cols, float* out, void* stream);

/* ---- parameter-server kernels ---- */
#define TFD_MAX_SLOTS 8 /* gradient staging slots (= workers) one shard reduces */
#define TFD_MAX_SEG 8   /* parameter-server shards a fused pull/push GEMM can address */
#define TFD_SEG_NONE 0
#define TFD_SEG_B_K 1   /* forward: B's K rows sharded over ps GPUs  */
#define TFD_SEG_B_N 2   /* dgrad:   B read as [N,K], N rows sharded  */
#define TFD_SEG_OUT_M 3 /* wgrad:   output rows sharded (fused push) */
/* One piece of a shard arena that has a bf16 shadow for tensor-core consumers. */
typedef struct {
  long long arena_off;  /* element offset of the piece in the fp32 arena */
  long long numel;      /* rows*cols                                      */
  int cols;             /* logical columns                                */
  int ld_shadow;        /* leading dimension of the bf16 shadow (>= cols, multiple of 8); 0 = no shadow */
  long long shadow_off; /* element offset in the bf16 shadow arena        */
} tfd_piece_t;

/* Gradient reduce + scale + TF-form Adam + bf16 shadow publish on the owning shard (push, X2/X3; m:208,222):
 *   g = scale * sum_{r<num_slots} slots[r][i];  m,v,w updated in place;  shadow = bf16(w).
 * slots: HOST array of num_slots (<= TFD_MAX_SLOTS) device pointers to fp32 gradient arenas (staging slots written by the workers,
 * possibly through NVLink).  lr_t is computed on the device from the Adam step counter t_counter[0] (u32,
 * incremented by this kernel), so the kernel can sit in a CUDA graph.  If wait_counter != NULL the kernel first
 * spins until *wait_counter >= wait_base + num_slots*epoch (gradient arrivals for this step); on completion it
 * publishes *version_flag = epoch+1 with release semantics.  epoch is read from t_counter. */
int tfd_adam_apply(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots, float scale,
                   long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1, float beta2,
                   float eps, unsigned* t_counter, const unsigned* wait_counter, unsigned* version_flag,
                   unsigned* global_step, void* stream);

/* ---- strict-sync cluster step without stand-alone rendezvous kernels ----
 * The reference's workers block in a token dequeue after pushing their gradients (m:279-299).  Here the hand-off is
 * folded into kernels that run anyway.  Every worker owns, in its own peer-mapped heap, a mirror of the small
 * variables and one version word per shard.  tfd_adam_apply_pub is tfd_adam_apply plus:
 *   signal_arrive[n_signal]: block 0 first adds 1 (release, system scope) to these arrival counters -- the signal of
 *       the worker that lives on the same GPU as this shard (its gradient kernels ran earlier in this stream and
 *       stored straight into the owners' slots); local_step (optional) is that worker's step word, advanced at the end;
 *   pieces[i] (parallel to the tfd_piece_t array): where variable i's values go inside EVERY worker's heap
 *       (byte offsets; fp32 mirror and/or bf16 mirror; -1 = not mirrored; mirrors are padded to 16 bytes);
 *   worker_base[nworkers] / worker_version[nworkers]: heap base and this shard's version word of every worker.
 * After the update the last block stores version = t into every worker_version word (release): the worker's next
 * step polls LOCAL memory (tfd_conv1_fwd gate list / tfd_gate_flags / the GEMM's wait_flags). */
typedef struct {
  long long f32_off;
  long long b16_off;
  int ld16;
  int pad;
} tfd_pub_piece_t;
typedef struct {
  const tfd_pub_piece_t* pieces;              /* num_pieces entries, or NULL                       */
  void* worker_base[8];                       /* TFD_MAX_SLOTS                                     */
  unsigned* worker_version[8];
  int nworkers;
  int n_signal;
  unsigned* signal_arrive[8];                 /* TFD_MAX_SEG                                       */
  unsigned* local_step;
} tfd_pub_t;
int tfd_adam_apply_pub(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots, float scale,
                       long long n, const tfd_piece_t* pieces, int num_pieces, float lr, float beta1, float beta2,
                       float eps, unsigned* t_counter, const unsigned* wait_counter, unsigned* version_flag,
                       unsigned* global_step, const tfd_pub_t* pub, void* stream);
/* The same signal from a worker GPU that hosts no shard (dedicated ps GPUs): one tiny block. */
int tfd_signal_arrive(unsigned* const* arrive, int narrive, unsigned* local_step, void* stream);
/* Stand-alone wait until every (local) flag has reached *want_ptr (the MLP step has no conv1 to fold it into). */
int tfd_gate_flags(const unsigned* const* flags, int n, const unsigned* want_ptr, void* stream);

/* Storage type of the gradient staging slots used by tfd_adam_apply / tfd_adam_serve / tfd_push_multi[_async] from
 * now on in this process: 0 = fp32 (default), 1 = bf16 (halves the bytes a push moves over NVLink; the shard still
 * accumulates and applies in fp32).  The push-fused wgrad GEMM writes bf16 slots with TFD_EPI_STORE_BF16. */
int tfd_set_slot_dtype(int is_bf16);

/* Gradient push: copy (and scale) a worker's local fp32 gradient arena segment into its staging slot on the
 * owning shard (dst may be a peer pointer), then signal arrival: atomicAdd(arrive_counter, 1) at system scope
 * after a release fence, executed once by the last block. */
int tfd_push_grads(const float* src, float* dst, long long n, float scale, unsigned* arrive_counter,
                   unsigned* block_counter, void* stream);
/* Parameter pull of small tensors: wait until *version_flag >= version, then copy n fp32 from src (peer) to dst. */
int tfd_pull_params(const float* src, float* dst, long long n, const unsigned* version_flag, unsigned version,
                    void* stream);

/* Table forms: one launch moves every segment (one per destination shard / source tensor).  push: after all
 * segments are written the last block adds 1 (release, system scope) to each arrive counter.  pull: each
 * segment first waits for *versions[i] >= want_version + *want_ptr (NULL flag = no wait; want_ptr optional).
 * push: local_step (optional, device) is incremented by the last block -- the worker's step counter that
 * want_ptr / wait_value_ptr of the next step read, so a captured CUDA graph advances by itself. */
int tfd_push_multi(const void* const* srcs, void* const* dsts, const long long* ns, int nseg, unsigned* const* arrive,
                   int narrive, float scale, unsigned* block_counter, unsigned* local_step, void* stream);
int tfd_pull_multi(const void* const* srcs, void* const* dsts, const long long* ns, const unsigned* const* versions,
                   int nseg, unsigned want_version, const unsigned* want_ptr, void* stream);

/* ---- asynchronous replicas: apply on arrival, in per-shard ticket order (m:72-75 with sync_replicas=False) ----
 * Worker: after its slot on a shard is complete it takes a ticket (atomic add on the shard's dispenser word) and
 * writes worker+1 into ring cell ticket % qlen (release); my_ticket[k] (local) receives ticket+1, the shard version
 * at which that gradient has been applied -- the worker waits for it before overwriting its slot again
 * (tfd_pull_multi_async: segment i waits for *versions[i] >= *want_segs[i]).
 * Shard: tfd_adam_serve applies exactly ONE gradient: the slot named by ring cell (t mod qlen), scale 1, then clears
 * the cell, bumps t / global_step and releases version = t+1.  Every arriving gradient is applied exactly once, in
 * ticket order, and no worker ever waits for another worker.  *shutdown != 0 makes a waiting serve kernel leave. */
int tfd_adam_serve(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots, long long n,
                   const tfd_piece_t* pieces, int num_pieces, float lr, float beta1, float beta2, float eps,
                   unsigned* t_counter, unsigned* queue, int qlen, const unsigned* shutdown, unsigned* version_flag,
                   unsigned* global_step, void* stream);
int tfd_push_multi_async(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                         unsigned* const* tickets, unsigned* const* queues, int nshards, int qlen, unsigned worker,
                         unsigned* my_ticket, unsigned* block_counter, unsigned* local_step, void* stream);
int tfd_pull_multi_async(const void* const* srcs, void* const* dsts, const long long* ns,
                         const unsigned* const* versions, const unsigned* const* want_segs, int nseg, void* stream);

/* ---- backup workers: aggregate the first R of N gradients, drop stale ones (m:62-65, m:210-220) ----
 * Worker: tfd_pull_multi_tag waits (per segment) for *versions[i] >= *want_segs[i] (= the tag of its previous push)
 * and stores tag = version+1 of segment tag_seg's shard in *tag_out; tfd_push_multi_stamp writes the slots and then
 * stamps every shard's word for this worker with *tag_ptr (release) and copies it to *tag_cur.
 * Shard: tfd_adam_apply_backup = a one-thread decide kernel that waits until >= R stamps equal the next update index
 * t and publishes them as a bit mask, followed by the Adam kernel that sums exactly those slots (scale 1/count).
 * A stamp older than t is a stale gradient: it never matches and is never read.  *shutdown != 0 releases the wait. */
int tfd_adam_apply_backup(float* w, float* m, float* v, void* shadow, const float* const* slots, int num_slots,
                          int replicas_to_aggregate, long long n, const tfd_piece_t* pieces, int num_pieces, float lr,
                          float beta1, float beta2, float eps, unsigned* t_counter, const unsigned* stamps,
                          unsigned* mask_word, const unsigned* shutdown, unsigned* version_flag, unsigned* global_step,
                          void* stream);
int tfd_push_multi_stamp(const void* const* srcs, void* const* dsts, const long long* ns, int nseg,
                         unsigned* const* stamps, int nshards, const unsigned* tag_ptr, unsigned* tag_cur,
                         unsigned* block_counter, unsigned* local_step, void* stream);
int tfd_pull_multi_tag(const void* const* srcs, void* const* dsts, const long long* ns, const unsigned* const* versions,
                       const unsigned* const* want_segs, int nseg, int tag_seg, unsigned* tag_out, void* stream);

/* ---- parameter server over NVLS multicast (EXPERIMENTAL: host-checked, not yet run on hardware) ----
 * Every GPU maps one symmetric heap at a unicast and at a multicast address (torch symmetric memory).  Workers
 * write gradients into their own copy and signal with one multimem.red; the owner shard reads the gradient SUM
 * with multimem.ld_reduce (ingress 1x instead of workers-x), applies TF-form Adam to its private master weights
 * and stores the result into every worker's mirror with multimem.st (egress 1x); workers read parameters locally.
 * One piece = the rows of one variable this shard owns. */
typedef struct {
  long long owner_off;  /* element offset in the owner's private w/m/v arenas                              */
  long long numel;
  int cols;
  int ld_shadow;        /* 0: published in fp32 at mirror_off; else bf16 rows of ld_shadow at shadow_off   */
  long long grad_off;   /* element offset in the (symmetric) gradient arena, single-shard layout            */
  long long mirror_off; /* element offset in the (symmetric) fp32 mirror                                    */
  long long shadow_off; /* element offset in the (symmetric) bf16 shadow mirror                             */
} tfd_nvls_piece_t;
/* adam = 1: wait until *arrive >= arrivals_per_step * (t+1), reduce, update, publish, version += 1 on every GPU.
 * adam = 0: publish the current parameters only (initialisation / restore); no counter moves, the caller
 *           synchronises on the host afterwards. */
int tfd_nvls_apply(float* w, float* m, float* v, const float* mc_grad, float* mc_mirror_w, void* mc_mirror_shadow,
                   const tfd_nvls_piece_t* pieces, int num_pieces, float scale, float lr, float beta1, float beta2,
                   float eps, unsigned* t_counter, const unsigned* arrive, unsigned arrivals_per_step,
                   unsigned* mc_version, unsigned* global_step, int adam, void* stream);
/* Worker step brackets: wait on the LOCAL version word until *version >= base + per_step * *step; and, after the
 * backward pass, add 1 to every shard's arrival counter through the multicast alias and advance *step. */
int tfd_nvls_gate(const unsigned* version, const unsigned* step, unsigned per_step, unsigned base, void* stream);
int tfd_nvls_signal(unsigned* mc_arrive, unsigned* step, void* stream);

/* ---- diagnostics: globaltimer phase stamps of the most recent Adam / conv1-forward kernel on the current GPU ----
 * adam: [0] start, [1] arrive signals issued, [2] arrivals complete, [3] block 0 done updating, [4] all blocks done,
 * [5] versions released.  conv1: [0] block (0,0) start, [1] gate passed, [2] done. */
int tfd_debug_adam_stamps(unsigned long long* out8);
int tfd_debug_conv1_stamps(unsigned long long* out4);

/* ---- peer memory (symmetric heap over cudaIpc) ---- */
/* cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDefault): copy-engine transfer (src may be peer-mapped); used for the
 * worker's prefetch of a big matrix from its owner shards so that no SM is spent on it. */
int tfd_memcpy_async(void* dst, const void* src, size_t bytes, void* stream);
int tfd_set_device(int device);
int tfd_ipc_alloc(size_t bytes, void** dev_ptr, void* handle_out64);    /* cudaMalloc + cudaIpcGetMemHandle */
int tfd_ipc_open(const void* handle64, void** dev_ptr);                  /* cudaIpcOpenMemHandle             */
int tfd_ipc_close(void* dev_ptr);
int tfd_free(void* dev_ptr);
int tfd_enable_peer_access(int device, int peer);

#ifdef __cplusplus
}
#endif
#endif /* TFD_B200_H_ */
