"""ctypes binding of ``libtfd_b200.so`` (C ABI in ``include/tfd_b200.h``).

The library is built in-tree by ``build_native.py`` (nvcc, sm_100a only) and
loaded from ``tensorflow_distributed_b200/_native/``.  There is no fallback: on
a machine with a GPU a missing library is an error, so a silent eager-PyTorch
path can never stand in for the kernels.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_longlong, c_size_t, c_uint, c_ulonglong, c_void_p
from typing import Optional

import torch

_LIB_PATH = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "_native", "libtfd_b200.so")
_lib: Optional[ctypes.CDLL] = None

EPI_STORE_F32, EPI_ATOMIC_F32, EPI_BIAS_RELU_BF16, EPI_STORE_BF16, EPI_RELUMASK_BF16 = range(5)
EPI_BIAS_RELU_F32, EPI_RELUMASK_F32 = 6, 7
DTYPE_BF16, DTYPE_TF32 = 0, 1

#: number of kernel launches issued through this binding (bench.py reports it as gpu_launches)
launch_count = 0


class NativeError(RuntimeError):
    pass


class Piece(ctypes.Structure):
    _fields_ = [("arena_off", c_longlong), ("numel", c_longlong), ("cols", c_int), ("ld_shadow", c_int),
                ("shadow_off", c_longlong)]


class PubPiece(ctypes.Structure):
    """tfd_pub_piece_t: where a variable's new values go inside every worker's heap (byte offsets, -1 = not mirrored)."""
    _fields_ = [("f32_off", c_longlong), ("b16_off", c_longlong), ("ld16", c_int), ("pad", c_int)]


class PubSpec(ctypes.Structure):
    """tfd_pub_t: publish / signal spec of tfd_adam_apply_pub (strict-sync cluster step without rendezvous kernels)."""
    _fields_ = [("pieces", ctypes.POINTER(PubPiece)), ("worker_base", c_void_p * 8), ("worker_version", c_void_p * 8),
                ("nworkers", c_int), ("n_signal", c_int), ("signal_arrive", c_void_p * 8), ("local_step", c_void_p)]


def make_pub(pieces, worker_bases, worker_versions, signal_arrive=(), local_step=None) -> PubSpec:
    """pieces: list of (f32_off, b16_off, ld16) parallel to the shard's Piece list.  The returned struct keeps the
    piece array alive through an attribute."""
    spec = PubSpec()
    arr = (PubPiece * max(1, len(pieces)))(*[PubPiece(int(a), int(b), int(c), 0) for a, b, c in pieces])
    spec._keep = arr
    spec.pieces = ctypes.cast(arr, ctypes.POINTER(PubPiece))
    spec.nworkers = len(worker_bases)
    for i, (b, v) in enumerate(zip(worker_bases, worker_versions)):
        spec.worker_base[i], spec.worker_version[i] = int(b), int(v)
    spec.n_signal = len(signal_arrive)
    for i, a in enumerate(signal_arrive):
        spec.signal_arrive[i] = int(a)
    spec.local_step = int(local_step) if local_step else None
    return spec


class NvlsPiece(ctypes.Structure):
    """tfd_nvls_piece_t: the rows of one variable a shard owns, with their places in the symmetric arenas."""
    _fields_ = [("owner_off", c_longlong), ("numel", c_longlong), ("cols", c_int), ("ld_shadow", c_int),
                ("grad_off", c_longlong), ("mirror_off", c_longlong), ("shadow_off", c_longlong)]


def lib_path() -> str:
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise NativeError("native library %s is missing: run `python build_native.py` "
                              "(or __graft_entry__.build()) first" % _LIB_PATH)
        _lib = ctypes.CDLL(_LIB_PATH)
        _lib.tfd_last_error.restype = c_char_p
        _lib.tfd_version.restype = c_int
    return _lib


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise NativeError("%s failed (rc=%d): %s" % (what, rc, lib().tfd_last_error().decode()))


def _p(t) -> c_void_p:
    if t is None:
        return c_void_p(0)
    if isinstance(t, int):
        return c_void_p(t)
    return c_void_p(t.data_ptr())


def _stream() -> c_void_p:
    return c_void_p(torch.cuda.current_stream().cuda_stream)


def _count(n: int = 1) -> None:
    global launch_count
    launch_count += n


def set_device(dev: int) -> None:
    _check(lib().tfd_set_device(c_int(dev)), "tfd_set_device")


def set_slot_dtype(is_bf16: bool) -> None:
    _check(lib().tfd_set_slot_dtype(c_int(int(is_bf16))), "tfd_set_slot_dtype")


def device_info(dev: int = 0) -> dict:
    sm, maj, mnr, sm_opt = c_int(), c_int(), c_int(), c_size_t()
    _check(lib().tfd_device_info(c_int(dev), ctypes.byref(sm), ctypes.byref(maj), ctypes.byref(mnr),
                                 ctypes.byref(sm_opt)), "tfd_device_info")
    return {"sm_count": sm.value, "cc": (maj.value, mnr.value), "smem_optin": sm_opt.value}


# ---------------------------------------------------------------------------------------------
def gemm_bf16(A, a_mn: int, B, b_mn: int, M: int, N: int, K: int, out, *, mode: int, bn: int = 128, splits: int = 1,
              alpha: float = 1.0, bias=None, relu_mask=None, lda: Optional[int] = None, ldb: Optional[int] = None,
              ld_out: Optional[int] = None, ld_mask: Optional[int] = None, wait_flag=None, wait_value: int = 0,
              split_stride: int = 0, tf32: bool = False) -> None:
    """D[M,N] = A B on tcgen05.  tf32=False: bf16 operands; tf32=True: fp32 operands read as tf32 (kind::tf32)."""
    lda = lda if lda is not None else A.stride(0)
    ldb = ldb if ldb is not None else B.stride(0)
    ld_out = ld_out if ld_out is not None else out.stride(0)
    ld_mask = ld_mask if ld_mask is not None else (relu_mask.stride(0) if relu_mask is not None else 0)
    if tf32:
        assert wait_flag is None
        _check(lib().tfd_gemm(c_int(DTYPE_TF32), _p(A), c_longlong(lda), c_int(a_mn), _p(B), c_longlong(ldb), c_int(b_mn),
                              c_int(M), c_int(N), c_int(K), c_int(bn), c_int(splits), c_int(mode), c_float(alpha), _p(out),
                              c_longlong(ld_out), _p(bias), _p(relu_mask), c_longlong(ld_mask), c_longlong(split_stride),
                              _stream()), "tfd_gemm(tf32)")
        _count()
        return
    _check(lib().tfd_gemm_bf16(_p(A), c_longlong(lda), c_int(a_mn), _p(B), c_longlong(ldb), c_int(b_mn), c_int(M),
                               c_int(N), c_int(K), c_int(bn), c_int(splits), c_int(mode), c_float(alpha), _p(out),
                               c_longlong(ld_out), _p(bias), _p(relu_mask), c_longlong(ld_mask),
                               c_longlong(split_stride), _p(wait_flag),
                               c_uint(wait_value), _stream()), "tfd_gemm_bf16")
    _count()


SEG_NONE, SEG_B_K, SEG_B_N, SEG_OUT_M = range(4)


def gemm_bf16_seg(A, a_mn: int, B_ptrs, b_mn: int, M: int, N: int, K: int, out, *, mode: int, lda: int, ldb: int,
                  ld_out: int, seg_mode: int, seg_rows, out_ptrs=None, wait_flags=None, wait_value: int = 0,
                  wait_value_ptr=None,
                  bn: int = 128, splits: int = 1, alpha: float = 1.0, bias=None, relu_mask=None, ld_mask: int = 0,
                  split_stride: int = 0, tf32: bool = False) -> None:
    """Sharded GEMM: B_ptrs / out_ptrs / wait_flags are per-shard raw device pointers (ints), possibly peer-mapped."""
    nseg = len(seg_rows) - 1
    BP = (c_void_p * nseg)(*[c_void_p(int(x)) for x in (B_ptrs if B_ptrs is not None else [0] * nseg)])
    OP = (c_void_p * nseg)(*[c_void_p(int(x)) for x in out_ptrs]) if out_ptrs is not None else None
    WF = (c_void_p * nseg)(*[c_void_p(int(x) if x else 0) for x in wait_flags]) if wait_flags is not None else None
    SR = (c_int * (nseg + 1))(*[int(r) for r in seg_rows])
    _check(lib().tfd_gemm_seg(c_int(DTYPE_TF32 if tf32 else DTYPE_BF16), _p(A), c_longlong(lda), c_int(a_mn), BP,
                              c_longlong(ldb), c_int(b_mn), c_int(M), c_int(N), c_int(K), c_int(bn), c_int(splits),
                              c_int(mode), c_float(alpha), _p(out), c_longlong(ld_out), _p(bias), _p(relu_mask),
                              c_longlong(ld_mask), c_longlong(split_stride), c_int(seg_mode), c_int(nseg), SR, OP, WF,
                              c_uint(wait_value), _p(wait_value_ptr), _stream()), "tfd_gemm_seg")
    _count()


def gemm_ref(A, B, M, N, K, out) -> None:
    _check(lib().tfd_gemm_ref(_p(A), c_longlong(A.stride(0)), _p(B), c_longlong(B.stride(0)), c_int(M), c_int(N),
                              c_int(K), _p(out), c_longlong(out.stride(0)), _stream()), "tfd_gemm_ref")
    _count()


def cast_bf16(src, dst, rows: int, cols: int, ld_src: Optional[int] = None, ld_dst: Optional[int] = None) -> None:
    _check(lib().tfd_cast_bf16(_p(src), c_longlong(ld_src if ld_src is not None else cols), _p(dst),
                               c_longlong(ld_dst if ld_dst is not None else cols), c_int(rows), c_int(cols), _stream()),
           "tfd_cast_bf16")
    _count()


def random_normal(out, seed: int, stddev: float = 1.0, elem_offset: int = 0) -> None:
    _check(lib().tfd_random_normal(_p(out), c_longlong(out.numel()), c_ulonglong(seed), c_float(stddev),
                                   c_longlong(elem_offset), _stream()),
           "tfd_random_normal")
    _count()


def _dt(tf32: bool) -> c_int:
    return c_int(DTYPE_TF32 if tf32 else DTYPE_BF16)


def scratch_floats(kind: str, B: int) -> int:
    """Size (floats) of the reduction scratch of a deterministic-gradient kernel: 'conv1_wgrad' / 'conv2_wgrad' /
    'pool2_bwd'.  The caller allocates it zeroed once; the kernel owns it from then on."""
    fn = getattr(lib(), "tfd_%s_scratch_floats" % kind)
    fn.restype = c_longlong
    return int(fn(c_int(B)))


def conv1_fwd(x, w, b, pool_out, argmax, B: int, tf32: bool = False, gate_flags=None, want_ptr=None,
              want_mul: int = 1) -> None:
    """gate_flags: LOCAL version words (ints or tensors) every block waits on (>= *want_ptr) before it reads w / b."""
    n = len(gate_flags) if gate_flags else 0
    G = (c_void_p * max(1, n))(*[c_void_p(int(f) if isinstance(f, int) else f.data_ptr()) for f in (gate_flags or [])])
    _check(lib().tfd_conv1_fwd(_dt(tf32), _p(x), _p(w), _p(b), _p(pool_out), _p(argmax), c_int(B), G, c_int(n),
                               _p(want_ptr), c_uint(want_mul), _stream()), "tfd_conv1_fwd")
    _count()


def conv1_wgrad(x, dpool, pool_out, argmax, dw, db, scratch, B: int, tf32: bool = False, nsplit: int = 1,
                split_stride: int = 0) -> None:
    _check(lib().tfd_conv1_wgrad(_dt(tf32), _p(x), _p(dpool), c_int(nsplit), c_longlong(split_stride), _p(pool_out),
                                 _p(argmax), _p(dw), _p(db), _p(scratch), c_int(B), _stream()), "tfd_conv1_wgrad")
    _count()


def conv2_fwd(inp, w, b, pool_out, argmax, B: int, ref: bool = False, tf32: bool = False) -> None:
    if ref:
        _check(lib().tfd_conv2_fwd_ref(_p(inp), _p(w), _p(b), _p(pool_out), _p(argmax), c_int(B), _stream()), "tfd_conv2_fwd_ref")
    else:
        _check(lib().tfd_conv2_fwd(_dt(tf32), _p(inp), _p(w), _p(b), _p(pool_out), _p(argmax), c_int(B), _stream()),
               "tfd_conv2_fwd")
    _count()


def conv2_dgrad(dy, w, dx, B: int, ref: bool = False, tf32: bool = False) -> None:
    """tf32: dx holds TWO partial planes (B*14*14*32 floats apart) that conv1_wgrad adds."""
    if ref:
        _check(lib().tfd_conv2_dgrad_ref(_p(dy), _p(w), _p(dx), c_int(B), _stream()), "tfd_conv2_dgrad_ref")
    else:
        _check(lib().tfd_conv2_dgrad(_dt(tf32), _p(dy), _p(w), _p(dx), c_int(B), _stream()), "tfd_conv2_dgrad")
    _count()


def conv2_wgrad(inp, dy, dw, scratch, B: int, ref: bool = False, tf32: bool = False) -> None:
    """dw is WRITTEN (deterministic plane reduction through `scratch`); the validation kernel (ref) accumulates."""
    if ref:
        dw.zero_()
        _check(lib().tfd_conv2_wgrad_ref(_p(inp), _p(dy), _p(dw), c_int(B), _stream()), "tfd_conv2_wgrad_ref")
    else:
        _check(lib().tfd_conv2_wgrad(_dt(tf32), _p(inp), _p(dy), _p(dw), _p(scratch), c_int(B), _stream()),
               "tfd_conv2_wgrad")
    _count()


def pool2_bwd(dpool, pool_out, argmax, dy, db, scratch, B: int, nsplit: int = 1, split_stride: int = 0,
              tf32: bool = False) -> None:
    _check(lib().tfd_pool2_bwd(_dt(tf32), _p(dpool), c_int(nsplit), c_longlong(split_stride), _p(pool_out), _p(argmax),
                               _p(dy), _p(db), _p(scratch), c_int(B), _stream()), "tfd_pool2_bwd")
    _count()


def head_rows(pre_f32, hid_in, hid_bias, out_w, out_b, labels, B: int, H: int, ld_h: int, keep_prob: float,
              seed: int, step: int, training: bool, hid_out, dhid, dlogits, loss_row, hit, keep_mask_out=None,
              step_ptr=None, pre_nsplit: int = 1, pre_split_stride: int = 0, tf32: bool = False) -> None:
    _check(lib().tfd_head_rows(_dt(tf32), _p(pre_f32), c_int(pre_nsplit), c_longlong(pre_split_stride), _p(hid_in),
                               _p(hid_bias), _p(out_w), _p(out_b), _p(labels), c_int(B), c_int(H), c_longlong(ld_h),
                               c_float(keep_prob), c_ulonglong(seed), c_ulonglong(step), _p(step_ptr),
                               c_int(int(training)), _p(hid_out), _p(dhid), _p(dlogits), _p(loss_row), _p(hit),
                               _p(keep_mask_out), _stream()), "tfd_head_rows")
    _count()


def head_cols(hid, dhid, ld_h: int, dlogits, loss_row, hit, B: int, H: int, d_out_w, d_hid_bias, d_out_b,
              scalars, tf32: bool = False) -> None:
    _check(lib().tfd_head_cols(_dt(tf32), _p(hid), _p(dhid), c_longlong(ld_h), _p(dlogits), _p(loss_row), _p(hit),
                               c_int(B), c_int(H), _p(d_out_w), _p(d_hid_bias), _p(d_out_b), _p(scalars), _stream()),
           "tfd_head_cols")
    _count()


def zero_multi(tensors) -> None:
    n = len(tensors)
    P = (c_void_p * n)(*[c_void_p(t.data_ptr()) for t in tensors])
    NS = (c_longlong * n)(*[t.numel() for t in tensors])
    _check(lib().tfd_zero_multi(P, NS, c_int(n), _stream()), "tfd_zero_multi")
    _count()


def colsum(x, ld: int, rows: int, cols: int, out, tf32: bool = False) -> None:
    _check(lib().tfd_colsum(_dt(tf32), _p(x), c_longlong(ld), c_int(rows), c_int(cols), _p(out), _stream()), "tfd_colsum")
    _count()


def colsum_bf16(x, ld: int, rows: int, cols: int, out) -> None:
    colsum(x, ld, rows, cols, out, False)


def adam_apply(w, m, v, shadow, slots_dev, num_slots: int, scale: float, n: int, pieces, lr: float, t_counter,
               wait_counter=None, version_flag=None, global_step=None, beta1: float = 0.9, beta2: float = 0.999,
               eps: float = 1e-8, pub: Optional[PubSpec] = None) -> None:
    arr = (Piece * max(1, len(pieces)))(*pieces)
    slots = (c_void_p * num_slots)(*[c_void_p(int(x)) for x in slots_dev[:num_slots]])  # host array of device pointers
    if pub is None:
        _check(lib().tfd_adam_apply(_p(w), _p(m), _p(v), _p(shadow), slots, c_int(num_slots), c_float(scale),
                                    c_longlong(n), arr, c_int(len(pieces)), c_float(lr), c_float(beta1), c_float(beta2),
                                    c_float(eps), _p(t_counter), _p(wait_counter), _p(version_flag), _p(global_step),
                                    _stream()), "tfd_adam_apply")
    else:
        _check(lib().tfd_adam_apply_pub(_p(w), _p(m), _p(v), _p(shadow), slots, c_int(num_slots), c_float(scale),
                                        c_longlong(n), arr, c_int(len(pieces)), c_float(lr), c_float(beta1),
                                        c_float(beta2), c_float(eps), _p(t_counter), _p(wait_counter), _p(version_flag),
                                        _p(global_step), ctypes.byref(pub), _stream()), "tfd_adam_apply_pub")
    _count()


def signal_arrive(arrive, local_step=None) -> None:
    """Worker without a shard on its GPU: add 1 (release) to every shard's arrival counter, advance the step word."""
    A = (c_void_p * len(arrive))(*[c_void_p(int(x)) for x in arrive])
    _check(lib().tfd_signal_arrive(A, c_int(len(arrive)), _p(local_step), _stream()), "tfd_signal_arrive")
    _count()


def gate_flags(flags, want_ptr) -> None:
    """Wait (on the device) until every flag word has reached *want_ptr."""
    F = (c_void_p * len(flags))(*[c_void_p(int(f) if isinstance(f, int) else f.data_ptr()) for f in flags])
    _check(lib().tfd_gate_flags(F, c_int(len(flags)), _p(want_ptr), _stream()), "tfd_gate_flags")
    _count()


def push_multi(srcs, dsts, ns, arrive, scale: float, block_counter, local_step=None) -> None:
    n = len(srcs)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    A = (c_void_p * max(1, len(arrive)))(*[c_void_p(int(x)) for x in arrive])
    _check(lib().tfd_push_multi(S, D, NS, c_int(n), A, c_int(len(arrive)), c_float(scale), _p(block_counter),
                                _p(local_step), _stream()), "tfd_push_multi")
    _count()


def pull_multi(srcs, dsts, ns, versions, want_version: int, want_ptr=None) -> None:
    n = len(srcs)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    V = (c_void_p * n)(*[c_void_p(int(x) if x else 0) for x in versions]) if versions is not None else None
    _check(lib().tfd_pull_multi(S, D, NS, V, c_int(n), c_uint(want_version), _p(want_ptr), _stream()),
           "tfd_pull_multi")
    _count()


def adam_serve(w, m, v, shadow, slots, num_slots: int, n: int, pieces, lr: float, t_counter, queue, qlen: int,
               shutdown, version_flag, global_step, beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8) -> None:
    arr = (Piece * max(1, len(pieces)))(*pieces)
    sl = (c_void_p * num_slots)(*[c_void_p(int(x)) for x in slots[:num_slots]])
    _check(lib().tfd_adam_serve(_p(w), _p(m), _p(v), _p(shadow), sl, c_int(num_slots), c_longlong(n), arr,
                                c_int(len(pieces)), c_float(lr), c_float(beta1), c_float(beta2), c_float(eps),
                                _p(t_counter), _p(queue), c_int(qlen), _p(shutdown), _p(version_flag), _p(global_step),
                                _stream()), "tfd_adam_serve")
    _count()


def push_multi_async(srcs, dsts, ns, tickets, queues, qlen: int, worker: int, my_ticket, block_counter,
                     local_step=None) -> None:
    n, k = len(srcs), len(tickets)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    T = (c_void_p * k)(*[c_void_p(int(x)) for x in tickets])
    Q = (c_void_p * k)(*[c_void_p(int(x)) for x in queues])
    _check(lib().tfd_push_multi_async(S, D, NS, c_int(n), T, Q, c_int(k), c_int(qlen), c_uint(worker), _p(my_ticket),
                                      _p(block_counter), _p(local_step), _stream()), "tfd_push_multi_async")
    _count()


def pull_multi_async(srcs, dsts, ns, versions, want_segs) -> None:
    n = len(srcs)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    V = (c_void_p * n)(*[c_void_p(int(x)) for x in versions])
    W = (c_void_p * n)(*[c_void_p(int(x)) for x in want_segs])
    _check(lib().tfd_pull_multi_async(S, D, NS, V, W, c_int(n), _stream()), "tfd_pull_multi_async")
    _count()


def adam_apply_backup(w, m, v, shadow, slots, num_slots: int, replicas_to_aggregate: int, n: int, pieces, lr: float,
                      t_counter, stamps, mask_word, shutdown, version_flag, global_step, beta1: float = 0.9,
                      beta2: float = 0.999, eps: float = 1e-8) -> None:
    arr = (Piece * max(1, len(pieces)))(*pieces)
    sl = (c_void_p * num_slots)(*[c_void_p(int(x)) for x in slots[:num_slots]])
    _check(lib().tfd_adam_apply_backup(_p(w), _p(m), _p(v), _p(shadow), sl, c_int(num_slots),
                                       c_int(replicas_to_aggregate), c_longlong(n), arr, c_int(len(pieces)), c_float(lr),
                                       c_float(beta1), c_float(beta2), c_float(eps), _p(t_counter), _p(stamps),
                                       _p(mask_word), _p(shutdown), _p(version_flag), _p(global_step), _stream()),
           "tfd_adam_apply_backup")
    _count(2)


def push_multi_stamp(srcs, dsts, ns, stamps, tag_ptr, tag_cur, block_counter, local_step=None) -> None:
    n, k = len(srcs), len(stamps)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    T = (c_void_p * k)(*[c_void_p(int(x)) for x in stamps])
    _check(lib().tfd_push_multi_stamp(S, D, NS, c_int(n), T, c_int(k), _p(tag_ptr), _p(tag_cur), _p(block_counter),
                                      _p(local_step), _stream()), "tfd_push_multi_stamp")
    _count()


def pull_multi_tag(srcs, dsts, ns, versions, want_segs, tag_seg: int, tag_out) -> None:
    n = len(srcs)
    S = (c_void_p * n)(*[c_void_p(int(x)) for x in srcs])
    D = (c_void_p * n)(*[c_void_p(int(x)) for x in dsts])
    NS = (c_longlong * n)(*[int(x) for x in ns])
    V = (c_void_p * n)(*[c_void_p(int(x)) for x in versions])
    W = (c_void_p * n)(*[c_void_p(int(x)) for x in want_segs])
    _check(lib().tfd_pull_multi_tag(S, D, NS, V, W, c_int(n), c_int(tag_seg), _p(tag_out), _stream()),
           "tfd_pull_multi_tag")
    _count()


def debug_stamps():
    """(adam[8], conv1[4]) globaltimer stamps of the last Adam / conv1-forward kernel on the current GPU (synchronises)."""
    import ctypes
    a, c = (ctypes.c_ulonglong * 8)(), (ctypes.c_ulonglong * 4)()
    _check(lib().tfd_debug_adam_stamps(a), "tfd_debug_adam_stamps")
    _check(lib().tfd_debug_conv1_stamps(c), "tfd_debug_conv1_stamps")
    return list(a), list(c)


def memcpy_async(dst: int, src: int, nbytes: int) -> None:
    """Copy-engine transfer on the current stream (src may be a peer-mapped address); a memcpy node under capture."""
    _check(lib().tfd_memcpy_async(c_void_p(int(dst)), c_void_p(int(src)), c_size_t(int(nbytes)), _stream()),
           "tfd_memcpy_async")


# ---- symmetric heap ----
def ipc_alloc(nbytes: int):
    ptr = c_void_p()
    handle = ctypes.create_string_buffer(64)
    _check(lib().tfd_ipc_alloc(c_size_t(nbytes), ctypes.byref(ptr), handle), "tfd_ipc_alloc")
    return ptr.value, handle.raw


def ipc_open(handle: bytes) -> int:
    ptr = c_void_p()
    _check(lib().tfd_ipc_open(ctypes.c_char_p(handle), ctypes.byref(ptr)), "tfd_ipc_open")
    return ptr.value


def ipc_close(ptr: int) -> None:
    _check(lib().tfd_ipc_close(c_void_p(ptr)), "tfd_ipc_close")


def free(ptr: int) -> None:
    _check(lib().tfd_free(c_void_p(ptr)), "tfd_free")


def tensor_from_ptr(ptr: int, numel: int, dtype: torch.dtype, device) -> torch.Tensor:
    """Wrap raw device memory (e.g. a peer-mapped ipc pointer) as a torch tensor without copying."""
    esz = torch.empty((), dtype=dtype).element_size()

    class _Holder:
        pass

    h = _Holder()
    typestr = {torch.float32: "<f4", torch.bfloat16: "<u2", torch.uint8: "|u1", torch.int32: "<i4",
               torch.uint32: "<u4"}[dtype]
    h.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2,
                                  "strides": None}
    t = torch.as_tensor(h, device=device)
    if dtype == torch.bfloat16:
        t = t.view(torch.int16).view(torch.bfloat16)
    elif dtype == torch.uint32:
        t = t.view(torch.int32)
    assert t.data_ptr() == ptr and t.element_size() == esz or dtype == torch.uint32
    return t


# ---- parameter server over NVLS multicast (experimental; see csrc/nvls_kernels.cu) ----
def nvls_apply(w, m, v, mc_grad, mc_mirror_w, mc_mirror_shadow, pieces, scale: float, lr: float, t_counter, arrive,
               arrivals_per_step: int, mc_version, global_step=None, adam: bool = True, beta1: float = 0.9,
               beta2: float = 0.999, eps: float = 1e-8) -> None:
    arr = (NvlsPiece * max(1, len(pieces)))(*pieces)
    _check(lib().tfd_nvls_apply(_p(w), _p(m), _p(v), _p(mc_grad), _p(mc_mirror_w), _p(mc_mirror_shadow), arr,
                                c_int(len(pieces)), c_float(scale), c_float(lr), c_float(beta1), c_float(beta2),
                                c_float(eps), _p(t_counter), _p(arrive), c_uint(arrivals_per_step), _p(mc_version),
                                _p(global_step), c_int(int(adam)), _stream()), "tfd_nvls_apply")
    _count()


def nvls_gate(version, step, per_step: int, base: int = 0) -> None:
    _check(lib().tfd_nvls_gate(_p(version), _p(step), c_uint(per_step), c_uint(base), _stream()), "tfd_nvls_gate")
    _count()


def nvls_signal(mc_arrive, step) -> None:
    _check(lib().tfd_nvls_signal(_p(mc_arrive), _p(step), _stream()), "tfd_nvls_signal")
    _count()
