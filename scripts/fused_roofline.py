"""Achieved fraction of roofline for the fused pull (peer-B GEMM) and fused push (peer-out wgrad) kernels.

Run with 2 ranks (one GPU each): rank 1 plays the ps GPU (owns the bf16 weights and the gradient slot), rank 0 the
worker.  Roofline time = max(FLOPs / measured bf16 peak, bytes over NVLink / measured per-direction peer bandwidth)
(B200_PROFILING.md: 770 GB/s measured peer copy; MEASURED_PEAKS.json for the GEMM peak).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 \
        scripts/fused_roofline.py > profiles/fused_roofline.json
"""
import json
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from tensorflow_distributed_b200.ops import native as nat  # noqa: E402

NVLINK_GBS = 770.0


def peaks():
    try:
        p = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return p["bf16_tflops_sustained"], p["hbm_gbs"], "measured"
    except Exception:
        return 1400.0, 6650.0, "fallback"


def timeit(fn, iters=30, warm=3, per_graph=10):
    """Median device time per launch (us).  The launches are replayed from a CUDA graph so the host-side cost of
    a ctypes call + tensor-map encoding (~15 us) does not mask kernels that only run for a few microseconds."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(per_graph):
            fn()
    g.replay()
    torch.cuda.synchronize()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(iters)]
    for s, e in evs:
        s.record(); g.replay(); e.record()
    torch.cuda.synchronize()
    ts = sorted(s.elapsed_time(e) for s, e in evs)
    return ts[len(ts) // 2] * 1e3 / per_graph


def main():
    tf32 = os.environ.get("TFD_PRECISION", "tf32") == "tf32"   # operand type of the GEMMs under test
    dt, esz = (torch.float32, 4) if tf32 else (torch.bfloat16, 2)
    rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda:%d" % local)
    dist.init_process_group("nccl", device_id=dev)
    nat.set_device(local)
    tflops, hbm, src = peaks()
    heap_bytes = 1 << 30
    ptr, handle = nat.ipc_alloc(heap_bytes)
    handles = [None, None]
    dist.all_gather_object(handles, handle)
    peer = nat.ipc_open(handles[1 - rank])
    dist.barrier()
    results = []
    if rank == 0:
        shapes = [("cnn fc1 (wd1)", 128, 1024, 3136, 8, 3), ("mlp layer 8192x8192", 128, 8192, 8192, 1, 1)]
        for name, M, N, K, fsplit, dsplit in shapes:
            x = torch.randn(M, K, device=dev).to(dt)
            dy = torch.randn(M, N, device=dev).to(dt)
            w_local = torch.randn(K, N, device=dev).to(dt)
            planes = torch.zeros(max(fsplit, 1), M, N, device=dev)
            dplanes = torch.zeros(max(dsplit, 1), M, K, device=dev)
            g_local = torch.zeros(K, N, device=dev)
            w_peer, slot_peer = peer, peer + (512 << 20)
            flops = 2.0 * M * N * K

            def fwd(bptr):
                nat.gemm_bf16_seg(x, 0, [bptr], 1, M, N, K, planes, mode=nat.EPI_STORE_F32, lda=K, ldb=N, ld_out=N,
                                  seg_mode=nat.SEG_B_K, seg_rows=[0, K], bn=64, splits=fsplit, split_stride=M * N, tf32=tf32)

            def dgrad(bptr):
                nat.gemm_bf16_seg(dy, 0, [bptr], 0, M, K, N, dplanes, mode=nat.EPI_STORE_F32, lda=N, ldb=N, ld_out=K,
                                  seg_mode=nat.SEG_B_N, seg_rows=[0, K], bn=64, splits=dsplit, split_stride=M * K, tf32=tf32)

            def wgrad(optr):
                nat.gemm_bf16_seg(x, 1, [dy.data_ptr()], 1, K, N, M, None, mode=nat.EPI_STORE_F32, lda=K, ldb=N,
                                  ld_out=N, seg_mode=nat.SEG_OUT_M, seg_rows=[0, K], out_ptrs=[optr], bn=128, tf32=tf32)

            for kind, fn, loc, rem, nbytes in (
                    ("pull-fused forward GEMM (B tiles TMA-loaded from the ps GPU)", fwd, w_local.data_ptr(), w_peer, K * N * esz),
                    ("pull-fused dgrad GEMM (B tiles TMA-loaded from the ps GPU)", dgrad, w_local.data_ptr(), w_peer, K * N * esz),
                    ("push-fused wgrad GEMM (tiles stored into the slot on the ps GPU)", wgrad, g_local.data_ptr(), slot_peer, K * N * 4)):
                t_loc = timeit(lambda: fn(loc))
                t_rem = timeit(lambda: fn(rem))
                peak = tflops * (0.5 if tf32 else 1.0)  # tf32 tensor throughput is half of bf16's
                roof_us = max(flops / (peak * 1e12), nbytes / (NVLINK_GBS * 1e9)) * 1e6
                roof_loc = max(flops / (peak * 1e12), nbytes / (hbm * 1e9)) * 1e6
                results.append({"dtype": "tf32" if tf32 else "bf16", "shape": name, "M": M, "N": N, "K": K, "kernel": kind, "nvlink_bytes": nbytes,
                                "flops": flops, "us_peer": round(t_rem, 2), "us_local": round(t_loc, 2),
                                "roofline_us_peer": round(roof_us, 2), "fraction_of_roofline_peer": round(roof_us / t_rem, 3),
                                "achieved_GBs_peer": round(nbytes / t_rem / 1e3, 1),
                                "roofline_us_local_hbm": round(roof_loc, 2),
                                "fraction_of_roofline_local": round(roof_loc / t_loc, 3)})
        sys.stdout.flush()
        print("JSON_BEGIN")
        print(json.dumps({"peaks": {"bf16_tflops_sustained": tflops, "hbm_gbs": hbm, "nvlink_gbs_per_dir": NVLINK_GBS,
                                    "source": src}, "results": results}, indent=1))
    torch.cuda.synchronize()
    dist.barrier()
    nat.ipc_close(peer)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
