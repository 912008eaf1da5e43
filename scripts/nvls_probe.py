"""NVLS capability probe, run under torchrun on >= 2 GPUs: does this box give torch symmetric memory a multicast
(NVSwitch) address?  Prints one line per rank; the output is kept under profiles/ as the evidence for / against the
multicast transport (parallel/ps_nvls.py)."""
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    device = torch.device("cuda:%d" % local)
    dist.init_process_group("nccl", device_id=device)
    info = {"rank": rank, "world": world}
    try:
        from cuda.bindings import driver as cu
        (err,) = cu.cuInit(0)
        err, dev = cu.cuDeviceGet(local)
        err, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
        info["cu_multicast_supported"] = int(v)
        err, v = cu.cuDeviceGetAttribute(cu.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev)
        info["cu_fabric_handles"] = int(v)
    except Exception as e:  # noqa: BLE001
        info["cuda_python_error"] = repr(e)[:200]
    try:
        import torch.distributed._symmetric_memory as symm
        buf = symm.empty(1 << 20, dtype=torch.uint8, device=device)
        buf.zero_()
        torch.cuda.synchronize()
        h = symm.rendezvous(buf, dist.group.WORLD)
        info["symm_has_multicast"] = bool(getattr(h, "has_multicast_support", False))
        info["symm_multicast_ptr"] = hex(int(getattr(h, "multicast_ptr", 0) or 0))
        info["symm_buffer_ptrs"] = [hex(int(p)) for p in h.buffer_ptrs]
    except Exception as e:  # noqa: BLE001
        info["symm_error"] = repr(e)[:300]
    print("NVLS_PROBE %s" % info, flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    sys.exit(main())
