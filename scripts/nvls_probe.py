#!/usr/bin/env python3
"""Is NVLS multicast exposed on this box (SURVEY §7.5 "probe once")?
torchrun --nproc-per-node N scripts/nvls_probe.py
Prints CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED and whether torch's symmetric memory hands out a
multicast pointer for a small buffer."""
import os

import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
rank, world = dist.get_rank(), dist.get_world_size()
out = {"rank": rank, "world": world}
try:
    from cuda.bindings import driver as drv
except Exception:  # noqa: BLE001
    from cuda import cuda as drv
try:
    drv.cuInit(0)
    err, dev = drv.cuDeviceGet(local)
    err, v = drv.cuDeviceGetAttribute(drv.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev)
    out["CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED"] = int(v)
    err, v = drv.cuDeviceGetAttribute(drv.CUdevice_attribute.CU_DEVICE_ATTRIBUTE_HANDLE_TYPE_FABRIC_SUPPORTED, dev)
    out["HANDLE_TYPE_FABRIC_SUPPORTED"] = int(v)
except Exception as e:  # noqa: BLE001
    out["driver_query_error"] = repr(e)[:200]
try:
    import torch.distributed._symmetric_memory as symm_mem

    t = symm_mem.empty(1024, dtype=torch.uint32, device=f"cuda:{local}")
    hdl = symm_mem.rendezvous(t, group=dist.group.WORLD)
    out["symm_mem_multicast_ptr"] = hex(int(getattr(hdl, "multicast_ptr", 0) or 0))
    out["symm_mem_buffer_ptrs"] = len(hdl.buffer_ptrs)
except Exception as e:  # noqa: BLE001
    out["symm_mem_error"] = repr(e)[:300]
gathered = [None] * world
dist.all_gather_object(gathered, out)
if rank == 0:
    for g in gathered:
        print(g, flush=True)
dist.destroy_process_group()
