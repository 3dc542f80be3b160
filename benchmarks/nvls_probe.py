#!/usr/bin/env python
"""Round-2 planning probe: does this box support NVLink multicast (NVLS), and does NCCL use it?
    torchrun --nproc-per-node 2 benchmarks/nvls_probe.py      (NCCL_DEBUG=INFO lines go to stderr)"""
import ctypes
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
import torch.distributed as dist

from distributed_llms_example_b200 import _lib

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
torch.zeros(1, device="cuda")
v = ctypes.c_int(-1)
_lib.check(_lib.load().b2_nvls_supported(ctypes.byref(v)))
print(f"rank {local}: CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED = {v.value}", flush=True)
if int(os.environ.get("WORLD_SIZE", "1")) > 1:
    os.environ["NCCL_DEBUG"] = "INFO"
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    x = torch.ones(64 << 20, device="cuda")
    for _ in range(3):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.destroy_process_group()
