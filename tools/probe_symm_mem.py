#!/usr/bin/env python
"""Does this host expose NVLS multicast to user processes?  (torch symmetric memory as the probe.)"""
import os
import sys

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import torch.distributed._symmetric_memory as sm
try:
    try:
        sm.enable_symm_mem_for_group(dist.group.WORLD.group_name)
    except Exception as e:      # noqa: BLE001
        print("enable_symm_mem_for_group:", e)
    t = sm.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
    h = sm.rendezvous(t, dist.group.WORLD.group_name)
    mc = getattr(h, "multicast_ptr", None)
    print(f"rank {rank}: buffer_ptrs={[hex(p) for p in h.buffer_ptrs]} multicast_ptr={hex(mc) if mc else mc} "
          f"signal_pads={len(h.signal_pad_ptrs)} backend={sm.get_backend(torch.device('cuda', local)) if hasattr(sm, 'get_backend') else '?'}")
except Exception as e:      # noqa: BLE001
    import traceback
    traceback.print_exc()
    print("symmetric memory unavailable:", e)
dist.destroy_process_group()
