"""Probe (torchrun): does torch's symmetric memory give us an NVLS multicast pointer here?"""
import os
import torch
import torch.distributed as dist

local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    import torch.distributed._symmetric_memory as symm_mem
    t = symm_mem.empty(4096, dtype=torch.int64, device=f"cuda:{local}")
    h = symm_mem.rendezvous(t, dist.group.WORLD)
    print(f"rank {dist.get_rank()}: multicast_ptr={getattr(h, 'multicast_ptr', None)} buffer_ptrs={[hex(p) for p in h.buffer_ptrs]} "
          f"signal_pads={len(h.signal_pad_ptrs)}", flush=True)
except Exception as e:  # noqa: BLE001
    print(f"rank {dist.get_rank()}: symmetric memory unavailable: {type(e).__name__}: {e}", flush=True)
dist.barrier()
dist.destroy_process_group()
