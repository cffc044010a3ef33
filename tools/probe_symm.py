import faulthandler, os, sys
faulthandler.dump_traceback_later(50, exit=True)
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
def log(*a): print(f"[r{rank}]", *a, flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    t = symm.empty(1024, dtype=torch.float32, device=dev)
    h = symm.rendezvous(t, dist.group.WORLD)
    log("rendezvous ok", type(h).__name__, "rank", h.rank, "ws", h.world_size)
    log("buffer_ptrs", [hex(p) for p in h.buffer_ptrs], "signal_pad_ptrs", [hex(p) for p in h.signal_pad_ptrs], "pad size", getattr(h, "signal_pad_size", None))
    t.fill_(rank + 1); torch.cuda.synchronize(); dist.barrier()
    peer = h.get_buffer((rank + 1) % h.world_size, (1024,), torch.float32)
    log("peer read", float(peer[0]))
    h.barrier(channel=0); log("handle barrier ok")
except Exception as e:
    import traceback; traceback.print_exc(); log("symm mem FAILED", repr(e))
log("can_access_peer", torch.cuda.can_device_access_peer(0, 1))
dist.barrier(); dist.destroy_process_group(); log("done")
