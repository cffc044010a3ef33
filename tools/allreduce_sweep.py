"""BASELINE config 4: smoke-dist send/recv + all_reduce bit-exactness and the message-size sweep (16 B ... 256 MiB,
plus the 1.72 MB gradient bucket).  Launch with the env-contract launcher or torchrun; rank 0 writes a JSON file."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from pytorch_operator_b200 import dist_sendrecv
rank = int(os.environ["RANK"]); ws = int(os.environ["WORLD_SIZE"])
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", rank)) % torch.cuda.device_count()); torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
inp, results = dist_sendrecv.run()
sizes = [4, 256, 4096, 65536, 431296, 1 << 20, 1 << 22, 1 << 24, 1 << 26]
rows = dist_sendrecv.allreduce_check(sizes=sizes, iters=20)
if rank == 0:
    out = {"world_size": ws, "sendrecv_bit_exact": all(torch.equal(r, torch.mul(inp, inp)) for r in results), "allreduce": rows}
    json.dump(out, open(sys.argv[1], "w"), indent=1)
    print(json.dumps(out))
dist.barrier(); dist.destroy_process_group()
