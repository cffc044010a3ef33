import faulthandler, os, sys
faulthandler.dump_traceback_later(40, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
import pytorch_operator_b200 as pkg
rank = int(os.environ["RANK"]); dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
torch.cuda.set_device(dev)
dist.init_process_group("nccl", device_id=dev)
def log(*a): print(f"[r{rank}]", *a, flush=True)
torch.manual_seed(1)
net = pkg.Net().to(dev); model = pkg.DistributedDataParallel(net); opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
x = torch.randn(64, 1, 28, 28, device=dev); t = torch.randint(0, 10, (64,), device=dev)
log("eager step", float(model.train_step(x, t, opt)))
mode = sys.argv[1]
if mode == "plain":   # capture only an all_reduce
    g = torch.cuda.CUDAGraph(); s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        dist.all_reduce(net.flat_grads)
    torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize(); log("warm ok")
    with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
        dist.all_reduce(net.flat_grads)
    log("captured"); g.replay(); torch.cuda.synchronize(); log("replayed")
else:
    log("building StepGraph")
    sg = pkg.StepGraph(model, opt, 64)
    log("captured")
    sg.x.copy_(x); sg.target.copy_(t)
    l = sg.replay(); torch.cuda.synchronize(); log("replayed", float(l))
net.release_graphs()
if mode == 'plain': g.reset()
dist.barrier(); dist.destroy_process_group(); log("done")
