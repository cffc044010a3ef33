"""Step time at batch 64 for a list of B200MNIST_PLAN overrides ("fc1_split,dg_split,c2w_split,c2dg_gx", 0 = keep), each in a
fresh process, UNROLL (default 8) steps per CUDA graph on a resident input ring:  python tools/sweep_plan.py 0,0,0,0 0,0,27,32"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
child = r'''
import os, sys
sys.path.insert(0, os.getcwd())
import torch
import pytorch_operator_b200 as pkg
dev = torch.device("cuda", 0); torch.cuda.set_device(dev)
B = 64
torch.manual_seed(1)
model = pkg.Net().to(dev)
opt = pkg.SGD(model.parameters(), lr=0.01, momentum=0.5)
g = torch.Generator(device=dev).manual_seed(7)
x = torch.randn(B, 1, 28, 28, device=dev, generator=g); t = torch.randint(0, 10, (B,), device=dev, generator=g)
rec = torch.empty(64, B*784*4 + B*8, dtype=torch.uint8, device=dev)
rec[:, :B*784*4].view(torch.float32).normal_(generator=g)
rec[:, B*784*4:].view(torch.int64).copy_(torch.randint(0, 10, (64, B), generator=g, device=dev))
ring = pkg.InputRing(B, records=rec)
model.train_step(x, t, opt, ring=ring)
U = int(os.environ.get("UNROLL", "8"))
gr = pkg.StepGraph(model, opt, B, x=x, target=t, ring=ring, unroll=U)
best = 1e9
for rep in range(3):
    for _ in range(300): gr.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(2000 // U): gr.replay()
    e1.record(); torch.cuda.synchronize()
    best = min(best, e0.elapsed_time(e1) / (2000 // U * U) * 1e3)
print("PLAN", os.environ.get("B200MNIST_PLAN"), "step_us", round(best, 2), "loss", float(gr.loss), flush=True)
'''
plans = sys.argv[1:] or ["0,0,0,0"]
for p in plans:
    env = dict(os.environ, B200MNIST_PLAN=p)
    r = subprocess.run([sys.executable, "-c", child], env=env, cwd=ROOT, capture_output=True, text=True)
    out = [l for l in r.stdout.splitlines() if l.startswith("PLAN")]
    print(out[0] if out else ("FAIL " + p + " " + r.stderr[-300:]), flush=True)
