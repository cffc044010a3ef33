"""Debug aid: per-tensor gradient error of the CUDA step vs the torch oracle for several B."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import pytorch_operator_b200 as pkg
from oracle import mnist_oracle as mo

def main():
    for B in [int(a) for a in sys.argv[1:]] or [1, 2, 3, 4, 5, 6, 7, 8, 64, 65]:
        torch.manual_seed(1)
        net = pkg.Net().to("cuda:0")
        x, t = mo.synthetic_batch(B, 500 + B)
        loss = float(net.forward_backward(x.cuda(), t.cuda()))
        lo, _, g = mo.loss_and_grads(mo.init_params(1), x, t)
        errs = []
        for n, v in net.named_views(net.flat_grads).items():
            ref = g[n].numpy(); e = np.abs(v.cpu().numpy() - ref).max() / np.abs(ref).max()
            errs.append(f"{n.replace('.weight','.w').replace('.bias','.b')}={e:.1e}")
        print(B, f"loss {abs(loss-float(lo)):.1e}", " ".join(errs), flush=True)
main()
