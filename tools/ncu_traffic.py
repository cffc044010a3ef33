"""Write profiles/r01_traffic.json: DRAM bytes (read + write) per launch of each kernel of the step, from ncu
--set full reports.  usage: python tools/ncu_traffic.py batch64=gpurun_out/prof_final.ncu-rep batch4096=..."""
import csv, io, json, os, subprocess, sys
LABEL = {  # ncu kernel base name -> bench.py label
    "conv1_relu_pool_fwd_kernel": "conv1_relu_pool_fwd", "conv2_relu_pool_fwd_kernel": "conv2_relu_pool_fwd",
    "gemm_kernel": "fc1 forward gemm", "head_kernel": "fc2/log_softmax/nll head",
    "fc_bwd_kernel": "fc backward (fc1 wgrad + fc1 dgrad + fc2 wgrad)", "conv2_bwd_kernel": "conv2 backward (wgrad + dgrad)",
    "conv1_wgrad_kernel": "conv1 wgrad", "sgd_fused_kernel": "sgd momentum", "grad_finalize_kernel": "grad finalize",
    "gemm_tc_kernel<0, 0, 0>": "fc1 forward gemm (tcgen05 3xTF32)", "gemm_tc_kernel<0, 1, 1>": "fc1 dgrad gemm (tcgen05 3xTF32)",
    "gemm_tc_kernel<1, 1, 2>": "fc1 wgrad gemm (tcgen05 3xTF32)"}
UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
out = {}
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "r01_traffic.json")
if os.path.exists(path):
    out = json.load(open(path))
for arg in sys.argv[1:]:
    key, rep = arg.split("=")
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw))); h, u = rows[0], rows[1]
    ir, iw, ik = h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum"), h.index("Kernel Name")
    acc = {}
    for r in rows[2:]:
        name = r[ik].replace("void ", "").split("(")[0]
        label = LABEL.get(name) or (LABEL["gemm_kernel"] if name.startswith("gemm_kernel<") else None)
        if not label: continue
        b = float(r[ir]) * UNIT[u[ir]] + float(r[iw]) * UNIT[u[iw]]
        acc.setdefault(label, []).append(b)
    out[key] = {k: sum(v) / len(v) for k, v in acc.items()}
json.dump(out, open(path, "w"), indent=1, sort_keys=True)
print(json.dumps(out, indent=1))
