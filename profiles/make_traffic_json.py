#!/usr/bin/env python
"""profiles/r02_ncu_traffic.json from an `ncu --set full` report of `bench.py --no-graph` (one launch per kernel):
dram__bytes_read.sum + dram__bytes_write.sum per launch, summed over the kernels of each C-ABI entry point.
    python profiles/make_traffic_json.py gpurun_out/prof.ncu-rep cfg3 4096 > profiles/r02_ncu_traffic.json"""
import csv
import io
import json
import subprocess
import sys

ENTRY = [  # (kernel-name substring, entry point)
    ("frontend_bwd_tc_kernel", "xnn_frontend_bwd"), ("frontend_bwd_finalize", "xnn_frontend_bwd"),
    ("frontend_tc_kernel", "xnn_frontend_fwd"), ("tc_split_weights", "xnn_frontend_fwd"),
    ("mlp_fc1_kernel<0", "xnn_mlp_bn2_train_prepare"), ("w1_image_kernel", "xnn_mlp_bn2_train_prepare"),
    ("mlp_img_kernel<1", "xnn_mlp_fwd"), ("mlp_fc1_kernel<1", "xnn_mlp_fwd"),
    ("mlp_img_kernel<2", "xnn_mlp_bwd"), ("mlp_bwd_pipe_kernel", "xnn_mlp_bwd"), ("mlp_bwd_finalize", "xnn_mlp_bwd"),
    ("mlp_bwd_reduce", "xnn_mlp_bwd"), ("cooc_kernel", "xnn_cooc_counts"),
]
raw = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
ix = {h: i for i, h in enumerate(hdr)}
mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
per_kernel, seen = {}, set()
for r in rows[2:]:
    name = r[ix["Kernel Name"]]
    if name in seen:
        continue                     # first launch of every kernel
    seen.add(name)
    tot = 0.0
    for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
        tot += float(r[ix[m]].replace(",", "")) * mult.get(units[ix[m]], 1.0)
    per_kernel[name.split("(")[0]] = {"dram_bytes": tot, "us": float(r[ix["gpu__time_duration.sum"]].replace(",", ""))
                                       * {"us": 1.0, "ms": 1e3, "ns": 1e-3}.get(units[ix["gpu__time_duration.sum"]], 1.0)}
entry = {}
for k, v in per_kernel.items():
    for sub, e in ENTRY:
        if sub in k:
            entry[e] = entry.get(e, 0.0) + v["dram_bytes"]
            break
print(json.dumps({"workload": sys.argv[2], "per_gpu_batch": int(sys.argv[3]), "source": sys.argv[1].split("/")[-1],
                  "how": "ncu --set full --clock-control none, first launch of every kernel of one eager training step",
                  "dram_bytes_per_call": entry, "per_kernel": per_kernel}, indent=1))
