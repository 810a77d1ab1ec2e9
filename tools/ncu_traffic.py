"""profiles/ncu_traffic.json from an `ncu --set full` capture of tools/profile_step.py: DRAM bytes (read + write) per frame of
the kernels of each stage, which bench.py reports as roofline.traffic.  Usage:
    python tools/ncu_traffic.py <report.ncu-rep> <config> <frames per launch> <capture name>"""
import csv
import io
import json
import os
import subprocess
import sys

rep, config, frames, name = sys.argv[1], sys.argv[2], int(sys.argv[3]), sys.argv[4]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
ki, ri, wi, ti = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
stage_of = {"seg_stream_kernel": "segment", "seg_emit_kernel": "segment", "cell_filter_kernel": "radius", "contact_work_kernel": "pairs"}
stage, kernels, seen = {}, {}, set()
for r in rows[2:]:
    k = r[ki]
    base = k.replace("void ", "").split("<")[0].split("(")[0]
    if base not in stage_of or k in seen:
        continue
    seen.add(k)                       # one launch per kernel: the capture holds one steady-state batch
    b = float(r[ri].replace(",", "")) * scale[units[ri]] + float(r[wi].replace(",", "")) * scale[units[wi]]
    kernels[k] = {"dram_bytes_per_frame": b / frames, "ncu_ms": float(r[ti].replace(",", "")) * {"ms": 1.0, "us": 1e-3, "ns": 1e-6, "s": 1e3}.get(units[ti], 1.0)}
    stage[stage_of[base]] = stage.get(stage_of[base], 0.0) + b / frames
path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "ncu_traffic.json")
try:
    allc = json.load(open(path))
except Exception:
    allc = {}
stream = [v["dram_bytes_per_frame"] for k, v in kernels.items() if "seg_stream_kernel" in k]
allc[str(config)] = {"capture": name, "frames_per_launch": frames, "stage_bytes_per_frame": stage,
                     "stream_kernel_bytes_per_frame": stream[0] if stream else None, "kernels": kernels}
json.dump(allc, open(path, "w"), indent=1)
print(json.dumps(allc[str(config)], indent=1))
