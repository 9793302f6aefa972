"""`ncu --set full` report of the engine's dominant kernel -> profiles/rNN_conv_traffic.json, the file bench.py reads
`roofline.traffic` from (DRAM bytes per launch = dram__bytes_read.sum + dram__bytes_write.sum of the LAST profiled launch
of the kernel: warm L2 state, like inside the step).

Capture on the GPU box (one GPU, after the same command has exited 0 without ncu):
    ncu --set full --clock-control none --import-source on -k regex:tc_conv_halo -s 1 -c 2 -o gpurun_out/r02_conv \
        python tools/one_conv.py 96 96 5 64 1 fprop
then here:
    python tools/ncu_traffic.py gpurun_out/r02_conv.ncu-rep "fprop 96->96 5x5 @64x64, N=256" profiles/r02_conv_traffic.json
"""
import csv
import io
import json
import subprocess
import sys

rep, shape, out = sys.argv[1:4]
raw = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
SCALE = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "second": 1e6, "byte": 1.0,
         "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def val(r, name):
    if name not in col:
        return None
    try:
        return float(r[col[name]].replace(",", "")) * SCALE.get(units[col[name]], 1.0)
    except ValueError:
        return None


launches = [r for r in rows[2:] if len(r) >= len(hdr)]
r = launches[-1]
tensor = val(r, "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active")
if tensor is None:
    tensor = val(r, "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active")
d = {"kernel": r[col["Kernel Name"]][:80], "shape": shape,
     "dram_bytes_per_launch": (val(r, "dram__bytes_read.sum") or 0) + (val(r, "dram__bytes_write.sum") or 0),
     "dram_bytes_read": val(r, "dram__bytes_read.sum"), "dram_bytes_write": val(r, "dram__bytes_write.sum"),
     "duration_us": val(r, "gpu__time_duration.sum"), "tensor_pipe_pct": tensor,
     "l2_to_sm_bytes": val(r, "l1tex__m_xbar2l1tex_read_bytes.sum"), "launches_profiled": len(launches),
     "source": rep, "how": "ncu --set full --clock-control none, last profiled launch"}
json.dump(d, open(out, "w"), indent=1)
print(json.dumps(d))
