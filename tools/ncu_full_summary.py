"""Summarise an `ncu --set full` report: per launch, duration, DRAM bytes (dram__bytes_read.sum + dram__bytes_write.sum),
L2 sector traffic, tensor-pipe and L1TEX utilisation.   python tools/ncu_full_summary.py <file.ncu-rep> > profiles/...txt"""
import csv, io, subprocess, sys

rep = sys.argv[1]
WANT = {
    "gpu__time_duration.sum": "dur",
    "dram__bytes_read.sum": "dr",
    "dram__bytes_write.sum": "dw",
    "l1tex__m_xbar2l1tex_read_bytes.sum": "l2",
    "sm__inst_executed_pipe_tensor.sum": "tensor_inst",
    "sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pct",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active": "tensor_pct2",
    "l1tex__throughput.avg.pct_of_peak_sustained_active": "l1_pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_pct",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed": "dram_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_pct",
    "launch__grid_size": "grid",
    "launch__registers_per_thread": "regs",
}
out = subprocess.check_output(["ncu", "-i", rep, "--page", "raw", "--csv"], text=True)
rows = list(csv.reader(io.StringIO(out)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}
SCALE = {"ns": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "nsecond": 1e-3, "second": 1e6,
         "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def val(r, name):
    if name not in col:
        return None
    s = r[col[name]].replace(",", "")
    try:
        v = float(s)
    except ValueError:
        return None
    return v * SCALE.get(units[col[name]], 1.0)


print(f"# {rep}: ncu --set full --clock-control none; one row per profiled launch")
print(f"{'dur us':>8} {'DRAM MB':>8} {'DRAM GB/s':>9} {'L2>SM MB':>8} {'tensor%':>7} {'l1tex%':>6} {'l2%':>5} {'dram%':>5} "
      f"{'grid':>6} {'regs':>4}  kernel")
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[col["Kernel Name"]][:70]
    d = {k: val(r, m) for m, k in WANT.items()}
    dram = (d["dr"] or 0) + (d["dw"] or 0)
    t = d["tensor_pct"] if d["tensor_pct"] is not None else d["tensor_pct2"]
    dur = d["dur"] or 0.0
    print(f"{dur:8.1f} {dram/1e6:8.2f} {dram/1e3/max(dur,1e-9):9.1f} {(d['l2'] or 0)/1e6:8.1f} "
          f"{(t if t is not None else float('nan')):7.1f} {(d['l1_pct'] or 0):6.1f} {(d['l2_pct'] or 0):5.1f} "
          f"{(d['dram_pct'] or 0):5.1f} {int(d['grid'] or 0):6d} {int(d['regs'] or 0):4d}  {name}")
