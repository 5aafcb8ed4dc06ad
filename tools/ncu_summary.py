"""Summarise an .ncu-rep (read on a CPU box with `ncu -i`) into a markdown table: one row per captured kernel with the handful of
numbers DESIGN / BASELINE cite (duration, tensor-pipe activity, DRAM bytes and throughput, registers, shared memory, grid).
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep [> profiles/ncu_x.md]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]
cols = [("gpu__time_duration.sum", "dur us", 1.0), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor %", 1.0),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1.0), ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1.0),
        ("dram__bytes_read.sum", "DRAM rd MB", 1.0), ("dram__bytes_write.sum", "DRAM wr MB", 1.0),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 %", 1.0), ("launch__grid_size", "grid", 1.0), ("launch__block_size", "block", 1.0),
        ("launch__registers_per_thread", "regs", 1.0), ("launch__shared_mem_per_block_dynamic", "dyn smem KB", 1.0),
        ("sm__warps_active.avg.per_cycle_active", "warps/SM", 1.0)]
units = dict(zip(hdr, rows[1]))


def find(name):
    for i, h in enumerate(hdr):
        if h == name or h.endswith(name):
            return i
    return None


idx = [(find(n), label, n) for n, label, _ in cols]
print("| kernel | " + " | ".join(label for _, label, _ in idx) + " |")
print("|---|" + "---:|" * len(idx))
ki = hdr.index("Kernel Name")
for r in rows[2:]:
    name = r[ki].replace("void ", "").replace("rtdc::", "").replace("bf16gemm::", "")
    name = name.split("(")[0]
    vals = []
    for i, label, n in idx:
        if i is None or r[i] == "":
            vals.append("-")
            continue
        v = float(r[i].replace(",", ""))
        u = units.get(hdr[i], "")
        if label == "dur us":
            v = v / 1000.0 if u in ("ns", "nsecond") else (v * 1000.0 if u in ("ms", "msecond") else v)
        if label.endswith("MB"):
            v = {"byte": v / 1e6, "Kbyte": v / 1e3, "Mbyte": v, "Gbyte": v * 1e3}.get(u, v)
        if label == "dyn smem KB":
            v = {"byte/block": v / 1e3, "Kbyte/block": v}.get(u, v)
        vals.append(f"{v:.1f}" if abs(v) < 1000 and v != int(v) else f"{v:.0f}")
    print(f"| `{name[:64]}` | " + " | ".join(vals) + " |")
