"""Dev tool: per-launch summary of selected metrics from an .ncu-rep (ncu -i ... --page raw --csv); also prints the summed DRAM
traffic of the build kernels in the form profiles/traffic_build.json keeps."""
import csv, io, json, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum"]
txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
h = rows[0]; units = rows[1]
ki = h.index("Kernel Name")
traffic = {}
for r in rows[2:]:
    name = r[ki].split("(")[0]
    print("---", name)
    for m in WANT:
        if m in h:
            i = h.index(m); print(f"  {m:78s} {r[i]} {units[i]}")
    try:
        rd = float(r[h.index("dram__bytes_read.sum")].replace(",", "")); wr = float(r[h.index("dram__bytes_write.sum")].replace(",", ""))
        ur = units[h.index("dram__bytes_read.sum")]; uw = units[h.index("dram__bytes_write.sum")]
        sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        traffic.setdefault(name, []).append(rd * sc[ur] + wr * sc[uw])
    except Exception:
        pass
print("# mean DRAM bytes per launch:", json.dumps({k: sum(v) / len(v) for k, v in traffic.items()}))
