"""Dev tool: FP64 floating-point operations executed by each kernel of an .ncu-rep that was captured with --import-source on
(--page source --csv per kernel): DFMA = 2 flop per predicated-on thread instruction, DMUL / DADD = 1, DMMA m8n8k4 = 512 per
warp instruction.  usage: fp64_flops.py report.ncu-rep kernel_regex [launches_in_report_per_kernel]"""
import csv, io, json, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
nl = float(sys.argv[3]) if len(sys.argv) > 3 else 1.0
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:" + rx], capture_output=True, text=True).stdout
out = {}; cur = None; hdr = None
for r in csv.reader(io.StringIO(txt)):
    if not r: continue
    if r[0] == "Kernel Name": cur = r[1].split("(")[0]; out.setdefault(cur, {"dfma": 0, "dmul": 0, "dadd": 0, "dmma": 0}); hdr = None; continue
    if r[0] == "Address": hdr = r; ii = hdr.index("Instructions Executed"); it = hdr.index("Predicated-On Thread Instructions Executed"); continue
    if hdr is None or not r[0].startswith("0x"): continue
    ops = r[1].split()
    op = ops[1] if ops and ops[0].startswith("@") else (ops[0] if ops else "")
    base = op.split(".")[0]
    if base == "DFMA": out[cur]["dfma"] += int(r[it] or 0)
    elif base == "DMUL": out[cur]["dmul"] += int(r[it] or 0)
    elif base == "DADD": out[cur]["dadd"] += int(r[it] or 0)
    elif base == "DMMA": out[cur]["dmma"] += int(r[ii] or 0)
for k, v in out.items():
    v["flop"] = (2 * v["dfma"] + v["dmul"] + v["dadd"] + 512 * v["dmma"]) / nl
    v["dmma_share"] = 512 * v["dmma"] / nl / max(v["flop"], 1)
print(json.dumps(out, indent=1))
