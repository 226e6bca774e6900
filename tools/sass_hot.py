"""Dev tool: join an ncu SASS source page (--page source --csv) with nvdisasm -g line info and report
instructions executed / stall samples per source line.  usage: sass_hot.py ncu_sass.csv cubin kernel_substr"""
import csv, re, subprocess, sys, collections
src_csv, cubin, kern = sys.argv[1:4]
rows = list(csv.reader(open(src_csv)))
hdr = rows[1]; ia = hdr.index('Address'); ii = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
data = [(int(r[ia], 16), int(r[ii] or 0), int(r[isamp] or 0), r[1]) for r in rows[2:] if len(r) > ii and r[ia].startswith('0x')]
base = data[0][0]
dis = subprocess.run(['nvdisasm', '-g', '-c', cubin], capture_output=True, text=True).stdout.splitlines()
inside = False; line = ('?', 0); off2line = {}
for l in dis:
    if l.startswith('//---') and '.text.' in l:
        inside = kern in l and 'slowpath' not in l
        continue
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m: line = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]+)\*/', l)
    if m: off2line[int(m.group(1), 16)] = line
agg = collections.defaultdict(lambda: [0, 0])
tot_i = tot_s = 0
for a, n, s, _ in data:
    k = off2line.get(a - base, ('?', 0))
    agg[k][0] += n; agg[k][1] += s; tot_i += n; tot_s += s
byfile = collections.defaultdict(lambda: [0, 0])
for (f, l), (n, s) in agg.items(): byfile[f][0] += n; byfile[f][1] += s
print("total inst", tot_i, "samples", tot_s)
for f, (n, s) in sorted(byfile.items(), key=lambda x: -x[1][1]): print(f"  {f:28s} inst {100*n/tot_i:5.1f}%  samples {100*s/max(tot_s,1):5.1f}%")
print("top lines by samples:")
for (f, l), (n, s) in sorted(agg.items(), key=lambda x: -x[1][1])[:int(sys.argv[4]) if len(sys.argv) > 4 else 30]:
    print(f"  {f}:{l:<5d} inst {100*n/tot_i:5.1f}%  samples {100*s/max(tot_s,1):5.1f}%")
