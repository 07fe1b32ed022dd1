"""Samples / instructions by SASS index range of an ncu report (source page): python tools/ncu_regions.py rep [bucket]"""
import csv, subprocess, sys
rep = sys.argv[1]; B = int(sys.argv[2]) if len(sys.argv) > 2 else 100
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines())); hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
tot = sum(int(r[ix['# Samples']]) for r in data)
for b in range(0, len(data), B):
    seg = data[b:b + B]
    s = sum(int(r[ix['# Samples']]) for r in seg); e = sum(int(r[ix['Instructions Executed']]) for r in seg)
    first = seg[0][ix['Source']].strip()[:40]
    print(f"{b:5d} samples {100*s/tot:5.1f}%  exec {e/1e6:9.1f}M  cyc/instr {s/max(e,1)*1e0:8.5f}  {first}")
