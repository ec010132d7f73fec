"""Per-source-line stall samples from `ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > X.csv`.
usage: python tools/ncu_lines.py X.csv [launch index] [min samples]"""
import csv, sys, collections

def num(x):
    try: return int(float(x.replace(',', '')))
    except Exception: return 0

def main():
    path = sys.argv[1]; which = int(sys.argv[2]) if len(sys.argv) > 2 else 0; mins = int(sys.argv[3]) if len(sys.argv) > 3 else 1000
    rows = list(csv.reader(open(path)))
    blocks = []; cur = None
    for r in rows:
        if r and r[0] == 'File Path': cur = {'file': r[1], 'rows': []}; blocks.append(cur)
        elif r and r[0] == 'Function Name': cur['fn'] = r[1]
        elif r and r[0] == 'Line No': cur['hdr'] = r
        elif cur is not None and r: cur['rows'].append(r)
    # launches are concatenated: split when a file name repeats
    launches = []; seen = set(); cur = []
    for b in blocks:
        if b['file'] in seen: launches.append(cur); cur = []; seen = set()
        seen.add(b['file']); cur.append(b)
    launches.append(cur)
    L = launches[which]
    print('launches:', len(launches), 'function:', L[0].get('fn', '')[:80])
    tot = 0; out = []
    for b in L:
        h = b['hdr']; iS = h.index('# Samples'); iI = h.index('Instructions Executed')
        sc = [i for i, x in enumerate(h) if x.startswith('stall_') and 'Not Issued' not in x]
        f = b['file'].split('/')[-1]
        for r in b['rows']:
            if not r[0].strip().isdigit(): continue
            s = num(r[iS]); tot += s
            st = sorted([(num(r[i]), h[i][6:]) for i in sc], reverse=True)[:3]
            out.append((f, int(r[0]), s, num(r[iI]), r[1].strip()[:80], st))
    print('total samples', tot)
    for f, ln, s, n, src, st in sorted(out, key=lambda x: (x[0], x[1])):
        if s >= mins:
            print(f"{f[:14]:14s}{ln:>4} {s:8d} {100.0*s/tot:5.1f}% inst={n:>11d} {src:80s} {[(a,b) for a,b in st if a]}")

main()
