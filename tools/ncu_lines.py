#!/usr/bin/env python
"""Samples / executed instructions per CUDA source line from `ncu --page source --csv --print-source cuda,sass`."""
import csv, collections, sys
def main(path, topn=25):
    rows = list(csv.reader(open(path)))
    h = next(i for i, r in enumerate(rows) if r and r[0] == 'Line No')
    hdr = rows[h]
    iS = hdr.index('# Samples'); iE = hdr.index('Instructions Executed')
    sc = [i for i, c in enumerate(hdr) if c.startswith('stall_') and 'Not Issued' not in c]
    agg = collections.defaultdict(lambda: [0, 0, '', collections.Counter()])
    cur = ''
    for r in rows:
        if r and r[0] == 'File Path': cur = r[1].split('/')[-1]
        if len(r) < len(hdr): continue
        try: ln = int(r[0]); sm = int(r[iS] or 0); ex = int(r[iE] or 0)
        except ValueError: continue
        a = agg[(cur, ln)]; a[0] += sm; a[1] += ex; a[2] = r[1][:90]
        for i in sc:
            try: a[3][hdr[i][6:]] += int(r[i] or 0)
            except ValueError: pass
    print('samples', sum(a[0] for a in agg.values()), 'warp instructions', sum(a[1] for a in agg.values()))
    for ln, a in sorted(agg.items(), key=lambda x: -x[1][0])[:topn]:
        print(ln[0], ln[1], a[0], a[1], a[2].strip(), a[3].most_common(3))
if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25)
