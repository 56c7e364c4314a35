#!/usr/bin/env python
"""Summarise an `ncu --page source --csv --print-source sass` export (optionally .gz): stall-reason totals and the
instructions that collect the most warp-stall samples.   usage: tools/ncu_src_stalls.py file.csv[.gz] [top_n]"""
import csv, gzip, sys

def main():
    path = sys.argv[1]
    topn = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    op = gzip.open if path.endswith('.gz') else open
    rows = list(csv.reader(op(path, 'rt')))
    print(rows[0][1][:120])
    hdr = rows[1]
    si = hdr.index('# Samples')
    st = [i for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    seen, data = set(), []
    for r in rows[2:]:
        if len(r) <= si or not r[si].isdigit() or r[0] in seen:
            continue
        seen.add(r[0])
        data.append(r)
    tot = sum(int(r[si]) for r in data)
    agg = {}
    for r in data:
        for i in st:
            if r[i].isdigit():
                agg[hdr[i]] = agg.get(hdr[i], 0) + int(r[i])
    print('instructions', len(data), 'samples', tot)
    print('  '.join(f'{k[6:]} {v / tot:.1%}' for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v / tot > 0.004))
    for r in sorted(data, key=lambda r: -int(r[si]))[:topn]:
        reasons = sorted([(int(r[i]), hdr[i][6:]) for i in st if r[i].isdigit() and int(r[i]) > 0], reverse=True)[:2]
        print(r[0][-5:], r[si].rjust(6), f'{int(r[si]) / tot:6.1%}', r[1][:64].ljust(64), reasons)

main()
