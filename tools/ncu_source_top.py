"""Top stall sites of a kernel from `ncu -i X.ncu-rep --page source --csv` output:
    python tools/ncu_source_top.py X.src.csv [N]"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
    hdr = rows[1]
    ia, isrc, ismp, iex = hdr.index('Address'), hdr.index('Source'), hdr.index('# Samples'), hdr.index('Instructions Executed')
    st = [(i, h) for i, h in enumerate(hdr) if h.startswith('stall_') and 'Not Issued' not in h]
    data = [r for r in rows[2:] if len(r) > ismp and r[ismp].isdigit()]
    tot = sum(int(r[ismp]) for r in data)
    print('total samples', tot, 'instructions', len(data))
    agg = {}
    for r in data:
        for i, h in st:
            if r[i].isdigit():
                agg[h] = agg.get(h, 0) + int(r[i])
    print('by reason:', sorted(agg.items(), key=lambda kv: -kv[1])[:8])
    for r in sorted(data, key=lambda r: -int(r[ismp]))[:n]:
        reasons = sorted([(int(r[i]) if r[i].isdigit() else 0, h) for i, h in st], reverse=True)[:2]
        print(r[ia][-5:], r[ismp].rjust(6), r[iex].rjust(9), r[isrc][:64].ljust(64), reasons)


if __name__ == '__main__':
    main()
