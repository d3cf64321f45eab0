"""Segments an `ncu --page source --csv` export into runs of equal execution count (loop bodies) and
prints each run's share of executed instructions and stall samples.  Usage: ncu_src.py src.csv [lo hi]"""
import csv
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr = rows[1]
    ia, isrc, iss = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
    out = [(int(r[ia]), r[isrc].strip(), int(r[iss])) for r in rows[2:] if len(r) > ia]
    tot, stot = sum(o[0] for o in out), sum(o[2] for o in out)
    base = out[0][0] or 1
    if len(sys.argv) > 3:
        for i in range(int(sys.argv[2]), int(sys.argv[3])):
            print(i, round(out[i][0] / base, 2), out[i][2], out[i][1][:110])
        return
    segs = []
    for i, (n, s, smp) in enumerate(out):
        if segs and abs(segs[-1][2] - n) < 0.02 * max(n, 1):
            segs[-1][1] = i; segs[-1][3] += n; segs[-1][4] += smp
        else:
            segs.append([i, i, n, n, smp])
    print('total warp instructions', tot, 'per first-instruction execution', round(tot / base, 1), 'samples', stot)
    for a, b, n, t, smp in segs:
        if t / tot > 0.004 or smp / stot > 0.01:
            print(f'{a:4d}-{b:4d} x{n / base:6.2f} instrs {b - a + 1:4d} share {t / tot * 100:5.1f}% samples {smp / stot * 100:5.1f}%')


if __name__ == '__main__':
    main()
