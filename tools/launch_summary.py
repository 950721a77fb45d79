"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: device time per kernel (name, grid)."""
import collections
import csv
import re
import sys


def main(path, top=40):
    with open(path) as f:
        lines = [l for l in f if not l.startswith('==')]
    agg = collections.defaultdict(lambda: [0, 0.0])
    samples = collections.defaultdict(list)
    for row in csv.DictReader(lines):
        name = re.sub(r'\(.*', '', row['Kernel Name'])
        v = float(row['Metric Value'].replace(',', ''))
        u = row['Metric Unit']
        v = v / 1e3 if u == 'ns' else v * 1e3 if u == 'ms' else v
        k = (name, row['Grid Size'])
        agg[k][0] += 1
        agg[k][1] += v
        samples[k].append(v)
    tot = sum(v[1] for v in agg.values())
    print(f"# {path}: {sum(v[0] for v in agg.values())} launches, {tot:.1f} us device time")
    # launches enqueued past convergence are gated no-ops (a few us): "live" = launches above 30% of the kernel's maximum
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
        mx = max(samples[k])
        live = [t for t in samples[k] if t > 0.3 * mx]
        print(f"{v[1]:11.1f} us {100 * v[1] / tot:5.1f}%  n={v[0]:5d}  live={len(live):5d} live_avg={sum(live) / len(live):9.1f} us  {k[0][:80]}  grid={k[1]}")


if __name__ == '__main__':
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
