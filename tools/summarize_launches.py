"""Turns an `ncu --metrics gpu__time_duration.sum --csv` launch list into the markdown table kept under profiles/.
usage: python tools/summarize_launches.py launches.csv "title" "command" > profiles/NAME.md"""
import collections
import csv
import io
import sys


def main():
    path, title, cmd = sys.argv[1], sys.argv[2], sys.argv[3]
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    agg = collections.defaultdict(lambda: [0.0, 0])
    tot, n = 0.0, 0
    for row in csv.DictReader(io.StringIO("".join(lines))):
        if row.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(row["Metric Value"].replace(",", ""))
        u = row["Metric Unit"]
        ms = v / 1e6 if u in ("ns", "nsecond") else v / 1e3 if u in ("us", "usecond") else v
        k = row["Kernel Name"]
        agg[k][0] += ms
        agg[k][1] += 1
        tot += ms
        n += 1
    own = sum(v[0] for k, v in agg.items() if "adkf::" in k)
    print(f"# {title}\n\nCommand (B200, after the same command exited 0 without ncu):\n\n```\n{cmd}\n```\n")
    print(f"{n} launches, {tot:.1f} ms summed device time (cold-cache, serialised -- compare SHARES). Own kernels (`adkf::*`): "
          f"{own:.1f} ms = {100 * own / tot:.1f}%.\n")
    print("| ms | share | launches | kernel |\n|---:|---:|---:|---|")
    for k, (ms, c) in sorted(agg.items(), key=lambda x: -x[1][0])[:28]:
        print(f"| {ms:.3f} | {100 * ms / tot:.2f}% | {c} | `{k[:110]}` |")


if __name__ == "__main__":
    main()
