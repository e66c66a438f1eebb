"""Turns an ncu launch list (--metrics gpu__time_duration.sum --csv) into a per-kernel table."""
import collections, csv, sys

def main(path, out=None, note=""):
    rows = list(csv.reader(l for l in open(path) if l.startswith('"')))
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        v = float(r[vi].replace(",", ""))
        v *= {"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6, "second": 1e6}.get(r[ui], 1.0)
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    lines = [f"# ncu launch list summary: {path}", "", note, "",
             "Per-launch times under ncu are cold-cache and serialised: compare SHARES, not absolutes.", "",
             "| kernel | launches | total us | share |", "|---|---:|---:|---:|"]
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        lines.append(f"| `{k[:110]}` | {a[0]} | {a[1]:.1f} | {100 * a[1] / tot:.2f}% |")
    text = "\n".join(lines) + "\n"
    if out:
        open(out, "w").write(text)
    print(text)

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else None, sys.argv[3] if len(sys.argv) > 3 else "")
