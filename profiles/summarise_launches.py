"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list: per-kernel count, total and share.
    python profiles/summarise_launches.py launches.csv [skip [count]]"""
import csv, sys, collections, re
def main(path, skip=0, count=None):
    rows = []
    with open(path) as f:
        lines = [l for l in f if not l.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") == "gpu__time_duration.sum":
            v = float(r["Metric Value"].replace(",", "")); u = r["Metric Unit"]
            ns = v * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(u, 1)
            rows.append((re.sub(r"\(.*", "", r["Kernel Name"])[:70], ns))
    rows = rows[skip:] if count is None else rows[skip:skip + count]
    agg = collections.OrderedDict()
    for k, ns in rows:
        c, t = agg.get(k, (0, 0.0)); agg[k] = (c + 1, t + ns)
    tot = sum(t for _, t in agg.values())
    print(f"# {path}: {len(rows)} launches, {tot/1e6:.3f} ms total (cold-cache, serialised: compare shares)")
    print(f"{'kernel':70s} {'n':>5s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}")
    for k, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:70s} {c:5d} {t/1e6:10.3f} {t/c/1e3:10.1f} {100*t/tot:6.1f}%")
if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 0, int(sys.argv[3]) if len(sys.argv) > 3 else None)
