"""Summarises an `ncu --metrics gpu__time_duration.sum --csv` launch list into kernel, launches, total ms, share."""
import csv, re, sys
from collections import defaultdict

def main(path, out, header):
    tot, cnt = defaultdict(float), defaultdict(int)
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ki, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
    for r in rows[1:]:
        name = re.sub(r"\(.*$", "", r[ki]).replace("void ", "")
        ns = float(r[vi].replace(",", "")) * {"ns": 1.0, "us": 1e3, "ms": 1e6}.get(r[ui], 1.0)
        tot[name] += ns
        cnt[name] += 1
    total = sum(tot.values())
    with open(out, "w") as f:
        f.write(header + "\n")
        for k in sorted(tot, key=tot.get, reverse=True):
            f.write(f"{k:<90} {cnt[k]:>5} {tot[k] / 1e6:>10.3f} ms {100 * tot[k] / total:>6.2f} %\n")

if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3])
