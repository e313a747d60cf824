"""Top stalled SASS instructions of the first kernel in an .ncu-rep (warp-state sampling), with stall reasons.
usage: python tools/ncu_hot.py report.ncu-rep [top] [lo hi]   (lo/hi: also dump the instruction range)"""
import collections
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    # one table per profiled launch, each starting with its own header row
    heads = [i for i, r in enumerate(rows) if "# Samples" in r]
    seen = set()
    for t, hi in enumerate(heads):
        end = heads[t + 1] if t + 1 < len(heads) else len(rows)
        body = rows[hi + 1:end]
        sig = (len(body), tuple(tuple(r) for r in body[:50]))
        if sig in seen:                      # the page repeats a launch once per source view
            continue
        seen.add(sig)
        title = next((" ".join(r) for r in rows[max(0, hi - 3):hi] if r and "Kernel" in " ".join(r)), "")
        print("=== launch table {} of the report {}".format(len(seen), title[:110]))
        table(rows[hi], body, top)


def table(h, body, top):
    si, so = h.index("# Samples"), h.index("Source")
    st = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    data = [r for r in body if len(r) > si and (r[si] or "0").isdigit()]
    tot = sum(int(r[si] or 0) for r in data)
    print("total samples", tot, "instructions", len(data))
    if tot == 0:
        return
    b = collections.Counter()
    for i, r in enumerate(data):
        b[i // 100] += int(r[si] or 0)
    print("per-100-instruction buckets:", {k * 100: v for k, v in sorted(b.items()) if v})
    ranked = sorted(((int(r[si] or 0), i) for i, r in enumerate(data)), reverse=True)[:top]
    for n, i in ranked:
        r = data[i]
        print(i, n, "%.1f%%" % (100.0 * n / tot), r[so].strip()[:70], {h[j][6:]: r[j] for j in st if r[j] not in ("0", "")})
    if len(sys.argv) > 4:
        for i in range(int(sys.argv[3]), int(sys.argv[4])):
            print(i, data[i][si], data[i][so].strip()[:100])


if __name__ == "__main__":
    main()
