"""Summarise .ncu-rep captures (ncu --set full) into a markdown table for profiles/."""
import csv
import subprocess
import sys

WANT = [
    ("gpu__time_duration.sum", "time (us)"),
    ("launch__grid_size", "grid"),
    ("launch__registers_per_thread", "regs"),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM thr %"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM thr %"),
    ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "L2 thr %"),
    ("dram__bytes_read.sum", "DRAM rd (MB)"),
    ("dram__bytes_write.sum", "DRAM wr (MB)"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %"),
]


SCALE = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6,
         "byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3, "Tbyte": 1e6}


def rows_of(path):
    """One dict per profiled launch; times in us and byte counts in MB whatever unit ncu picked for the report."""
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = {}
        for h, u, v in zip(hdr, units, r):
            if u in SCALE and (h.endswith(".sum") and ("bytes" in h or "time_duration" in h)):
                try:
                    v = str(float(v.replace(",", "")) * SCALE[u])
                except ValueError:
                    pass
            d[h] = v
        yield d


def main(paths):
    print("| kernel | " + " | ".join(n for _, n in WANT) + " |")
    print("|---|" + "---|" * len(WANT))
    for p in paths:
        for r in rows_of(p):
            name = r.get("Kernel Name", "?").split("(")[0].replace("void <unnamed>::", "")
            vals = []
            for k, _ in WANT:
                v = r.get(k, "")
                try:
                    vals.append("{:.1f}".format(float(v.replace(",", ""))))
                except ValueError:
                    vals.append(v)
            print("| `{}` | ".format(name) + " | ".join(vals) + " |")


if __name__ == "__main__":
    main(sys.argv[1:])
