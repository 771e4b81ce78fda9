"""Turns ncu CSV exports into the short tables kept under profiles/.

    python tools/ncu_summarize.py launches <launches.csv>       # --metrics gpu__time_duration.sum --csv log
    python tools/ncu_summarize.py full <raw.csv>                 # ncu -i x.ncu-rep --page raw --csv
    python tools/ncu_summarize.py traffic <raw.csv> <kernel regex> <out.json> [note [batch kv_heads_local context_len]]
        # dram__bytes_read.sum + dram__bytes_write.sum per launch of the matching kernels -> the
        # file bench.py's roofline.traffic reads (profiles/paged_decode_ncu_traffic.json)
"""
import collections
import csv
import re
import sys


def launches(path):
    lines = [l for l in open(path) if not l.startswith("==")]
    agg = collections.OrderedDict()
    for row in csv.DictReader(lines):
        name = re.sub(r"\(.*", "", row["Kernel Name"]).replace("void ", "").replace("<unnamed>::", "")
        v = float(row["Metric Value"].replace(",", ""))
        v = v / 1000 if row["Metric Unit"] == "ns" else (v * 1000 if row["Metric Unit"] == "ms" else v)
        key = (name, row["Grid Size"], row["Block Size"])
        a = agg.setdefault(key, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print("| kernel | grid | block | launches | total us | avg us | share |")
    print("|---|---|---|---:|---:|---:|---:|")
    for (name, grid, block), (n, t) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name[:60]}` | {grid} | {block} | {n} | {t:.1f} | {t / n:.2f} | {t / tot:.3f} |")
    print(f"\ntotal {tot:.1f} us over {sum(a[0] for a in agg.values())} launches "
          "(cold-cache, serialised under ncu: compare shares, not absolutes)")


def full(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
            "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
            "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "launch__registers_per_thread", "launch__waves_per_multiprocessor", "lts__t_bytes.sum"]
    idx = [(w, hdr.index(w)) for w in want if w in hdr]
    print("| " + " | ".join(w.split(".")[0].replace("__", " ") for w, _ in idx) + " |")
    print("|" + "---|" * len(idx))
    for d in data:
        cells = []
        for w, i in idx:
            v = d[i]
            if w == "Kernel Name":
                v = "`" + re.sub(r"\(.*", "", v).replace("void ", "").replace("<unnamed>::", "")[:48] + "`"
            else:
                v = f"{v} {units[i]}".strip()
            cells.append(v)
        print("| " + " | ".join(cells) + " |")


def _bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[unit]


def traffic(path, pattern, out, note="", batch=None, kv_heads_local=None, context_len=None):
    import json
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik, ir, iw, it = (hdr.index(n) for n in ("Kernel Name", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__time_duration.sum"))
    sel = [d for d in data if re.search(pattern, d[ik])]
    rd = [_bytes(d[ir], units[ir]) for d in sel]
    wr = [_bytes(d[iw], units[iw]) for d in sel]
    json.dump({"kernel": pattern, "launches": len(sel), "grid": sel[0][hdr.index("Grid Size")],
               "dram_bytes_read_per_launch": sum(rd) / len(rd), "dram_bytes_write_per_launch": sum(wr) / len(wr),
               "dram_bytes_per_launch": (sum(rd) + sum(wr)) / len(sel),
               "gpu_time_us_under_ncu": [float(d[it].replace(",", "")) for d in sel], "note": note,
               # the shape the capture was taken on: bench.py's roofline.traffic only uses a capture of its own shape
               "batch": None if batch is None else int(batch),
               "kv_heads_local": None if kv_heads_local is None else int(kv_heads_local),
               "context_len": None if context_len is None else int(context_len)},
              open(out, "w"), indent=1)
    print(open(out).read())


if __name__ == "__main__":
    {"launches": launches, "full": full, "traffic": traffic}[sys.argv[1]](*sys.argv[2:])
