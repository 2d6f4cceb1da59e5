"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum[,dram__bytes_read.sum,dram__bytes_write.sum] --csv` launch
list: launches, total time, share of device time, mean DRAM bytes per launch.
    python tools/launch_summary.py profiles/r02_launches_eager_2steps.csv [--json profiles/r02_dram_traffic_per_launch.json]
The JSON holds the per-launch DRAM traffic of the tcgen05 GEMM, block_ffn and attention kernels over the ESTIMATOR launches
only (front-end and vocoder launches of the same kernels are listed separately), which is what bench.py's roofline.traffic reads."""
import collections
import csv
import json
import re
import sys


def short(name: str) -> str:
    name = re.sub(r"^void\s+", "", name).replace("(anonymous namespace)::", "").replace("<unnamed>::", "")
    name = re.sub(r"\(.*$", "", name)
    name = re.sub(r"<.*$", "", name)
    return name.split("::")[-1][:48]


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    iid, ik, im, iv, ig = hdr.index("ID"), hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Grid Size")
    launches = collections.OrderedDict()
    for r in rows[1:]:
        d = launches.setdefault(int(r[iid]), dict(name=short(r[ik]), grid=r[ig]))
        d[r[im]] = float(r[iv].replace(",", ""))
    agg = collections.OrderedDict()
    for d in launches.values():
        a = agg.setdefault(d["name"], dict(n=0, ns=0.0, rd=0.0, wr=0.0))
        a["n"] += 1
        a["ns"] += d.get("gpu__time_duration.sum", 0.0)
        a["rd"] += d.get("dram__bytes_read.sum", 0.0)
        a["wr"] += d.get("dram__bytes_write.sum", 0.0)
    tot = sum(a["ns"] for a in agg.values())
    print(f"# {path}: {len(launches)} launches, {tot / 1e6:.2f} ms of kernel time (serialised, cold caches: compare shares)")
    print(f"{'kernel':50s} {'launches':>8s} {'ms':>9s} {'share':>7s} {'us/launch':>10s} {'DRAM rd MB':>11s} {'DRAM wr MB':>11s}")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ns"]):
        print(f"{k:50s} {a['n']:8d} {a['ns'] / 1e6:9.3f} {100 * a['ns'] / tot:6.1f}% {a['ns'] / a['n'] / 1e3:10.1f} "
              f"{a['rd'] / a['n'] / 1e6:11.2f} {a['wr'] / a['n'] / 1e6:11.2f}")
    if "--json" in sys.argv:
        out = {}
        # the estimator's launches: those between the first and the last block_ffn / attention launch of the list
        ids = [i for i, d in launches.items() if d["name"].startswith(("block_ffn", "attention_tc"))]
        lo, hi = min(ids), max(ids)
        for key, pref in (("gemm", "tapgemm_tc"), ("block_ffn", "block_ffn"), ("attn", "attention_tc")):
            sel = [d for i, d in launches.items() if lo - 8 <= i <= hi and d["name"].startswith(pref)]
            if not sel:
                continue
            out[key] = dict(launches=len(sel),
                            dram_read_bytes_per_launch=sum(d.get("dram__bytes_read.sum", 0.0) for d in sel) / len(sel),
                            dram_write_bytes_per_launch=sum(d.get("dram__bytes_write.sum", 0.0) for d in sel) / len(sel),
                            seconds_per_launch_under_ncu=sum(d["gpu__time_duration.sum"] for d in sel) / len(sel) * 1e-9)
        with open(sys.argv[sys.argv.index("--json") + 1], "w") as f:
            json.dump(out, f, indent=1)
        print(json.dumps(out))


if __name__ == "__main__":
    main()
