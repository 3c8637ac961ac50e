"""Per-kernel table of ONE LBVH build from an `ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,
dram__bytes_write.sum,lts__t_sectors.sum --csv` log of profiles/build_only.py (three builds; the LAST one is tabulated).

usage: python profiles/build_table.py gpurun_out/r1_build10m_ncu.csv N_PRIMS > profiles/r1_build10m_table.txt

The algorithmic figure beside it is SURVEY 8(d)'s ~250 B per primitive (bounds 76, sort 64, hierarchy + refit ~110).
"""
import csv
import re
import sys
from collections import OrderedDict


def to_bytes(v, unit):
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    return float(v.replace(",", "")) * scale.get(unit, 1)


def main():
    path, n = sys.argv[1], int(sys.argv[2])
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr, rows = rows[0], rows[1:]
    ix = {h: i for i, h in enumerate(hdr)}
    launches = OrderedDict()
    for r in rows:
        L = launches.setdefault(int(r[ix["ID"]]), {"name": re.sub(r"\(.*", "", r[ix["Kernel Name"]]).replace("prt::", "")})
        m, u, v = r[ix["Metric Name"]], r[ix["Metric Unit"]], r[ix["Metric Value"]]
        if m == "gpu__time_duration.sum":
            L["us"] = float(v.replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}.get(u, 1e-3)
        elif m.startswith("dram__bytes"):
            L[m.split(".")[0][-4:].strip("_")] = to_bytes(v, u)
        elif m == "lts__t_sectors.sum":
            L["l2"] = float(v.replace(",", "")) * 32
    L = list(launches.values())
    # the last build = everything from the last k_init_bounds on
    start = max(i for i, l in enumerate(L) if l["name"].startswith("k_init_bounds"))
    build = L[start:]
    agg = OrderedDict()
    for l in build:
        a = agg.setdefault(l["name"], {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0, "l2": 0.0})
        a["n"] += 1
        a["us"] += l.get("us", 0.0)
        a["rd"] += l.get("read", 0.0)
        a["wr"] += l.get("rite", l.get("write", 0.0))
        a["l2"] += l.get("l2", 0.0)
    print(f"LBVH build of {n} primitives, last of three builds (ncu: serialised launches, cold cache)")
    print(f"{'kernel':<22}{'launches':>9}{'us':>10}{'dram_rd_MB':>12}{'dram_wr_MB':>12}{'B/prim':>9}{'GB/s':>9}{'l2_MB':>10}")
    tot = {"us": 0.0, "rd": 0.0, "wr": 0.0, "l2": 0.0}
    for k, a in agg.items():
        b = a["rd"] + a["wr"]
        print(f"{k:<22}{a['n']:>9}{a['us']:>10.1f}{a['rd'] / 1e6:>12.1f}{a['wr'] / 1e6:>12.1f}{b / n:>9.1f}"
              f"{(b / (a['us'] * 1e-6) / 1e9 if a['us'] else 0):>9.0f}{a['l2'] / 1e6:>10.1f}")
        for t in tot:
            tot[t] += a[t]
    b = tot["rd"] + tot["wr"]
    print(f"{'total':<22}{len(build):>9}{tot['us']:>10.1f}{tot['rd'] / 1e6:>12.1f}{tot['wr'] / 1e6:>12.1f}{b / n:>9.1f}"
          f"{b / (tot['us'] * 1e-6) / 1e9:>9.0f}{tot['l2'] / 1e6:>10.1f}")
    print(f"algorithmic (SURVEY 8d): 250 B/prim = {250 * n / 1e6:.0f} MB; measured DRAM traffic / algorithmic = {b / (250.0 * n):.2f}")


if __name__ == "__main__":
    main()
