"""One line per captured launch of an .ncu-rep (ncu --set full): duration, DRAM bytes, issue-slot utilisation, active
lanes per instruction, L1 / L2 throughput and hit rates - the numbers DESIGN.md quotes.
usage: python profiles/wave_table.py REPORT.ncu-rep [--traffic-json WORKLOAD WAVES_PER_STEP]"""
import csv
import io
import json
import os
import subprocess
import sys

COLS = [("gpu__time_duration.sum", "dur_us", 1e-3), ("dram__bytes_read.sum", "dram_rd_MB", None), ("dram__bytes_write.sum", "dram_wr_MB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%", 1), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%", 1),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst", 1), ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1wave%", 1),
        ("l1tex__t_sector_hit_rate.pct", "l1hit%", 1), ("lts__t_sector_hit_rate.pct", "l2hit%", 1),
        ("lts__throughput.avg.pct_of_peak_sustained_elapsed", "l2%", 1), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%", 1),
        ("launch__registers_per_thread", "regs", 1), ("smsp__inst_executed.sum", "Minst", 1e-6)]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TO_NS = {"ns": 1.0, "us": 1e3, "ms": 1e6, "s": 1e9}


def main(path, extra):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    H, U = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(H)}
    print("%-22s" % "kernel" + "".join("%11s" % c[1] for c in COLS))
    tot_ns, tot_b, tot_l2 = 0.0, 0.0, 0.0
    w_issue = w_lanes = w_t = 0.0                                        # time-weighted means over the captured launches
    for r in rows[2:]:
        if "smsp__issue_active.avg.pct_of_peak_sustained_active" in ix:
            d = float(r[ix["gpu__time_duration.sum"]].replace(",", "")) * TO_NS.get(U[ix["gpu__time_duration.sum"]], 1.0)
            w_t += d
            w_issue += d * float(r[ix["smsp__issue_active.avg.pct_of_peak_sustained_active"]].replace(",", ""))
            w_lanes += d * float(r[ix["smsp__thread_inst_executed_per_inst_executed.ratio"]].replace(",", ""))
        if "lts__t_sectors.sum" in ix:                                   # L2-level traffic: every sector the L2 served, 32 B each
            tot_l2 += float(r[ix["lts__t_sectors.sum"]].replace(",", "")) * 32.0
        line = "%-22s" % r[ix["Kernel Name"]].split("(")[0].replace("void ", "").replace("prt::", "")[:22]
        for name, label, scale in COLS:
            if name not in ix:
                line += "%11s" % "-"; continue
            v = float(r[ix[name]].replace(",", "")); u = U[ix[name]]
            if label == "dur_us":
                v = v * TO_NS.get(u, 1.0); tot_ns += v; v *= 1e-3
            elif scale is None:
                v = v * TO_BYTES.get(u, 1.0); tot_b += v; v *= 1e-6
            else:
                v *= scale
            line += "%11.1f" % v
        print(line)
    print("total: %.3f ms (ncu: cold cache, serialised launches), %.1f MB DRAM traffic, %.1f MB L2-level traffic (lts__t_sectors.sum x 32 B)"
          % (tot_ns * 1e-6, tot_b * 1e-6, tot_l2 * 1e-6))
    if extra[:1] == ["--traffic-json"]:
        wl, waves = extra[1], float(extra[2])
        p = os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json")
        j = json.load(open(p)) if os.path.exists(p) else {}
        j[wl] = {"dram_bytes_per_step": tot_b * waves, "dram_bytes_per_wave": tot_b, "l2_bytes_per_step": tot_l2 * waves,
                 "issue_slot_pct": (w_issue / w_t) if w_t else None, "lanes_per_inst": (w_lanes / w_t) if w_t else None,
                 "source": "profiles/wave_table.py %s: dram__bytes_read.sum+dram__bytes_write.sum over the captured launches of one wave x %g waves per step; issue-slot %% and lanes/instruction are time-weighted means of the same launches"
                           % (os.path.basename(path), waves)}
        json.dump(j, open(p, "w"), indent=1)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
