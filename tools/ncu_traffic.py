#!/usr/bin/env python3
"""profiles/traffic.json from an `ncu --set full` capture of one bench step.

usage: ncu -i X.ncu-rep --page raw --csv > X.csv ; ncu_traffic.py X.csv <workload> <capture name> [traffic.json]
Per captured kernel: time, DRAM read/write bytes, L2 and L1 hit rates, sectors per request of global
loads, occupancy, issue slots; per kernel GROUP of the library (cmsbwt_stats.group_ms): the sum of
dram__bytes_read.sum + dram__bytes_write.sum of the captured launches (bench.py's roofline.traffic)."""
import csv
import json
import sys

GROUPS = [("pack", ("k_pack",)), ("ms", ("k_ms_fast", "k_ms(")), ("gather", ("k_msf_gather", "k_msf_fix", "k_compact")),
          ("rank", ("k_link", "k_tail_", "Onesweep", "RadixSort", "k_tr_", "k_kappa", "k_rsort", "k_rs_")),
          ("tiles", ("k_tile_hist", "k_tile_scatter", "k_tile_cov", "k_tile_bin", "k_tile_deepest")),
          ("place", ("k_tile_place", "k_tile_rank", "k_term_syms", "k_place_sentinels")),
          ("offsets", ("k_vrank", "k_cnt_sa", "k_fill_part")), ("fill", ("k_fill",)), ("rle", ("k_rle",))]
COLS = [("gpu__time_duration.sum", "ms"), ("dram__bytes_read.sum", "rdGB"), ("dram__bytes_write.sum", "wrGB"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("l1tex__t_sector_hit_rate.pct", "L1hit%"),
        ("l1tex__average_t_sectors_per_request_pipe_lsu_mem_global_op_ld.ratio", "sect/req"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"),
        ("launch__registers_per_thread", "regs")]


def conv(v, u):
    try:
        v = float(v.replace(",", ""))
    except ValueError:
        return float("nan")
    return v * {"us": 1e-3, "usecond": 1e-3, "ns": 1e-6, "nsecond": 1e-6, "s": 1e3, "second": 1e3, "msecond": 1, "ms": 1,
                "Gbyte": 1, "Mbyte": 1e-3, "Kbyte": 1e-6, "byte": 1e-9}.get(u, 1)


def group_of(name):
    for g, pats in GROUPS:
        if any(p in name for p in pats):
            return g
    return None


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    h, units = rows[0], rows[1]
    kn = h.index("Kernel Name")
    idx = [(h.index(c) if c in h else None, nm) for c, nm in COLS]
    per_group = {}
    print("%-30s" % "kernel" + "".join("%9s" % nm for _, nm in COLS) + "   DRAM GB/s")
    for r in rows[2:]:
        vals = [conv(r[i], units[i]) if i is not None else float("nan") for i, _ in idx]
        name = r[kn].replace("void ", "")
        bw = (vals[1] + vals[2]) / (vals[0] / 1e3) if vals[0] else 0
        print("%-30s" % name[:29] + "".join("%9.3f" % v for v in vals) + "   %8.0f" % bw)
        g = group_of(name)
        if g:
            per_group[g] = per_group.get(g, 0.0) + (vals[1] + vals[2]) * 1e9
    if len(sys.argv) > 3:
        out = sys.argv[4] if len(sys.argv) > 4 else "profiles/traffic.json"
        try:
            tj = json.load(open(out))
        except Exception:
            tj = {}
        tj[sys.argv[2]] = {g: int(v) for g, v in per_group.items()}
        tj["_capture"] = sys.argv[3]
        tj["_comment"] = ("dram__bytes_read.sum + dram__bytes_write.sum per step, summed over the captured launches of each "
                          "kernel group (tools/ncu_traffic.py on the named ncu --set full capture of `bench.py --steps 1`)")
        json.dump(tj, open(out, "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
