"""ncu -i X.ncu-rep --page raw --csv > raw.csv ; python tools/ncu_summary.py raw.csv [--json out.json --source NAME]
Prints a markdown table of the captured launches and (with --json) writes profiles/roofline_traffic.json:
dram bytes per launch of each kernel (the LAST captured launch of a name: warm caches)."""
import csv, json, re, sys

COLS = [("gpu__time_duration.sum", "ms", 1e-6), ("dram__bytes_read.sum", "dram rd MB", None), ("dram__bytes_write.sum", "dram wr MB", None),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram %", 1), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue %", 1),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps act %", 1), ("launch__registers_per_thread", "regs", 1),
        ("smsp__thread_inst_executed_per_inst_executed.ratio", "thr/inst", 1), ("l1tex__t_sector_hit_rate.pct", "L1 hit %", 1),
        ("lts__t_sector_hit_rate.pct", "L2 hit %", 1), ("smsp__inst_executed.sum", "warp inst", 1),
        ("l1tex__throughput.avg.pct_of_peak_sustained_active", "L1 tput %", 1), ("launch__grid_size", "grid", 1)]

def to_bytes(v, unit):
    m = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return v * m.get(unit, 1)

def main():
    path = sys.argv[1]
    rows = list(csv.reader(l for l in open(path) if not l.startswith("==")))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ix = {h: i for i, h in enumerate(hdr)}
    def num(s):
        try: return float(s.replace(",", ""))
        except ValueError: return float("nan")
    print("| kernel | " + " | ".join(c[1] for c in COLS) + " |")
    print("|---|" + "---|" * len(COLS))
    out = {}
    for d in data:
        name = re.sub(r"\(.*", "", d[ix["Kernel Name"]]).replace("void ", "")
        cells, rec = [], {}
        for key, label, scale in COLS:
            if key not in ix:
                cells.append("-"); continue
            v = num(d[ix[key]]); u = units[ix[key]]
            if label == "ms":
                v = v * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(u, 1e-6)
            elif label.startswith("dram") and label.endswith("MB"):
                v = to_bytes(v, u) / 1e6
            rec[label] = v
            cells.append("%.4g" % v)
        print("| %s | %s |" % (name, " | ".join(cells)))
        out[name.split("<")[0]] = rec
    if "--json" in sys.argv:
        # profiles/roofline_traffic.json: per kernel name (the LONGEST captured launch: the full frame, not a band) what
        # bench.py's roofline needs: warp instructions, L1 wavefronts and DRAM bytes per launch
        src = sys.argv[sys.argv.index("--source") + 1] if "--source" in sys.argv else path
        def raw(d, key):
            return num(d[ix[key]]) if key in ix else None
        best = {}
        for d in data:
            name = re.sub(r"\(.*", "", d[ix["Kernel Name"]]).replace("void ", "").split("<")[0]
            ms = num(d[ix["gpu__time_duration.sum"]]) * {"ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3}.get(units[ix["gpu__time_duration.sum"]], 1e-6)
            if name in best and best[name][0] >= ms:
                continue
            wav = raw(d, "l1tex__data_pipe_lsu_wavefronts.sum")
            if wav is None:  # the raw page carries it as a per-SM average under a section prefix
                key = next((h for h in hdr if h.endswith("l1tex__data_pipe_lsu_wavefronts.avg")), None)
                n_l1 = (raw(d, "l1tex__cycles_elapsed.sum") or 0) / (raw(d, "l1tex__cycles_elapsed.avg") or 1)
                wav = raw(d, key) * round(n_l1) if key and n_l1 else None
            rd = to_bytes(raw(d, "dram__bytes_read.sum"), units[ix["dram__bytes_read.sum"]])
            wr = to_bytes(raw(d, "dram__bytes_write.sum"), units[ix["dram__bytes_write.sum"]])
            best[name] = (ms, {"dram_bytes_per_launch": rd + wr, "ms_under_ncu": ms, "warp_inst": raw(d, "smsp__inst_executed.sum"),
                               "threads_per_inst": raw(d, "smsp__thread_inst_executed_per_inst_executed.ratio"),
                               "l1_wavefronts": wav, "issue_active_pct": raw(d, "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                               "l1_throughput_pct": raw(d, "l1tex__throughput.avg.pct_of_peak_sustained_active"),
                               "l2_hit_pct": raw(d, "lts__t_sector_hit_rate.pct"), "registers": raw(d, "launch__registers_per_thread"),
                               "grid": raw(d, "launch__grid_size"), "source": src})
        json.dump({k: v[1] for k, v in best.items()}, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)

main()
