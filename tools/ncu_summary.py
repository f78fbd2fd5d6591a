"""Summarise ncu reports (read on the CPU box) into profiles/: per-kernel launch shares and the key metrics of the
`--set full` captures.  Usage: python tools/ncu_summary.py <round-tag> <launches.csv> <rep> [<rep> ...]"""
import collections, csv, json, re, subprocess, sys

PIPE = re.compile(r"pipe_tensor|tensor_op|dmma|pipe_fp64|pipe_fmaheavy.*pct|inst_executed_pipe_uniform")

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct"]


def launches(path):
    rows = list(csv.reader(open(path, errors="replace")))
    hdr, agg = None, collections.OrderedDict()
    for r in rows:
        if "Kernel Name" in r:
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            try:
                v = float(d["Metric Value"].replace(",", ""))
            except ValueError:
                continue
            unit = d.get("Metric Unit", "ns")
            v_us = v / 1e3 if unit in ("ns", "nsecond") else (v if unit in ("us", "usecond") else v * 1e3)
            agg.setdefault(d["Kernel Name"].split("(")[0], []).append(v_us)
    tot = sum(sum(v) for v in agg.values())
    return {k: {"launches": len(v), "avg_us": sum(v) / len(v), "total_us": sum(v), "share": sum(v) / tot} for k, v in agg.items()}


def full(path):
    # a .ncu-rep report, or its raw page already exported as CSV on the GPU box (tools/profile_r2.sh)
    out = open(path, errors="replace").read() if path.endswith(".csv") else \
        subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    res = []
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        e = {"kernel": d.get("Kernel Name", "?")}
        for k in KEYS:
            if k in d:
                e[k] = f"{d[k]} {units[hdr.index(k)]}".strip()
        # the FP64 tensor (DMMA) sub-pipe: every tensor-pipe counter the report carries, whatever this ncu calls them
        for k in hdr:
            if PIPE.search(k) and k not in e and d.get(k, "") not in ("", "n/a"):
                e[k] = f"{d[k]} {units[hdr.index(k)]}".strip()
        res.append(e)
    return res


if __name__ == "__main__":
    tag, lpath, reps = sys.argv[1], sys.argv[2], sys.argv[3:]
    summary = {"launch_list": {"file": lpath, "note": "cold-cache, serialised under ncu: compare SHARES, not absolutes",
                               "kernels": launches(lpath)}, "full_set": {}}
    for r in reps:
        summary["full_set"][r] = full(r)
    json.dump(summary, open(f"profiles/{tag}_ncu_summary.json", "w"), indent=1)
    for k, v in summary["launch_list"]["kernels"].items():
        print(f"{k[:60]:60s} n={v['launches']:4d} avg={v['avg_us']:10.1f} us share={v['share']*100:5.1f}%")
