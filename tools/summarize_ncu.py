"""Turn ncu outputs into the small tracked summaries under profiles/.
  launch list (csv from --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum) -> per-kernel totals and shares
  full capture (.ncu-rep)                                                                      -> one row per launch, key metrics"""
import collections, csv, subprocess, sys


def launch_summary(raw_csv, out_csv):
    rows = [r for r in csv.reader(open(raw_csv)) if len(r) > 8]
    hdr = rows[0]
    ik, im, iu, iv, iid = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Unit"), hdr.index("Metric Value"), hdr.index("ID")
    per = collections.defaultdict(lambda: {"n": set(), "us": 0.0, "rd": 0.0, "wr": 0.0})
    for r in rows[1:]:
        k = r[ik].split("(")[0]
        v = float(r[iv].replace(",", ""))
        u = r[iu]
        d = per[k]
        d["n"].add(r[iid])
        if r[im] == "gpu__time_duration.sum":
            d["us"] += v / 1e3 if u == "ns" else v * 1e3 if u == "ms" else v
        else:
            b = v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            d["rd" if "read" in r[im] else "wr"] += b
    tot = sum(d["us"] for d in per.values())
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["kernel", "launches", "total_us", "share_of_time", "avg_us", "dram_read_MB_per_launch", "dram_write_MB_per_launch"])
        for k, d in sorted(per.items(), key=lambda x: -x[1]["us"]):
            n = len(d["n"])
            w.writerow([k, n, f"{d['us']:.1f}", f"{d['us'] / tot:.3f}", f"{d['us'] / n:.2f}", f"{d['rd'] / n / 1e6:.3f}", f"{d['wr'] / n / 1e6:.3f}"])


KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor_subpipe_imma.avg.pct_of_peak_sustained_active",
        "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__pcsamp_sample_count"]


def full_summary(rep, out_csv):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    with open(out_csv, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(KEYS + ["top stall reasons (share of warp samples)"])
        w.writerow([units[hdr.index(k)] if k in hdr else "" for k in KEYS] + [""])
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            tot = float(d.get("smsp__pcsamp_sample_count") or 0) or 1.0
            st = sorted(((k.replace("smsp__pcsamp_warps_issue_stalled_", ""), float(v)) for k, v in d.items()
                         if k.startswith("smsp__pcsamp_warps_issue_stalled_") and not k.endswith("not_issued")), key=lambda x: -x[1])
            w.writerow([d.get(k, "") for k in KEYS] + [" ".join(f"{k}:{100 * v / tot:.0f}%" for k, v in st[:6])])


if __name__ == "__main__":
    mode, src, dst = sys.argv[1:4]
    (launch_summary if mode == "launches" else full_summary)(src, dst)
