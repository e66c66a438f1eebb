"""Summarises .ncu-rep files (ncu --set full) into one JSON: python tools/summarize_ncu.py out.json a.ncu-rep b.ncu-rep ..."""
import csv, io, json, subprocess, sys

KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__average_warp_latency_per_inst_issued.ratio", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__inst_executed.sum",
        "smsp__inst_executed.avg.per_cycle_active", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]


def main(out, reps):
    res = {}
    for rep in reps:
        txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(txt)))
        if len(rows) < 3:
            res[rep] = "unreadable"
            continue
        hdr, units = rows[0], rows[1]
        for r in rows[2:]:
            d = dict(zip(hdr, r))
            entry = {"kernel": d.get("Kernel Name", "?")[:120]}
            for k in KEEP:
                if k in d:
                    entry[k] = f"{d[k]} {units[hdr.index(k)]}".strip()
            res.setdefault(rep.split("/")[-1], []).append(entry)
    json.dump(res, open(out, "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2:])
