"""Compact text summary of an .ncu-rep (run where ncu is installed; no GPU needed).
usage: python tools/ncu_summary.py report.ncu-rep [max_kernels]
       python tools/ncu_summary.py --traffic report.ncu-rep     (updates profiles/dram_traffic.json:
           per kernel name, dram__bytes_read.sum + dram__bytes_write.sum averaged over its launches
           in the report -- what bench.py reports as roofline.traffic)"""
import csv
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_dmma_cycles_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
]


def to_bytes(v, unit):
    x = float(v.replace(",", ""))
    return x * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}.get(unit, 1.0)


def traffic(rep):
    import json
    import os
    import re
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    u = dict(zip(hdr, units))
    acc = {}
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = re.split(r"[<(]", d.get("Kernel Name", "?"))[0].split("::")[-1].strip().split()[-1]
        b = to_bytes(d["dram__bytes_read.sum"], u["dram__bytes_read.sum"]) + \
            to_bytes(d["dram__bytes_write.sum"], u["dram__bytes_write.sum"])
        acc.setdefault(name, []).append(b)
    path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "dram_traffic.json")
    try:
        cur = json.load(open(path))
    except Exception:
        cur = {}
    for k, v in acc.items():
        cur[k] = {"dram_bytes_per_launch": sum(v) / len(v), "launches_in_capture": len(v),
                  "source": "ncu --set full, " + os.path.basename(rep)}
    json.dump(cur, open(path, "w"), indent=1, sort_keys=True)
    print(json.dumps({k: cur[k] for k in acc}, indent=1))


def main():
    if sys.argv[1] == "--traffic":
        return traffic(sys.argv[2])
    rep = sys.argv[1]
    maxk = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:2 + maxk]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        print("== %s  (id %s)" % (d.get("Kernel Name", "?")[:100], d.get("ID")))
        for k in WANT:
            if k in d:
                print("  %-75s %s %s" % (k, d[k], u[k]))
        stalls = [(k, d[k]) for k in hdr if "issue_stalled" in k and k.endswith("_per_warp_active.pct") and d.get(k)]
        stalls = sorted(stalls, key=lambda kv: -float(kv[1].replace(",", "") or 0))[:8]
        for k, v in stalls:
            print("  stall %-69s %s" % (k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_warp_active.pct", ""), v))


if __name__ == "__main__":
    main()
