"""Turn an `ncu --set full` capture into the small tracked files under profiles/:

    python tools/ncu_extract.py gpurun_out/<name>.ncu-rep <kernel key> <out prefix> [launch note]

writes  <out prefix>_ncu_raw.csv      the raw page (every metric of the captured launch)
        <out prefix>_ncu_metrics.json the metrics quoted in profiles/*_summary.md
        <out prefix>_stalls.json      warp-stall samples per reason, summed over the SASS lines
and prints the DRAM bytes per launch (what bench.py reports as roofline.traffic).
Runs here (no GPU needed): `ncu -i` only reads the report."""
import csv
import io
import json
import subprocess
import sys

KEEP = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
        "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_static",
        "sm__warps_active.avg.per_cycle_active", "sm__cycles_active.avg", "sm__cycles_elapsed.avg",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.per_cycle_active", "smsp__inst_executed.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_barrier.ratio",
        "smsp__average_warp_latency_issue_stalled_wait.ratio",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard.ratio",
        "smsp__average_warp_latency_issue_stalled_short_scoreboard.ratio",
        "smsp__average_warp_latency_issue_stalled_math_pipe_throttle.ratio",
        "smsp__average_warp_latency_issue_stalled_not_selected.ratio",
        "smsp__average_warp_latency_issue_stalled_no_instruction.ratio",
        "smsp__average_warp_latency_issue_stalled_lg_throttle.ratio",
        "smsp__average_warp_latency_issue_stalled_mio_throttle.ratio")
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def page(rep, name):
    return subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True,
                          check=True).stdout


def main():
    rep, key, prefix = sys.argv[1:4]
    note = sys.argv[4] if len(sys.argv) > 4 else ""
    raw = page(rep, "raw")
    open(prefix + "_ncu_raw.csv", "w").write(raw)
    rows = list(csv.reader(io.StringIO(raw)))
    names, units, vals = rows[0], rows[1], rows[2]
    met = {n: {"unit": u, "value": v} for n, u, v in zip(names, units, vals) if n in KEEP}
    json.dump({key: met, "note": note}, open(prefix + "_ncu_metrics.json", "w"), indent=1, sort_keys=True)
    dram = sum(float(met[m]["value"]) * SCALE[met[m]["unit"]] for m in ("dram__bytes_read.sum", "dram__bytes_write.sum"))
    src = list(csv.reader(io.StringIO(page(rep, "source"))))
    hdr = next(r for r in src if r and r[0] == "Address")
    body = [r for r in src[src.index(hdr) + 1:] if len(r) == len(hdr)]
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    stalls = {hdr[i]: sum(int(r[i] or 0) for r in body) for i in stall_cols}
    total = sum(stalls.values()) or 1
    json.dump({key: {"samples": stalls, "share": {k: v / total for k, v in stalls.items()}}, "note": note},
              open(prefix + "_stalls.json", "w"), indent=1, sort_keys=True)
    print(json.dumps({key: {"dram_bytes_per_launch": dram,
                            "duration_ms_under_ncu": float(met["gpu__time_duration.sum"]["value"])}}))


if __name__ == "__main__":
    main()
