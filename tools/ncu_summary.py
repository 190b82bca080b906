"""Key metrics of one kernel launch from an `ncu --set full` report -> JSON (for profiles/).

    python tools/ncu_summary.py REPORT.ncu-rep OUT.json --units N [--unit-name frames] [--alg-bytes-per-unit 962]

`--units` is the number of work units (fbank frames, STFT frames ...) the captured launch processed;
DRAM bytes per unit are derived from it.  bench.py reads profiles/fbank_ncu_summary.json for
`roofline.traffic`.
"""
import argparse
import csv
import io
import json
import subprocess

KEYS = [
    "gpu__time_duration.sum",
    "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
    "sm__warps_active.avg.pct_of_peak_sustained_active",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
    "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
]
TO_BYTES = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report")
    ap.add_argument("out")
    ap.add_argument("--units", type=float, required=True)
    ap.add_argument("--unit-name", default="frames")
    ap.add_argument("--alg-bytes-per-unit", type=float, default=None)
    ap.add_argument("--command", default="")
    a = ap.parse_args()
    txt = subprocess.run(["ncu", "-i", a.report, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    ix = {h: i for i, h in enumerate(hdr)}
    out = {"kernel": vals[ix["Kernel Name"]] if "Kernel Name" in ix else "", "command": a.command,
           a.unit_name + "_per_launch": a.units, "metrics": {}}
    for k in KEYS:
        if k in ix:
            out["metrics"][k] = [vals[ix[k]], units[ix[k]]]

    def nbytes(k):
        v, u = out["metrics"].get(k, ["0", "byte"])
        return float(v.replace(",", "")) * TO_BYTES.get(u, 1.0)

    dram = nbytes("dram__bytes_read.sum") + nbytes("dram__bytes_write.sum")
    out["dram_bytes_per_launch"] = dram
    out["dram_bytes_per_" + a.unit_name[:-1]] = dram / a.units
    if a.alg_bytes_per_unit is not None:
        out["algorithmic_bytes_per_" + a.unit_name[:-1]] = a.alg_bytes_per_unit
    inst = out["metrics"].get("smsp__inst_executed.sum")
    if inst:
        out["warp_inst_per_" + a.unit_name[:-1]] = float(inst[0].replace(",", "")) / a.units
    wf = out["metrics"].get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum")
    if wf:
        out["smem_wavefronts_per_" + a.unit_name[:-1]] = float(wf[0].replace(",", "")) / a.units
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
