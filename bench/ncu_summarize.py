"""Turn `ncu -i <rep> --page raw --csv` into the markdown table kept under profiles/.

    ncu -i gpurun_out/prof_all.ncu-rep --page raw --csv > /tmp/raw.csv
    python bench/ncu_summarize.py /tmp/raw.csv > profiles/ncu_all_kernels.md

Columns: duration; DRAM bytes read + written (and % of peak DRAM throughput); tensor-pipe % (the guide's
`sm__pipe_tensor_cycles_active`); achieved warps; registers; the largest warp-stall reason per issue.
"""
import csv
import sys

STALLS = {
    "long_scoreboard": "long scoreboard (global/local memory)",
    "short_scoreboard": "short scoreboard (shared memory / MUFU)",
    "barrier": "barrier",
    "membar": "memory barrier",
    "wait": "fixed-latency wait",
    "math_pipe_throttle": "math pipe throttle",
    "mio_throttle": "MIO throttle",
    "lg_throttle": "LG throttle (global memory queue)",
    "tex_throttle": "TEX throttle",
    "sleeping": "sleeping (nanosleep / mbarrier try_wait)",
    "branch_resolving": "branch resolving",
    "dispatch_stall": "dispatch",
    "drain": "drain",
    "imc_miss": "constant-cache miss",
    "no_instruction": "no instruction",
    "not_selected": "not selected",
    "selected": "selected",
    "gmma": "tensor-core wait",
}


def fnum(d, key, default=0.0):
    try:
        return float(d.get(key, "").replace(",", ""))
    except ValueError:
        return default


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units = rows[0], rows[1]
    unit_of = dict(zip(hdr, units))
    out = []
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        name = d.get("Kernel Name", "?").replace("rdp::", "")
        name = name.split("(")[0] if len(name) > 70 else name
        dur = fnum(d, "gpu__time_duration.sum")
        if unit_of.get("gpu__time_duration.sum", "").startswith("ns"):
            dur /= 1e3
        elif unit_of.get("gpu__time_duration.sum", "").startswith("ms"):
            dur *= 1e3

        def to_mb(key):
            v, u = fnum(d, key), unit_of.get(key, "")
            return v * {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
        rd, wr = to_mb("dram__bytes_read.sum"), to_mb("dram__bytes_write.sum")
        dram_pct = fnum(d, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
                        fnum(d, "dram__throughput.avg.pct_of_peak_sustained_elapsed"))
        tensor = 0.0
        for key in ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
                    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
                    "sm__ops_path_tensor_op_hmma_src_bf16_dst_fp32_sparsity_off.avg.pct_of_peak_sustained_elapsed"):
            if key in d and d[key] not in ("", "n/a"):
                tensor = fnum(d, key)
                break
        warps = fnum(d, "sm__warps_active.avg.pct_of_peak_sustained_active")
        regs = int(fnum(d, "launch__registers_per_thread"))
        grid = d.get("launch__grid_size", "?")
        stall, stall_v = "-", 0.0
        for k in hdr:
            if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio"):
                key = k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]
                if key in ("selected", "not_selected"):
                    continue
                v = fnum(d, k)
                if v > stall_v:
                    stall, stall_v = STALLS.get(key, key), v
        out.append((name, dur, rd, wr, dram_pct, tensor, warps, regs, grid, f"{stall} ({stall_v:.1f})"))
    print("| kernel | grid | regs | time (us) | DRAM read (MB) | DRAM written (MB) | DRAM % of peak | tensor pipe % | "
          "warps active % | top stall (warps per issue) |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for o in out:
        print(f"| `{o[0]}` | {o[8]} | {o[7]} | {o[1]:.1f} | {o[2]:.1f} | {o[3]:.1f} | {o[4]:.0f} | {o[5]:.0f} | "
              f"{o[6]:.0f} | {o[9]} |")


if __name__ == "__main__":
    main()
