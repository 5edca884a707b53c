"""ncu_summary.py -- turns one `ncu --set full --import-source on` report of a kernel into the text summary
committed under profiles/ (selected raw metrics + per-opcode instruction and stall-sample table).

    python tools/ncu_summary.py gpurun_out/r01c_project.ncu-rep "header line ..." > profiles/r01c_project_kernel_ncu_full.txt
    python tools/ncu_summary.py rep.ncu-rep --json profiles/project_kernel_profile.json --key <cubin key> \
        --workload 10000x50000x2048 --source profiles/<summary>.txt "header ..." > profiles/<summary>.txt

--json writes the machine-readable profile bench.py reads (instruction count, pipe utilisations, DRAM bytes of one
launch), keyed by the sha256 of the projection cubin that was profiled (tools/cubin_key.py, recorded in the same
gpurun call as the capture).  Runs in the build container (ncu reads reports without a GPU).  Tooling, not product code."""
import json
import collections
import csv
import io
import re
import subprocess
import sys

KEEP = re.compile(
    r"^(gpu__time_duration\.sum|dram__bytes_(read|write)\.sum(\.per_second)?|dram__bytes\.sum\.per_second|"
    r"gpu__dram_throughput\.avg\.pct_of_peak_sustained_elapsed|"
    r"sm__inst_executed_pipe_(alu|fma|fmaheavy|fmalite|xu|lsu|uniform|fp64|tc)\.avg\.pct_of_peak_sustained_active|"
    r"sm__pipe_(alu|fma|fmaheavy|fmalite|xu)_cycles_active\.avg\.pct_of_peak_sustained_(active|elapsed)|"
    r"smsp__issue_active\.avg\.pct_of_peak_sustained_active|smsp__inst_executed\.sum|"
    r"smsp__warps_eligible\.avg\.per_cycle_active|sm__warps_active\.avg\.pct_of_peak_sustained_active|"
    r"smsp__average_warps_issue_stalled_[a-z_]+_per_issue_active\.ratio|"
    r"sm__cycles_(elapsed|active)\.avg|sm__cycles_elapsed\.avg\.per_second|"
    r"l1tex__data_bank_conflicts_pipe_lsu_mem_shared(_op_(ld|st|ldgsts))?\.sum|"
    r"l1tex__data_pipe_lsu_wavefronts_mem_shared(_op_(ld|st))?\.sum|"
    r"l1tex__t_bytes_pipe_lsu_mem_global_op_ldgsts_cache_access\.sum|"
    r"launch__(grid_size|block_size|registers_per_thread|shared_mem_per_block_dynamic|"
    r"occupancy_limit_(registers|shared_mem|warps|blocks)|waves_per_multiprocessor)|"
    r"sm__maximum_warps_per_active_cycle_pct|lts__t_sector_hit_rate\.pct)$")


def ncu_csv(rep, page):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def to_bytes(value, unit):
    mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    return float(value.replace(",", "")) * mult[unit]


def write_json(path, key, workload, source, hdr, units, vals):
    ix = {h: i for i, h in enumerate(hdr)}

    def num(name):
        return float(vals[ix[name]].replace(",", ""))

    dur_unit = units[ix["gpu__time_duration.sum"]]
    dur_ms = num("gpu__time_duration.sum") * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(dur_unit, 1e-6)
    prof = {
        "kernel_sass_sha256": key, "kernel": vals[ix["Kernel Name"]], "grid": vals[ix["Grid Size"]], "workload": workload,
        "source": source,
        "inst_executed": num("smsp__inst_executed.sum"),
        "sm_cycles_elapsed_avg": num("sm__cycles_elapsed.avg"),
        "sm_clock_ghz": num("sm__cycles_elapsed.avg.per_second"),
        "duration_ms_under_ncu": dur_ms,
        "issue_active_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
        "pipe_alu_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
        "pipe_fmaheavy_pct": num("sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed"),
        "pipe_xu_pct": num("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
        "pipe_lsu_pct": num("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
        "dram_bytes_read": to_bytes(vals[ix["dram__bytes_read.sum"]], units[ix["dram__bytes_read.sum"]]),
        "dram_bytes_write": to_bytes(vals[ix["dram__bytes_write.sum"]], units[ix["dram__bytes_write.sum"]]),
        "shared_bank_conflicts": num("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"),
        "registers_per_thread": num("launch__registers_per_thread"),
    }
    json.dump(prof, open(path, "w"), indent=1)
    open(path, "a").write("\n")


def main():
    argv = sys.argv[1:]
    opts = {}
    for flag in ("--json", "--key", "--workload", "--source"):
        if flag in argv:
            i = argv.index(flag)
            opts[flag] = argv[i + 1]
            del argv[i:i + 2]
    rep = argv[0]
    for line in argv[1:]:
        print("# " + line)
    raw = ncu_csv(rep, "raw")
    hdr, units, vals = raw[0], raw[1], raw[2]
    ix = {h: i for i, h in enumerate(hdr)}
    if "--json" in opts:
        write_json(opts["--json"], opts.get("--key", ""), opts.get("--workload", ""), opts.get("--source", ""), hdr, units, vals)
        print("# cubin key of the profiled library (tools/cubin_key.py): %s" % opts.get("--key", ""))
    print("# kernel: %s  grid %s block %s" % (vals[ix["Kernel Name"]], vals[ix["Grid Size"]], vals[ix["Block Size"]]))
    print("#\n# ---- selected raw metrics (ncu --page raw) ----")
    for h, u, v in zip(hdr, units, vals):
        if KEEP.match(h):
            print("%-100s %-16s %s" % (h, u, v))

    src = ncu_csv(rep, "source")
    shdr = src[1]
    six = {h: i for i, h in enumerate(shdr)}
    cols = ["# Samples", "Instructions Executed", "stall_dispatch", "stall_math", "stall_wait", "stall_not_selected",
            "stall_selected", "stall_short_sb", "stall_long_sb", "stall_mio"]
    agg = collections.defaultdict(collections.Counter)
    for r in src[2:]:
        if len(r) < len(shdr):
            continue
        m = re.match(r"\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)", r[six["Source"]])
        op = m.group(2) if m else "?"
        if op.startswith("IMAD"):
            op = "IMAD.WIDE" if "WIDE" in op else ("IMAD.IADD" if "IADD" in op else ("IMAD.X" if ".X" in op else "IMAD"))
        else:
            op = op.split(".")[0]
        for k in cols:
            try:
                agg[op][k] += int(float(r[six[k]] or 0))
            except (ValueError, KeyError):
                pass
    total_inst = sum(c["Instructions Executed"] for c in agg.values())
    print("#\n# ---- per-opcode warp instructions and stall samples (ncu --page source, SASS view) ----")
    print("%-10s %10s %7s %8s %8s %8s %8s %8s %8s %8s %8s %8s" % (
        "opcode", "inst(M)", "share", "samples", "dispatch", "math", "wait", "not_sel", "selected", "short_sb",
        "long_sb", "mio"))
    for op, c in sorted(agg.items(), key=lambda kv: -kv[1]["Instructions Executed"])[:18]:
        print("%-10s %10.1f %6.1f%% %8d %8d %8d %8d %8d %8d %8d %8d %8d" % (
            op, c["Instructions Executed"] / 1e6, 100.0 * c["Instructions Executed"] / max(1, total_inst), c["# Samples"],
            c["stall_dispatch"], c["stall_math"], c["stall_wait"], c["stall_not_selected"], c["stall_selected"],
            c["stall_short_sb"], c["stall_long_sb"], c["stall_mio"]))
    print("%-10s %10.1f" % ("total", total_inst / 1e6))


if __name__ == "__main__":
    main()
