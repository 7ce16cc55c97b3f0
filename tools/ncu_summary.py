"""Summarise an .ncu-rep (one kernel launch) into a small text file for profiles/:
key raw metrics, SASS opcode mix and the hottest straight-line blocks (from the source page).
With --json OUT.json --meta META.json it also writes the machine-readable counts bench.py's roofline block is
computed from: executed MUFU and FMA-pipe work per kernel evaluation, DRAM bytes, keyed by the hash of the kernel's
sources (META is what the profiled target wrote: kernel family, evaluations per launch, source hash).
Usage: python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/out.txt [--json profiles/out.json --meta gpurun_out/meta.json]"""
import collections
import csv
import io
import json
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
json_out = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None
meta_in = sys.argv[sys.argv.index("--meta") + 1] if "--meta" in sys.argv else None
# SASS opcodes that execute on the FMA pipe (B300_MICROARCH.md: FFMA/FMUL/IMAD/HFMA2 on the fma pipe); the FP32x2
# forms occupy it for two passes
FMA_SCALAR = ("FFMA", "FMUL", "FADD", "IMAD", "HFMA2", "HADD2", "HMUL2")
FMA_PACKED = ("FFMA2", "FMUL2", "FADD2")
KEYS = ["gpu__time_duration.sum", "sm__cycles_elapsed.max", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__inst_issued.avg.per_cycle_active", "smsp__inst_executed.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__average_warps_issue_stalled", "sm__sass_thread_inst_executed_op_f", "smsp__sass_thread_inst_executed_op_f"]


def run(page):
    return subprocess.run(["ncu", "-i", rep, "--page", page, "--csv"], capture_output=True, text=True).stdout


with open(out, "w") as f:
    rows = list(csv.reader(io.StringIO(run("raw"))))
    hdr, units, vals = rows[0], rows[1], rows[-1]
    f.write("# ncu --set full --clock-control none: %s\n# kernel: %s\n\n## raw metrics\n" % (rep, vals[hdr.index("Kernel Name")]))
    for h, u, v in zip(hdr, units, vals):
        if any(h.startswith(k) for k in KEYS) and not any(s in h for s in (".max.", ".min.", ".sum.pct", ".sum.per")):
            f.write("%-90s %-14s %s\n" % (h, u, v))
    rows = list(csv.reader(io.StringIO(run("source"))))
    hdr, data = rows[1], rows[2:]
    isrc, iex, ism = hdr.index("Source"), hdr.index("Instructions Executed"), hdr.index("# Samples")
    tot = sum(int(r[iex]) for r in data)
    by = collections.Counter()
    for r in data:
        t = r[isrc].split()
        op = t[1] if t[0].startswith("@") else t[0]
        by[op.split(".")[0]] += int(r[iex])
    f.write("\n## SASS opcode mix (warp instructions executed, total %d)\n" % tot)
    for k, v in by.most_common(16):
        f.write("%-10s %6.2f%%\n" % (k, 100.0 * v / tot))
    f.write("\n## hottest straight-line blocks (same execution count)\n")
    blocks, cur = [], None
    for i, r in enumerate(data):
        c = int(r[iex])
        if cur and abs(c - cur[2]) <= 0.02 * max(c, cur[2], 1):
            cur[1] = i; cur[3] += c; cur[4] += int(r[ism])
        else:
            if cur:
                blocks.append(cur)
            cur = [i, i, c, c, int(r[ism])]
    blocks.append(cur)
    for b in sorted(blocks, key=lambda b: -b[3])[:8]:
        mufu = sum(1 for r in data[b[0]:b[1] + 1] if "MUFU" in r[isrc])
        f.write("instr %5d-%5d  n=%4d (MUFU %2d)  exec/instr=%12d  share=%5.2f%%  stall samples=%d\n" % (
            b[0], b[1], b[1] - b[0] + 1, mufu, b[2], 100.0 * b[3] / tot, b[4]))
print("wrote", out)

if json_out:
    meta = json.load(open(meta_in)) if meta_in else {}

    def raw(name):
        for h, v in zip(hdr0, vals0):
            if h == name:
                try:
                    return float(v.replace(",", ""))
                except ValueError:
                    return None
        return None

    rows0 = list(csv.reader(io.StringIO(run("raw"))))
    hdr0, units0, vals0 = rows0[0], rows0[1], rows0[-1]
    unit = dict(zip(hdr0, units0))

    def in_bytes(name):
        v = raw(name)
        u = unit.get(name, "byte")
        return None if v is None else v * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1.0)

    def in_ms(name):
        v = raw(name)
        u = unit.get(name, "ms")
        return None if v is None else v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1.0)

    mufu = sum(v for k, v in by.items() if k == "MUFU")
    packed = sum(v for k, v in by.items() if k in FMA_PACKED)
    scalar = sum(v for k, v in by.items() if k in FMA_SCALAR)
    evals = float(meta.get("evals_per_launch", 0)) or None
    doc = {
        "kernel": vals0[hdr0.index("Kernel Name")], "report": rep, "source_hash": meta.get("source_hash"),
        "workload": meta, "duration_ms": in_ms("gpu__time_duration.sum"),
        "dram_bytes_read": in_bytes("dram__bytes_read.sum"), "dram_bytes_write": in_bytes("dram__bytes_write.sum"),
        "warp_inst": {"total": tot, "mufu": mufu, "fma_pipe_packed_f32x2": packed, "fma_pipe_scalar": scalar},
        "pipe_pct_of_peak": {"xu": raw("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                             "fma_inst": raw("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                             "issue": raw("smsp__issue_active.avg.pct_of_peak_sustained_active")},
        "registers_per_thread": raw("launch__registers_per_thread"),
        "note": "per_eval: thread-level operations per kernel evaluation = warp instructions x 32 / evaluations of the launch; "
                "an FP32x2 instruction occupies the FMA pipe for two passes, so fma_pipe_lane_slots = (scalar + 2 x packed) x 32 / evals",
    }
    if evals:
        doc["per_eval"] = {"mufu_lane_ops": mufu * 32.0 / evals, "fma_pipe_lane_slots": (scalar + 2.0 * packed) * 32.0 / evals,
                           "warp_inst_x32": tot * 32.0 / evals}          # = warp instructions per warp-wide (32-lane) evaluation
    json.dump(doc, open(json_out, "w"), indent=1)
    print("wrote", json_out)
