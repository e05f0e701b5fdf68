"""Summaries of an `ncu --set full --import-source on` capture of window_lm_kernel, written under profiles/.

usage: python tools/ncu_summary.py gpurun_out/r1_win592_final.ncu-rep r1 [windows_in_capture]

Writes  profiles/<round>_window_kernel_ncu_full.json   selected raw metrics of the launch
        profiles/<round>_window_kernel_phases.txt      samples / warp instructions per phase of the kernel
        profiles/traffic.json                          DRAM bytes per launch scaled to the default bench launch
Needs the ncu CLI (reads the report, runs nothing on a GPU).
"""
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = re.compile(r"^(dram__bytes_(read|write)\.sum(\.per_second)?|gpu__time_duration\.sum|gpu__dram_throughput.*elapsed|"
                  r"l1tex__t_sector_hit_rate\.pct|lts__t_sector_hit_rate\.pct|launch__(registers_per_thread|occupancy_limit_.*|"
                  r"shared_mem_per_block_dynamic|grid_size|block_size)|sm__cycles_elapsed\.max|smsp__inst_executed\.sum|"
                  r"sm__inst_executed_pipe_tensor_subpipe_dmma.*active|sm__pipe_fp64_cycles_active.*active|"
                  r"sm__warps_active.*active|smsp__issue_active\.avg\.pct_of_peak_sustained_active|"
                  r"l1tex__data_pipe_lsu_wavefronts_mem_shared\.sum\.pct_of_peak_sustained_elapsed|"
                  r"smsp__average_warps_issue_stalled_.*_per_issue_active\.ratio)$")


def ncu(*args):
    return subprocess.run(["ncu", *args], capture_output=True, text=True, check=True).stdout


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return x


def raw_metrics(rep):
    rows = list(csv.reader(ncu("-i", rep, "--page", "raw", "--csv").splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    out = {"Kernel Name": vals[hdr.index("Kernel Name")]}
    for h, u, v in zip(hdr, units, vals):
        if KEEP.match(h):
            out[h] = {"value": num(v), "unit": u}
    return out


def I(x):
    try:
        return int(x)
    except ValueError:
        return 0


def phase_table(rep, windows, iters=10, obs=15000):
    """samples and executed warp instructions per source-line range of ba_window.cu (phase markers are the
    `// ----- <name>` banner comments of the kernel)."""
    src = open(os.path.join(ROOT, "visualodometry_b200/csrc/ba_window.cu")).read().split("\n")
    marks = [(i + 1, re.sub(r"[-/ ]+$", "", re.sub(r"^\s*// -*\s*", "", l)).strip()[:44])
             for i, l in enumerate(src)
             if re.match(r"^\s*// -{20,} \S", l) or re.match(r"^\s*// (loop 1:|loop 2:|Z Z\^T accumulate|the next batch's landmark)", l)]
    # address -> (file, line, samples, warp instructions, thread instructions) from the cuda,sass view ...
    rows = list(csv.reader(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass").splitlines()))
    cur = hdr = line = None
    addr = {}
    for r in rows:
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
            continue
        if len(r) > 5 and r[0] == "Line No":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            if r[0] != "":
                line = int(r[0])
                continue
            addr[r[2]] = (cur, line, I(r[6]), I(r[7]), I(r[8]))
    # ... then walk the SASS in address order: inlined helpers (ba_math.cuh) and the lambdas belong to the phase
    # of the last kernel-body line (>= first banner) seen before them
    sass = list(csv.reader(ncu("-i", rep, "--page", "source", "--csv", "--print-source", "sass").splitlines()))[2:]
    per = {}
    name = "prologue"
    first = marks[0][0] if marks else 0
    for r in sass:
        if not r or r[0] not in addr:
            continue
        f, l, smp, wi, ti_ = addr[r[0]]
        if f == "ba_window.cu" and l >= first:
            for ln, nm in marks:
                if l >= ln:
                    name = nm
        d = per.setdefault(name, [0, 0, 0])
        d[0] += smp; d[1] += wi; d[2] += ti_
    ts = sum(v[0] for v in per.values()); ti = sum(v[1] for v in per.values())
    unit = windows * iters * obs
    lines = [f"# {os.path.basename(rep)}: {windows} windows x {iters} LM iterations x {obs} observations; phases = banner comments of ba_window.cu"]
    for k, v in per.items():
        lines.append(f"{k[:44]:44s} samples {100 * v[0] / max(ts, 1):5.1f}%  warp-inst {100 * v[1] / max(ti, 1):5.1f}%  = "
                     f"{v[1] / unit:6.2f} winst/obs  thread-inst/obs {v[2] / unit:7.1f}")
    lines.append(f"total warp instructions per observation*iteration {ti / unit:.2f}")
    return "\n".join(lines) + "\n"


def main():
    rep, rnd = sys.argv[1], sys.argv[2]
    windows = int(sys.argv[3]) if len(sys.argv) > 3 else 592
    m = raw_metrics(rep)
    kernel = m.pop("Kernel Name")
    summ = {"kernel": kernel, "round": rnd,
            "command": "ncu --set full --clock-control none --import-source on -k regex:window_lm -s 1 -c 1 python bench.py "
                       f"--windows {windows} --distinct 64 --steps 2 --warmup 1 --no-cpu",
            "workload": f"{windows} C1-shaped windows x 10 LM iterations (one wave: 4 CTAs per SM x 148 SMs)", "metrics": m}
    out = os.path.join(ROOT, "profiles", f"{rnd}_window_kernel_ncu_full.json")
    json.dump(summ, open(out, "w"), indent=1)
    open(os.path.join(ROOT, "profiles", f"{rnd}_window_kernel_phases.txt"), "w").write(phase_table(rep, windows))

    def gb(k):
        v = m[k]
        return v["value"] * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}[v["unit"]]
    per_launch = (gb("dram__bytes_read.sum") + gb("dram__bytes_write.sum")) * 4096 / windows
    json.dump({"c4": per_launch,
               "note": f"dram__bytes_read.sum+dram__bytes_write.sum of one window_lm_kernel launch ({windows} windows, ncu --set "
                       f"full), scaled x4096/{windows} to the default bench launch",
               "source": os.path.relpath(out, ROOT)}, open(os.path.join(ROOT, "profiles", "traffic.json"), "w"), indent=1)
    print(open(os.path.join(ROOT, "profiles", f"{rnd}_window_kernel_phases.txt")).read())
    for k in ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__warps_active.avg.pct_of_peak_sustained_active",
              "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active"):
        print(k, m.get(k))


if __name__ == "__main__":
    main()
