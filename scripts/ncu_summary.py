#!/usr/bin/env python
"""Turns the ncu artefacts a gpurun call brought back (gpurun_out/) into the text summaries kept
under profiles/.   python scripts/ncu_summary.py <tag>   e.g. r01a"""
import collections
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
PROF = os.path.join(ROOT, "profiles")

KEEP = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
    "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__t_bytes.sum", "l1tex__t_sector_hit_rate.pct", "launch__grid_size", "launch__block_size",
    "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem",
    "launch__waves_per_multiprocessor", "launch__shared_mem_per_block_static", "launch__shared_mem_per_block_dynamic",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__cycles_active.avg",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_ld.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_st.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__issue_active.avg.per_cycle_active", "sm__cycles_elapsed.max",
]


def stall_table(f, d):
    """Warp-state section of --set full: every metric whose name says 'stall' (the names differ between ncu
    versions: smsp__average_warp_latency_issue_stalled_*, smsp__average_warps_issue_stalled_*_per_issue_active,
    smsp__warp_issue_stalled_*_per_warp_active), largest first."""
    rows = []
    for h, (v, u) in d.items():
        if "stall" not in h:
            continue
        try:
            rows.append((float(v.replace(",", "")), h, u))
        except ValueError:
            pass
    f.write(f"# warp stall reasons ({len(rows)} metrics with 'stall' in the name; zero rows omitted)\n")
    for x, h, u in sorted(rows, reverse=True):
        if x != 0:
            f.write(f"{h:100s} {x:12.3f} {u}\n")


def launches(tag):
    src = os.path.join(OUT, "launches.csv")
    if not os.path.exists(src):
        return
    rows = [r for r in csv.reader(open(src)) if len(r) > 10]
    hdr = rows[0]
    ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = collections.OrderedDict()
    for r in rows[1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        a = agg.setdefault(r[ki], [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    with open(os.path.join(PROF, f"{tag}_launches.txt"), "w") as f:
        f.write("# ncu --metrics gpu__time_duration.sum --clock-control none (per-launch, cold-cache, serialised)\n")
        f.write("# command: python bench.py --nt 40 --steps 1 --warmup 3 --no-cpu --no-e2e (first 400 launches)\n")
        f.write(f"# total {tot / 1e3:.1f} us over {sum(a[0] for a in agg.values())} launches\n")
        f.write("# count   total_us   share  avg_us  kernel\n")
        for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"{a[0]:6d} {a[1] / 1e3:10.1f} {100 * a[1] / tot:6.1f}% {a[1] / a[0] / 1e3:7.2f}  {k[:110]}\n")


def full(tag, name):
    rep = os.path.join(OUT, name + ".ncu-rep")
    if not os.path.exists(rep):
        return None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
    with open(os.path.join(PROF, f"{tag}_{name}.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on, one launch; {d.get('Kernel Name', ('', ''))[0]}\n")
        for k in KEEP:
            if k in d:
                f.write(f"{k:70s} {d[k][0]:>16s} {d[k][1]}\n")
        stall_table(f, d)
    # the whole raw page of this launch, for anything the summary leaves out
    open(os.path.join(PROF, f"{tag}_{name}.raw.csv"), "w").write(raw)

    def num(k):
        v, u = d[k]
        x = float(v.replace(",", ""))
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        return x * mult
    return num("dram__bytes_read.sum") + num("dram__bytes_write.sum")


def full_multi(tag, name, note):
    """Several kernels in one report (the elastic time step): one block of KEEP metrics per launch."""
    rep = os.path.join(OUT, name + ".ncu-rep")
    if not os.path.exists(rep):
        return None
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    out = {}
    with open(os.path.join(PROF, f"{tag}_{name}.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on; {note}\n")
        for vals in rows[2:]:
            d = {h: (v, u) for h, u, v in zip(hdr, units, vals)}
            kn = d.get("Kernel Name", ("?", ""))[0]
            f.write(f"\n## {kn}\n")
            for k in KEEP:
                if k in d:
                    f.write(f"{k:70s} {d[k][0]:>16s} {d[k][1]}\n")
            stall_table(f, d)

            def num(k):
                v, u = d[k]
                return float(v.replace(",", "")) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
            out.setdefault(kn.split("(")[0], []).append(num("dram__bytes_read.sum") + num("dram__bytes_write.sum"))
    return {k: sum(v) / len(v) for k, v in out.items()}        # mean per launch of each kernel


if __name__ == "__main__":
    tag = sys.argv[1]
    os.makedirs(PROF, exist_ok=True)
    if len(sys.argv) > 2 and sys.argv[2] == "small":
        print(full_multi(tag, "small_full", "C1 cluster-resident time loop: forward (200 steps) and adjoint kernel "
                         "(bench.py --workload c1 --nt 200)"))
        for n in ("bench_c1.log", "bench_c5.log"):
            pth = os.path.join(OUT, n)
            if os.path.exists(pth):
                open(os.path.join(PROF, f"{tag}_{n[:-4]}_1gpu.log.txt"), "w").write(open(pth).read())
        sys.exit(0)
    if len(sys.argv) > 2 and sys.argv[2] == "elastic":
        src = os.path.join(OUT, "launches_c3.csv")
        if os.path.exists(src):
            os.replace(src, os.path.join(OUT, "launches.csv"))
            launches(tag + "_c3")
        tf = full_multi(tag, "el_fwd_full", "C3 elastic forward step: sample, f1, f2, inject (bench.py --workload c3 --nt 40)")
        ta = full_multi(tag, "el_adj_full", "C3 elastic adjoint step: sample, b2_fused, b1_fused, inject (bench.py --workload c3 --nt 40)")
        print(tf, ta)
        tj = os.path.join(PROF, "traffic.json")
        cur = json.load(open(tj)) if os.path.exists(tj) else {}
        if tf and ta:
            cur["c3"] = {"forward_bytes_per_launch": sum(tf.values()), "adjoint_bytes_per_launch": sum(ta.values()),
                         "pair_bytes_per_launch": sum(tf.values()) + sum(ta.values()),
                         "per_kernel": {**tf, **ta}, "source": f"profiles/{tag}_el_*_full.txt"}
            json.dump(cur, open(tj, "w"), indent=1)
        sys.exit(0)
    launches(tag)
    # forward: one launch of a 3-shot group (MODE 1); adjoint: an advance-only launch (MODE 0) and the paired-imaging
    # launch after it (MODE 3), 6 shots each - scripts/gpu_r02c.sh
    tf = full(tag, "fwd_full")
    ta = full_multi(tag, "adj_full", "C2 adjoint: advance-only launches (MODE 0) and paired-imaging launches (MODE 3) of the two "
                    "concurrent 3-shot groups (4 consecutive launches; ncu serialises them)")
    print(tf, ta)
    for n in ("bench.log", "pytest_gpu.log", "smoke.log"):
        p = os.path.join(OUT, n)
        if os.path.exists(p):
            open(os.path.join(PROF, f"{tag}_{n}.txt"), "w").write(open(p).read())
    if tf and ta:
        tj = os.path.join(PROF, "traffic.json")
        cur = json.load(open(tj)) if os.path.exists(tj) else {}
        cells = 751 * 2301
        m0 = [v for k, v in ta.items() if "<4, 0>" in k or "<(int)4, (int)0>" in k]
        m3 = [v for k, v in ta.items() if "<4, 3>" in k or "<(int)4, (int)3>" in k]
        entry = {"forward_bytes_per_launch": tf, "forward_shots_per_launch": 3,
                 "forward_bytes_per_cell_update": tf / (3 * cells), "adjoint_shots_per_launch": 3,
                 "per_kernel_mean_bytes_per_launch": ta, "source": f"profiles/{tag}_*_full.txt"}
        if m0 and m3:
            entry["adjoint_bytes_per_cell_update"] = (m0[0] + m3[0]) / (2 * 3 * cells)
        elif m3 or m0:
            entry["adjoint_bytes_per_cell_update"] = (m3 or m0)[0] / (3 * cells)
        cur["c2"] = entry
        json.dump(cur, open(tj, "w"), indent=1)
    sys.exit(0)
