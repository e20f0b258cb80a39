"""Turns the ncu outputs of tools/prof_round.sh (in gpurun_out/) into the tracked summaries under profiles/.

usage: python tools/summarise_profiles.py <tag>          (tag = r01, r02, ... ; reads gpurun_out/*_<suffix> files)
  gpurun_out/launches_<s>.csv      -> profiles/<tag>_launches.csv (copy) + a per-kernel table on stdout
  gpurun_out/prof_<s>_wf.ncu-rep   -> profiles/<tag>_wf_kernel_ncu_metrics.json (selected raw metrics per launch),
                                      profiles/<tag>_wf_kernel_ncu_details.csv (--page details),
                                      profiles/<tag>_wf_kernel_stalls.json (warp-stall samples of the step loop)
  gpurun_out/prof_<s>_chain.ncu-rep-> the same three files for the chained launch (profiles/<tag>_chain_kernel_*)
  gpurun_out/bench_<s>.json        -> profiles/<tag>_bench.json
"""
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "profiles")
SRC = os.path.join(ROOT, "gpurun_out")

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size",
           "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
           "sm__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum"]


def ncu(args):
    return subprocess.run(["ncu"] + args, check=True, capture_output=True, text=True).stdout


def launch_table(path):
    rows = [r for r in csv.reader(open(path)) if r and r[0].isdigit()]
    hdr = next(r for r in csv.reader(open(path)) if r and r[0] == "ID")
    ik, iv = hdr.index("Kernel Name"), hdr.index("Metric Value")
    agg = {}
    for r in rows:
        name = r[ik].replace("fsb::<unnamed>::", "").replace("(bool)", "").replace("(int)", "")
        unit = r[hdr.index("Metric Unit")]
        val = float(r[iv].replace(",", ""))
        ms = val / 1e6 if unit in ("ns", "nsecond") else (val / 1e3 if unit in ("us", "usecond") else val)
        a = agg.setdefault(name, [0, 0.0])
        a[0] += 1
        a[1] += ms
    tot = sum(v[1] for v in agg.values())
    print("| kernel | launches | avg ms | total ms | share |\n|---|---|---|---|---|")
    for name, (n, ms) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name[:90]}` | {n} | {ms / n:.3f} | {ms:.2f} | {100 * ms / tot:.1f}% |")


def main():
    tag, suffix = sys.argv[1], sys.argv[2] if len(sys.argv) > 2 else "r1d"
    os.makedirs(OUT, exist_ok=True)
    b = os.path.join(SRC, f"bench_{suffix}.json")
    if os.path.exists(b):
        line = open(b).read().strip().splitlines()[-1]
        json.dump(json.loads(line), open(os.path.join(OUT, f"{tag}_bench.json"), "w"), indent=1)
    l = os.path.join(SRC, f"launches_{suffix}.csv")
    if os.path.exists(l):
        shutil.copy(l, os.path.join(OUT, f"{tag}_launches.csv"))
        launch_table(l)
    kname = "wf_kernel"
    rep = os.path.join(SRC, f"prof_{suffix}_wf.ncu-rep")
    if not os.path.exists(rep):
        kname = "chain_kernel"
        rep = os.path.join(SRC, f"prof_{suffix}_chain.ncu-rep")
    if not os.path.exists(rep):
        return
    raw = ncu(["-i", rep, "--page", "raw", "--csv"])
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = []
    for r in data:
        rec = {"Kernel Name": r[hdr.index("Kernel Name")]}
        for m in METRICS:
            if m in hdr:
                rec[m] = r[hdr.index(m)]
                rec.setdefault("units", {})[m] = units[hdr.index(m)]
        out.append(rec)
    json.dump(out, open(os.path.join(OUT, f"{tag}_{kname}_ncu_metrics.json"), "w"), indent=1)
    open(os.path.join(OUT, f"{tag}_{kname}_ncu_details.csv"), "w").write(ncu(["-i", rep, "--page", "details", "--csv"]))
    # warp-stall samples inside the step loop of the Gauss-Seidel launch (instructions executed once per step pair)
    src = ncu(["-i", rep, "--page", "source", "--csv", "--print-source", "sass"])
    blocks = src.split('"Kernel Name",')
    stalls_out = []
    for blk in blocks[1:]:
        lines = blk.splitlines()
        rr = list(csv.reader(lines[1:]))
        h = rr[0]
        rr = [r for r in rr[1:] if len(r) == len(h)]
        ix = {n: i for i, n in enumerate(h)}
        names = [n for n in h if n.startswith("stall_") and "Not Issued" not in n]
        execs = sorted(int(r[ix["Instructions Executed"]]) for r in rr)
        # the step bodies: the most common large execution count
        from collections import Counter
        cnt = Counter(e for e in execs if e > 1000)
        cand = sorted(((n * e, e) for e, n in cnt.items() if n >= 100), reverse=True)   # loops by dynamic instructions
        if not cand:
            continue
        allsamples = sum(int(r[ix["# Samples"]] or 0) for r in rr) or 1
        for _, e0 in cand[:3]:                                        # the step loops of the sweep phases come first
            sel = [r for r in rr if int(r[ix["Instructions Executed"]]) == e0]
            tot = sum(int(r[ix["# Samples"]]) for r in sel) or 1
            st = {n: sum(int(r[ix[n]] or 0) for r in sel) for n in names}
            stalls_out.append({"kernel": lines[0].strip('",'), "loop_instructions": len(sel), "executions_each": e0,
                               "samples": tot, "share_of_all_samples": round(tot / allsamples, 4),
                               "stall_share": {k: round(v / tot, 4) for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v}})
    json.dump(stalls_out, open(os.path.join(OUT, f"{tag}_{kname}_stalls.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
