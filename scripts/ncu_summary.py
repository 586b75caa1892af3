"""Summarise an .ncu-rep (read here, no GPU): headline metrics per kernel, pipe
utilisation, and the source lines with the most stall samples.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [--lines N]"""
import collections
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
nlines = int(sys.argv[sys.argv.index("--lines") + 1]) if "--lines" in sys.argv else 12
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "sm__mio_pq_read_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct"]
for r in rows[2:]:
    print("=====", r[hdr.index("Kernel Name")][:90])
    for w in WANT:
        if w in hdr:
            print(f"  {w} = {r[hdr.index(w)]} {units[hdr.index(w)]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
seen = set()
for blk in src.split('"Kernel Name",')[1:]:
    lines = blk.split("\n")
    name = lines[0][:90]
    if name in seen:
        continue
    seen.add(name)
    rr = list(csv.reader(lines[1:]))
    h = rr[0]
    iS, iN, iX = h.index("Source"), h.index("# Samples"), h.index("Instructions Executed")
    stall_cols = [(i, c) for i, c in enumerate(h) if c.startswith("stall_") and "Not Issued" not in c]
    tot = collections.Counter()
    cnt = collections.Counter()
    stalls = collections.Counter()
    samples = 0
    per = []
    for r in rr[1:]:
        if len(r) <= iN or not r[iS].split():
            continue
        try:
            ns, nx = int(r[iN]), int(r[iX])
        except ValueError:
            continue
        op = r[iS].split()
        o = (op[1] if op[0].startswith("@") and len(op) > 1 else op[0]).split(".")[0]
        tot[o] += ns
        cnt[o] += nx
        samples += ns
        for i, c in stall_cols:
            try:
                stalls[c] += int(r[i])
            except ValueError:
                pass
        per.append((ns, nx, r[iS].strip()))
    print("=====", name, "samples", samples, "warp-instr", sum(cnt.values()))
    print("  stalls:", ", ".join(f"{c[6:]} {100 * v / max(1, samples):.1f}%" for c, v in stalls.most_common(7)))
    for o, v in tot.most_common(nlines):
        print(f"  {o:10s} samples {100 * v / max(1, samples):5.1f}%  warp-instr {cnt[o]:>12d}")
    print("  -- hottest instructions")
    for ns, nx, t in sorted(per, reverse=True)[:nlines]:
        print(f"  {100 * ns / max(1, samples):5.1f}%  x{nx:<10d} {t[:100]}")
