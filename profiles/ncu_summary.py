#!/usr/bin/env python
"""Summarise an .ncu-rep (read on the CPU box): headline metrics + the hottest SASS instructions.

    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep [--top 40] [--regions a,b,c]

Writes nothing; redirect the output into profiles/ to commit it."""
import csv
import io
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared_op_atom.sum",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared_op_atom.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tensor.sum", "sm__cycles_elapsed.max"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    top = int(sys.argv[sys.argv.index("--top") + 1]) if "--top" in sys.argv else 40
    regions = [int(x) for x in sys.argv[sys.argv.index("--regions") + 1].split(",")] if "--regions" in sys.argv else None
    rows = page(rep, "raw")
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        name = r[hdr.index("Kernel Name")] if "Kernel Name" in hdr else "?"
        print(f"== {name[:100]}")
        for h, u, v in zip(hdr, units, r):
            if h in WANT:
                print(f"  {h:70s} {v} {u}")
    rows = page(rep, "source")
    if len(rows) < 3:
        return
    hdr, data = rows[1], rows[2:]
    ia, isamp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
    iw = hdr.index("L1 Wavefronts Shared") if "L1 Wavefronts Shared" in hdr else None
    st = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    data = [r for r in data if len(r) == len(hdr) and r[ia].strip().isdigit()]     # multi-kernel reports repeat the header
    tot = sum(int(r[ia]) for r in data) or 1
    ts = sum(int(r[isamp]) for r in data) or 1
    print(f"-- source page: {len(data)} SASS instructions, {tot} warp-instructions executed, {ts} samples")
    agg = {hdr[i]: sum(int(r[i]) for r in data) for i in st}
    print("   stall samples:", ", ".join(f"{k[6:]}={100 * v / ts:.1f}%" for k, v in sorted(agg.items(), key=lambda x: -x[1])[:8]))
    if regions:
        b = [0] + regions + [len(data)]
        for lo, hi in zip(b[:-1], b[1:]):
            e = sum(int(r[ia]) for r in data[lo:hi]); s = sum(int(r[isamp]) for r in data[lo:hi])
            w = sum(int(r[iw]) for r in data[lo:hi]) if iw is not None else 0
            print(f"   region [{lo},{hi}): inst {100 * e / tot:5.1f}%  samples {100 * s / ts:5.1f}%  smem wavefronts {w / 1e6:.0f}M")
    order = sorted(range(len(data)), key=lambda n: -int(data[n][isamp]))[:top]
    for n in sorted(order):
        r = data[n]
        stalls = sorted(((int(r[i]), hdr[i][6:]) for i in st), reverse=True)[:2]
        print(f"   {n:5d} inst {100 * int(r[ia]) / tot:5.2f}% smp {100 * int(r[isamp]) / ts:5.2f}% "
              f"[{stalls[0][1]},{stalls[1][1]}] {r[isrc].strip()[:80]}")


if __name__ == "__main__":
    main()
