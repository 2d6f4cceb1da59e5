"""Text summary of an `ncu --set full --import-source on` report (run where ncu is installed; no GPU needed):
    python tools/ncu_summary.py gpurun_out/prof_x.ncu-rep > profiles/r01_ncu_x.md
Key throughput / occupancy / DRAM metrics of each captured launch, the executed-instruction mix and the SASS lines
with the most stall samples."""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "l1tex__m_xbar2l1tex_read_bytes.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed"]


def page(rep, name):
    out = subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def main():
    rep = sys.argv[1]
    raw = page(rep, "raw")
    hdr, units = raw[0], raw[1]
    print(f"# ncu summary of `{rep}`\n")
    for r in raw[2:]:
        print(f"## {r[hdr.index('Kernel Name')][:110]}\n")
        print("| metric | value | unit |\n|---|---|---|")
        for k in KEYS:
            if k in hdr:
                print(f"| {k} | {r[hdr.index(k)]} | {units[hdr.index(k)]} |")
        stalls = [(h.replace("smsp__pcsamp_warps_issue_stalled_", ""), int(float(r[i] or 0))) for i, h in enumerate(hdr)
                  if h.startswith("smsp__pcsamp_warps_issue_stalled_") and not h.endswith("_not_issued")]
        stalls = sorted([s for s in stalls if s[1] > 0], key=lambda s: -s[1])
        print("\nstall samples: " + ", ".join(f"{n} {v}" for n, v in stalls[:10]) + "\n")
    src = page(rep, "source")
    if len(src) > 2:
        h = src[1]
        i_s, i_e, i_src = h.index("# Samples"), h.index("Instructions Executed"), h.index("Source")
        data = src[2:]
        tot = sum(int(r[i_e]) for r in data) or 1
        mix = collections.Counter()
        for r in data:
            t = r[i_src].strip().split()
            if t:
                mix[t[1] if t[0].startswith("@") else t[0]] += int(r[i_e])
        print("## executed instruction mix (first launch, warp-level)\n")
        print(", ".join(f"{k} {100 * v / tot:.1f}%" for k, v in mix.most_common(14)) + f"  (total {tot})\n")
        print("## SASS lines with the most stall samples\n\n| samples | executed | instruction |\n|---|---|---|")
        for r in sorted(data, key=lambda r: -int(r[i_s]))[:16]:
            print(f"| {r[i_s]} | {r[i_e]} | `{r[i_src].strip()[:100]}` |")


if __name__ == "__main__":
    main()
