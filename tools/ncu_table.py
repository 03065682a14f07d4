"""Markdown table from a tools/ncu_summary.py dump: one row per captured kernel launch.
Usage: python tools/ncu_table.py summary.txt [algorithmic_bytes ...] (one per launch, optional)."""
import sys

CONV = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def val(block, key):
    for line in block.splitlines():
        if line.strip().startswith(key + " "):
            p = line.split()
            return float(p[1]), (p[2] if len(p) > 2 else "")
    return None, None


def main():
    blocks = open(sys.argv[1]).read().split("## ")[1:]
    alg = [float(a) for a in sys.argv[2:]]
    print("| kernel | grid x block | µs | dram read + write (MB) | algorithmic GB/s | dram / algorithmic | DRAM busy % | issue slots busy % | warps active % | registers |")
    print("|---|---|---|---|---|---|---|---|---|---|")
    for i, b in enumerate(blocks):
        name = b.splitlines()[0].split("(")[0].replace("void ", "").strip()
        t, _ = val(b, "gpu__time_duration.sum")
        r, ru = val(b, "dram__bytes_read.sum")
        w, wu = val(b, "dram__bytes_write.sum")
        tot = r * CONV[ru] + w * CONV[wu]
        iss, _ = val(b, "smsp__issue_active.avg.pct_of_peak_sustained_active")
        wa, _ = val(b, "sm__warps_active.avg.pct_of_peak_sustained_active")
        dr, _ = val(b, "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed")
        rg, _ = val(b, "launch__registers_per_thread")
        gs, _ = val(b, "launch__grid_size")
        bs, _ = val(b, "launch__block_size")
        a = alg[i] if i < len(alg) else None
        gbs = f"{a / t / 1e3:.0f}" if a else "-"
        ratio = f"{tot / a:.3f}" if a else "-"
        print(f"| `{name}` | {int(gs)} x {int(bs)} | {t:.1f} | {tot / 1e6:.1f} | {gbs} | {ratio} | {dr:.0f} | {iss:.0f} | {wa:.0f} | {int(rg)} |")


if __name__ == "__main__":
    main()
