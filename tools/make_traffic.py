#!/usr/bin/env python3
"""profiles/rNN_traffic.json from ncu reports (--set full, or just the dram / duration metrics).

    python tools/make_traffic.py U=gpurun_out/prof_U.ncu-rep N=gpurun_out/prof_N.ncu-rep --frames 4 [--csv-prefix profiles/r01_ncu_full_4k]

Per image family and kernel: dram__bytes_read.sum + dram__bytes_write.sum of one launch divided by
the frames that launch processed (bench.py scales it back by its own frames per launch)."""
import csv, io, json, subprocess, sys, re

WANT = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed"]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def shown(name):
    name = name.replace("void ", "").replace("rct::", "")
    base = name.split("(")[0]
    base = re.sub(r"\(bool\)|\(int\)", "", base)
    m = re.match(r"k_apply_warp<(\d), *(\d+)", base)
    if m:
        return "k_apply_warp<%s,%s>" % ("bulk" if m.group(1) == "1" else "cp.async", m.group(2))
    return re.sub(r"<.*>", "", base)


def main():
    args = [a for a in sys.argv[1:] if "=" in a]
    frames = int(sys.argv[sys.argv.index("--frames") + 1]) if "--frames" in sys.argv else 4
    prefix = sys.argv[sys.argv.index("--csv-prefix") + 1] if "--csv-prefix" in sys.argv else None
    out = {}
    for a in args:
        fam, rep = a.split("=", 1)
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, data = rows[0], rows[1], rows[2:]
        idx = [hdr.index(w) for w in WANT if w in hdr]
        if prefix:
            with open(f"{prefix}_{fam}.csv", "w", newline="") as f:
                w = csv.writer(f)
                w.writerow([hdr[i] for i in idx]); w.writerow([units[i] for i in idx])
                for r in data:
                    w.writerow([r[i] for i in idx])
        ir, iw, it = hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum"), hdr.index("gpu__time_duration.sum")
        for r in data:
            b = float(r[ir]) * SCALE[units[ir]] + float(r[iw]) * SCALE[units[iw]]
            if not r[ir] or not r[iw]:
                continue
            out.setdefault(fam, {})[shown(r[hdr.index("Kernel Name")])] = {
                "dram_bytes_per_frame": b / frames, "ncu_duration": r[it] + " " + units[it], "frames_in_capture": frames}
    json.dump(out, sys.stdout, indent=1)


if __name__ == "__main__":
    main()
