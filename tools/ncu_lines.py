#!/usr/bin/env python
"""Per-source-line stall samples of one kernel of an .ncu-rep (needs -lineinfo and --import-source on):
    python tools/ncu_lines.py gpurun_out/x.ncu-rep k_wide_front [top]
Reads `ncu --page source --csv --print-source cuda,sass`; the CUDA-line rows carry the per-line sums."""
import csv
import subprocess
import sys


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-name", f"regex:{kern}"], capture_output=True, text=True).stdout
    cur_file, hdr, agg = None, None, []
    for r in csv.reader(raw.splitlines()):
        if not r:
            continue
        if r[0] == "File Path":
            cur_file = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
        elif hdr and r[0].isdigit():
            try:
                n = int(r[hdr.index("# Samples")])
            except ValueError:
                continue
            stalls = {}
            for i, h in enumerate(hdr):
                if h.startswith("stall_") and i < len(r) and r[i].isdigit() and int(r[i]):
                    stalls[h[6:]] = int(r[i])
            agg.append((n, cur_file, int(r[0]), r[1].strip(), stalls))
    tot = sum(a[0] for a in agg) or 1
    print(f"{kern}: {tot} samples")
    for n, f, ln, src, st in sorted(agg, key=lambda a: -a[0])[:top]:
        why = ", ".join(f"{k} {v}" for k, v in sorted(st.items(), key=lambda kv: -kv[1])[:3])
        print(f"{n / tot:6.3f} {f}:{ln:<4d} {src[:90]}   [{why}]")


if __name__ == "__main__":
    main()
