"""Per-source-line share of executed warp instructions and stall samples from an .ncu-rep (needs -lineinfo).
    python profiles/ncu_lines.py gpurun_out/prof.ncu-rep <kernel-regex> [launch-index] [top-N]"""
import csv, io, subprocess, sys

def main(path, kre, idx="1", top="40"):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass",
                          "--kernel-id", f"::regex:{kre}:{idx}"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    fname, hdr, data = "", None, []
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            fname = r[1].split("/")[-1]
        elif r[0] == "Line No":
            hdr = r
            ie, ss = hdr.index("Instructions Executed"), hdr.index("# Samples")
        elif hdr and len(r) > ie and r[2] == "-" and r[0].isdigit():  # a CUDA source line (aggregated over its SASS)
            try:
                data.append((int(r[ie]), int(r[ss]), fname, r[0], r[1].strip()))
            except ValueError:
                pass
    ti, ts = sum(d[0] for d in data) or 1, sum(d[1] for d in data) or 1
    print(f"total warp instructions {ti}, stall samples {ts}")
    for d in sorted(data, reverse=True)[:int(top)]:
        print(f"{100 * d[0] / ti:5.1f}% inst {100 * d[1] / ts:5.1f}% samp  {d[2]}:{d[3]:>4s}  {d[4][:100]}")

if __name__ == "__main__":
    main(*sys.argv[1:])
