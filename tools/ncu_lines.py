"""Stall samples of an ncu capture aggregated by CUDA source line (run here, no GPU needed).

    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top] [inst]      # "inst": rank by warp instructions executed instead

Needs a capture made with ``--import-source on`` of a binary compiled with ``-lineinfo``.
"""
import csv
import subprocess
import sys


def main():
    rep, top = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
    by_inst = len(sys.argv) > 3 and sys.argv[3] == "inst"
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                         capture_output=True, text=True).stdout
    cur, hdr, agg = None, None, {}
    for r in csv.reader(out.splitlines()):
        if len(r) == 2 and r[0] == "File Path":
            cur = r[1].split("/")[-1]
        elif r and r[0] == "Line No" and "# Samples" in r:
            hdr, si, ei = r, r.index("# Samples"), r.index("Instructions Executed")
        elif hdr and len(r) == len(hdr) and r[0].isdigit():
            try:
                s, e = float(r[si] or 0), float(r[ei] or 0)
            except ValueError:
                continue
            a = agg.setdefault((cur, int(r[0]), r[1][:100]), [0.0, 0.0])
            a[0] += s
            a[1] += e
    tot = sum(v[0] for v in agg.values()) or 1.0
    tot_i = sum(v[1] for v in agg.values()) or 1.0
    print(f"total samples {tot:.0f}, warp instructions executed {tot_i:.0f}")
    col = 1 if by_inst else 0
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][col])[:top]:
        print(f"{100 * v[0] / tot:5.1f}% inst={int(v[1]):>9} ({100 * v[1] / tot_i:4.1f}%) {k[0]}:{k[1]} {k[2]}")


if __name__ == "__main__":
    main()
