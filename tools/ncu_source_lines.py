"""Per-source-line stall samples / instruction counts of an .ncu-rep captured with --import-source on.
Usage: python tools/ncu_source_lines.py rep.ncu-rep [top_n]"""
import csv, subprocess, sys
def main(path, top=40):
    out = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                         capture_output=True, text=True).stdout
    hdr = None; lines = []
    for raw in out.splitlines():
        if raw.startswith('"Line No"'):
            hdr = next(csv.reader([raw])); continue
        if hdr is None or not raw.startswith('"'): continue
        # source text may hold quotes: split the fixed numeric tail off from the right
        tail = raw.rsplit('","', len(hdr) - 2)
        if len(tail) != len(hdr) - 1: continue
        head = tail[0].split('","', 1)
        if len(head) != 2: continue
        r = [head[0].lstrip('"'), head[1]] + [t.rstrip('"') for t in tail[1:]]
        if r[0] == "": continue            # SASS rows
        lines.append(r)
    iS = hdr.index("# Samples"); iI = hdr.index("Instructions Executed")
    stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[iS] or 0) for r in lines) or 1
    totI = sum(int(r[iI] or 0) for r in lines) or 1
    print(f"total samples {tot}, warp instructions {totI}")
    lines.sort(key=lambda r: -int(r[iS] or 0))
    for r in lines[:top]:
        st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:3]
        print(f"{int(r[iS]) / tot * 100:5.1f}% smp {int(r[iI] or 0) / totI * 100:5.1f}% inst  L{r[0]:>4}  {r[1].strip()[:90]}   | " +
              ", ".join(f"{n}:{c}" for c, n in st if c))
if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
