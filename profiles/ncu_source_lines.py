"""Per-source-line summary of an .ncu-rep captured with --import-source on (kernels built with -lineinfo):
    python profiles/ncu_source_lines.py rep [top_n] [all_lines.csv]
Prints, per CUDA source line, stall samples, executed instructions, shared-memory wavefronts (LSU) and the excessive
(bank-conflict) ones, sorted by samples."""
import csv, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
cur_file, hdr, lines = None, None, []
for r in rows:
    if len(r) >= 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) >= 2 and r[0] == "Line No":
        hdr = r; continue
    if hdr is None or len(r) < len(hdr) or r[0] == "":
        continue
    d = dict(zip(hdr, r))
    def num(k):
        try: return float(d.get(k, "0") or 0)
        except ValueError: return 0.0
    lines.append((cur_file, r[0], r[1][:110], num("Warp Stall Sampling (All Samples)"), num("Instructions Executed"),
                  num("L1 Wavefronts Shared"), num("L1 Wavefronts Shared Excessive")))
tot = [sum(l[i] for l in lines) for i in (3, 4, 5, 6)]
print(f"totals: samples {tot[0]:.0f}  inst {tot[1]:.0f}  smem wavefronts {tot[2]:.0f}  excessive {tot[3]:.0f}")
if len(sys.argv) > 3:                      # every line in file order (file, line, samples, instructions, wavefronts, source)
    with open(sys.argv[3], "w", newline="") as fh:
        w = csv.writer(fh)
        w.writerow(["file", "line", "samples", "inst", "smem_wavefronts", "source"])
        for l in lines:
            if l[3] or l[4]:
                w.writerow([l[0], l[1], int(l[3]), int(l[4]), int(l[5]), l[2]])
for key, name in ((3, "stall samples"), (4, "instructions executed"), (5, "shared wavefronts"), (6, "excessive (bank-conflict) wavefronts")):
    print(f"\n== top lines by {name}")
    for l in sorted(lines, key=lambda l: -l[key])[:top]:
        if l[key] <= 0: break
        print(f"{l[0]}:{l[1]:>4s} samp {100*l[3]/max(tot[0],1):5.1f}%  inst {100*l[4]/max(tot[1],1):5.1f}%  wf {100*l[5]/max(tot[2],1):5.1f}%  exc {100*l[6]/max(tot[3],1):5.1f}%  | {l[2]}")
