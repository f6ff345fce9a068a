"""Per-source-line stall samples of one kernel in an ncu report (needs -lineinfo + --import-source on).

    python tools/ncu_lines.py report.ncu-rep <kernel-id> [top]
"""
import csv
import subprocess
import sys

rep, kid = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", ] + (["--kernel-id", ":::" + kid] if kid != "0" else []),
                     capture_output=True, text=True).stdout
rows, fname, hdr, total = [], "", None, 0
for r in csv.reader(out.splitlines()):
    if not r:
        continue
    if r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0] not in ("", "Function Name", "Kernel Name") and len(r) == len(hdr):
        d = dict(zip(hdr[2:], r[2:]))
        try:
            s = int(d["# Samples"]); ins = int(d["Instructions Executed"]); thr = int(d["Thread Instructions Executed"])
        except (KeyError, ValueError):
            continue
        total += s
        rows.append((s, ins, thr, fname, r[0], r[1].strip()[:110], int(d.get("stall_long_sb", 0) or 0)))
seen, uniq = set(), []
for x in rows:
    if (x[3], x[4]) in seen:
        continue
    seen.add((x[3], x[4])); uniq.append(x)
total = sum(x[0] for x in uniq)
print("total samples", total)
for s, ins, thr, f, ln, src, lsb in sorted(uniq, key=lambda x: -x[0])[:top]:
    print("%5.1f%% %7d inst %5.1f thr/inst lsb %5d  %s:%s  %s" % (100.0 * s / max(total, 1), ins, thr / max(ins, 1), lsb, f, ln, src))
