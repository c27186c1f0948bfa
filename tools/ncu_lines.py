"""Aggregate an ncu report's source page by CUDA source line and by kernel phase.
    python tools/ncu_lines.py gpurun_out/prof.ncu-rep [top_n] [kernel-name-substring]"""
import csv
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 30
only = sys.argv[3] if len(sys.argv) > 3 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))


def toi(s):
    try:
        return int(s)
    except ValueError:
        return 0


agg, cur, first_kernel, nk, skip = {}, None, True, 0, False
stall_cols = {}
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if r[0] == "Function Name":
        skip = only is not None and only not in r[1]
        continue
    if r[0] == "Line No":
        i_s, i_i, i_t = r.index("# Samples"), r.index("Instructions Executed"), r.index("Thread Instructions Executed")
        stall_cols = {n: i for i, n in enumerate(r) if n.startswith("stall_") and "Not Issued" not in n}
        continue
    if r[0].isdigit() and not skip:
        key = (cur, int(r[0]))
        a = agg.setdefault(key, [r[1], 0, 0, 0, {}])
        a[1] += toi(r[i_s]); a[2] += toi(r[i_i]); a[3] += toi(r[i_t])
        for n, i in stall_cols.items():
            a[4][n] = a[4].get(n, 0) + toi(r[i])
ts = sum(a[1] for a in agg.values()); ti = sum(a[2] for a in agg.values())
print(f"total samples {ts} warp-inst {ti}")
for key, a in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    st = sorted(a[4].items(), key=lambda x: -x[1])[:2]
    sts = " ".join(f"{n[6:]}:{100 * v / max(a[1], 1):.0f}%" for n, v in st)
    print(f"{key[0][:14]:14s}:{key[1]:4d} samp {100 * a[1] / ts:5.1f}% inst {100 * a[2] / ti:5.1f}% thr {a[3] / max(a[2], 1):4.1f} [{sts}] | {a[0].strip()[:70]}")
# phases by marker comments in update_depth.cu
src = [(k[1], a) for k, a in agg.items() if k[0] == "update_depth.cu"]
marks = []
for ln, line in enumerate(open("tadataka_b200/csrc/update_depth.cu"), 1):
    if "----- phase" in line or "----- store" in line:
        marks.append((ln, line.strip(" /-\n")[:28]))
    if "template <int G>" in line:
        marks.append((ln, "prologue"))
marks.sort()
print("\nby phase (update_depth.cu lines only; inlined tdk_common.cuh lines listed separately):")
for i, (ln, name) in enumerate(marks):
    hi = marks[i + 1][0] if i + 1 < len(marks) else 10 ** 9
    s = sum(a[1] for l, a in src if ln <= l < hi); ins = sum(a[2] for l, a in src if ln <= l < hi)
    print(f"  {name:30s} samp {100 * s / ts:5.1f}% inst {100 * ins / ti:5.1f}%")
oth = [(k, a) for k, a in agg.items() if k[0] != "update_depth.cu"]
print(f"  {'other files (inlined helpers)':30s} samp {100 * sum(a[1] for k, a in oth) / ts:5.1f}% inst {100 * sum(a[2] for k, a in oth) / ti:5.1f}%")
