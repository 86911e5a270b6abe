"""Per-source-line hot spots of one kernel from an `ncu --set full --import-source on` report (needs -lineinfo):
stall samples and executed warp instructions per CUDA line. usage: python profiles/ncu_lines.py rep.ncu-rep kernel [top]"""
import csv, io, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass,cuda", "--csv", "--kernel-name", kern], capture_output=True, text=True).stdout
cur, hdr, acc = None, None, {}
for r in csv.reader(io.StringIO(out)):
    if len(r) >= 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No":
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].strip().isdigit():
        i_s, i_i = hdr.index("# Samples"), hdr.index("Instructions Executed")
        try:
            s, i = int(r[i_s] or 0), int(r[i_i] or 0)
        except ValueError:
            continue
        k = (cur, int(r[0]))
        a = acc.setdefault(k, [0, 0, r[1].strip()[:105]])
        a[0] += s; a[1] += i
ts, ti = sum(a[0] for a in acc.values()) or 1, sum(a[1] for a in acc.values()) or 1
print(f"kernel {kern}: {ts} samples, {ti} warp instructions")
for (f, ln), (s, i, src) in sorted(acc.items(), key=lambda kv: -(kv[1][1] if len(sys.argv) > 4 else kv[1][0]))[:top]:
    print(f"{100*s/ts:5.1f}% smp {100*i/ti:5.1f}% inst  {f}:{ln}  {src}")
