"""Top source lines of one kernel by executed warp instructions, from `ncu -i rep --page source --csv --print-source cuda,sass`."""
import csv, sys, collections, subprocess
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kern, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
agg, samp, total, hdr, curfile = {}, {}, 0, None, ""
for r in rows:
    if len(r) == 2 and r[0] == "File Path": curfile = r[1]; continue
    if len(r) > 5 and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) == len(hdr):
        try: n = int(r[hdr.index("Instructions Executed")])
        except ValueError: continue
        if not r[1].strip(): continue
        key = (curfile.split("/")[-1], r[0], r[1].strip()[:100])
        agg[key] = agg.get(key, 0) + n
        try: samp[key] = samp.get(key, 0) + int(r[hdr.index("# Samples")])
        except ValueError: pass
        total += n
ts = sum(samp.values()) or 1
print("total warp instructions", total)
for (f, l, s), n in sorted(agg.items(), key=lambda kv: -kv[1])[:top]:
    print("%5.1f%% inst %5.1f%% samples  %s:%s  %s" % (100.0 * n / total, 100.0 * samp.get((f, l, s), 0) / ts, f, l, s))
