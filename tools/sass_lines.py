"""Aggregates an ncu SASS source page per CUDA source line (line info from nvdisasm -g on the same cubin).
usage: sass_lines.py <report.ncu-rep> <libokrt.so> <kernel-substring> [topN]"""
import csv, os, re, subprocess, sys, tempfile, collections
rep, lib, kname = sys.argv[1], sys.argv[2], sys.argv[3]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(lib)], cwd=tmp, capture_output=True)
dis = []
for cubin in sorted(f for f in os.listdir(tmp) if "sm_100" in f):
    dis += subprocess.run(["nvdisasm", "-g", "-c", os.path.join(tmp, cubin)], capture_output=True, text=True).stdout.splitlines()
lines, cur, insec = [], None, False
for l in dis:
    if l.startswith(".text."):
        insec = kname in l
        continue
    if not insec:
        continue
    m = re.search(r'//## File ".*/([^/"]+)", line (\d+)(.*)', l)
    if m:
        cur = (m.group(1), int(m.group(2)), m.group(3).strip())
        continue
    if re.match(r"\s+/\*[0-9a-f]{4,}\*/", l):
        lines.append((cur, l.strip()))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hi = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
hdr = rows[hi]
body = []
for r in rows[hi + 1:]:
    if not r or r[0] in ("Kernel Name", "Address") or len(r) < len(hdr) // 2:
        break
    body.append(r)
ci = {n: hdr.index(n) for n in ("# Samples", "Instructions Executed", "Thread Instructions Executed", "Source")}
print("sass instructions: nvdisasm %d, ncu %d" % (len(lines), len(body)))
agg = collections.defaultdict(lambda: [0, 0, 0, 0])
for k, r in enumerate(body):
    key = lines[k][0] if k < len(lines) else None
    a = agg[key]
    a[0] += int(r[ci["# Samples"]] or 0); a[1] += int(r[ci["Instructions Executed"]] or 0); a[2] += int(r[ci["Thread Instructions Executed"]] or 0); a[3] += 1
tot_s = sum(a[0] for a in agg.values()); tot_i = sum(a[1] for a in agg.values()); tot_t = sum(a[2] for a in agg.values())
print("total samples %d, warp inst %.3g, thread inst %.3g, avg threads %.2f" % (tot_s, tot_i, tot_t, tot_t / max(1, tot_i)))
src = {}
def text(key):
    if key is None: return ""
    f = os.path.join(os.path.dirname(os.path.abspath(lib)), "csrc", key[0])
    if f not in src:
        try: src[f] = open(f).read().splitlines()
        except Exception: src[f] = []
    L = src[f]
    return L[key[1] - 1].strip()[:90] if 0 < key[1] <= len(L) else ""
print("%7s %7s %6s %5s %9s  %s" % ("samp%", "inst%", "thr/i", "sass", "Mexec", "line"))
for key, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%6.2f%% %6.2f%% %6.1f %5d %9.2f  %s:%s %s | %s" % (100.0 * a[0] / max(1, tot_s), 100.0 * a[1] / max(1, tot_i), a[2] / max(1, a[1]), a[3], a[1] / max(1, a[3]) / 1e6,
          key[0] if key else "?", key[1] if key else "?", (key[2][:30] if key else ""), text(key)))
