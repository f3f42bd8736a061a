"""Dev helper: STATIC instruction footprint of the executed code of a kernel, per source
function, from an .ncu-rep captured with -lineinfo/--import-source (i-cache budgeting:
B200's L1.5 instruction cache is ~32 KB = 2048 SASS instructions).
usage: python scripts/ncu_footprint.py REP [warps_per_launch]"""
import bisect, csv, io, os, re, subprocess, sys
rep = sys.argv[1]
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ut_multirobot_sim_b200", "csrc")
funcs = {}
for fn in ("umrs_kernels.cu", "umrs_device.cuh"):
    starts = []
    for n, line in enumerate(open(os.path.join(root, fn)), 1):
        m = re.match(r"^(?:__device__|__global__|template|static|cudaError_t|size_t).*?(\w+)\s*\(", line)
        if m and not line.startswith(" "):
            starts.append((n, m.group(1)))
    funcs[fn] = starts
def fname(f, ln):
    if f in funcs and funcs[f]:
        k = bisect.bisect_right([s[0] for s in funcs[f]], ln) - 1
        return f.split("_")[-1] + ":" + (funcs[f][k][1] if k >= 0 else "?")
    return f
cur_file, cur_line = None, None
seen = {}
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "File Path":
        cur_file = r[1].split("/")[-1]; continue
    if len(r) > 8 and r[0].isdigit():
        cur_line = int(r[0]); continue
    if len(r) > 8 and r[2].startswith("0x"):
        a = int(r[2], 16)
        if a not in seen:      # first (outermost file order) attribution only
            seen[a] = (fname(cur_file, cur_line), int(r[7]), r[3].strip())
tot = sum(v[1] for v in seen.values())
warps = float(sys.argv[2]) if len(sys.argv) > 2 else max(v[1] for v in seen.values() if "EXIT" in v[2] or True) / 1.0
W = float(sys.argv[2]) if len(sys.argv) > 2 else 65536.0
by = {}
for a, (f, n, s) in seen.items():
    b = by.setdefault(f, [0, 0, 0, 0])
    b[0] += 1
    if n >= 0.25 * W: b[1] += 1
    elif n > 0: b[2] += 1
    b[3] += n
print("%-34s %7s %7s %7s %8s" % ("function", "static", "hot", "warm", "dyn %"))
for f, b in sorted(by.items(), key=lambda kv: -kv[1][1]):
    print("%-34s %7d %7d %7d %8.2f" % (f, b[0], b[1], b[2], 100.0 * b[3] / tot))
print("total static %d, hot %d (%.1f KB), executed %d" % (
    len(seen), sum(b[1] for b in by.values()), sum(b[1] for b in by.values()) * 16 / 1024.0,
    sum(b[1] + b[2] for b in by.values())))
