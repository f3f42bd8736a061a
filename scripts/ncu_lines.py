"""Dev helper: per-source-line instruction / stall-sample shares of a kernel from an
.ncu-rep (needs -lineinfo).  usage: python scripts/ncu_lines.py gpurun_out/prof.ncu-rep [N]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_bytes.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
for i, h in enumerate(hdr):
    if h in want:
        print("%-75s %-12s %s" % (h, units[i], [r[i] for r in rows[2:]]))
for i, h in enumerate(hdr):
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio"):
        v = float(rows[2][i])
        if v > 0.15:
            print("  stall %-60s %.2f" % (h.split("issue_stalled_")[1].split("_per_issue")[0], v))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
cur, agg, seen = None, {}, set()
kern = 0
for r in csv.reader(io.StringIO(src)):
    if len(r) == 2 and r[0] == "Function Name":
        pass
    if len(r) == 2 and r[0] == "File Path":
        cur = r[1].split("/")[-1]
        continue
    if len(r) > 8 and r[0].isdigit() and r[2] == "-":
        key = (cur, int(r[0]))
        a = agg.setdefault(key, [r[1], 0, 0])
        a[1] += int(r[7]); a[2] += int(r[6])
tot = sum(v[1] for v in agg.values()) or 1
ts = sum(v[2] for v in agg.values()) or 1
print("sum over lines: inst %d samples %d (all profiled launches, nested inlines counted per level)" % (tot, ts))
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1])[:top]:
    print("%-20s %4d %6.2f%% inst %6.2f%% smp  %s" % (k[0], k[1], 100 * v[1] / tot, 100 * v[2] / ts, v[0][:90]))

# ---- per-function shares (file, first line of each function found by a crude scan) ----
import os, re, bisect
root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "ut_multirobot_sim_b200", "csrc")
funcs = {}
for fn in ("umrs_kernels.cu", "umrs_device.cuh"):
    starts = []
    for n, line in enumerate(open(os.path.join(root, fn)), 1):
        m = re.match(r"^(?:__device__|__global__|template|static|cudaError_t|size_t).*?(\w+)\s*\(", line)
        if m and not line.startswith(" "):
            starts.append((n, m.group(1)))
    funcs[fn] = starts
by = {}
for (f, ln), v in agg.items():
    name = f
    if f in funcs and funcs[f]:
        k = bisect.bisect_right([s[0] for s in funcs[f]], ln) - 1
        name = f + ":" + (funcs[f][k][1] if k >= 0 else "?")
    b = by.setdefault(name, [0, 0]); b[0] += v[1]; b[1] += v[2]
print("---- per function ----")
for k, v in sorted(by.items(), key=lambda kv: -kv[1][0])[:25]:
    print("%-45s %6.2f%% inst %6.2f%% smp" % (k, 100 * v[0] / tot, 100 * v[1] / ts))
