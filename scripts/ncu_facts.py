"""Dev helper: profiles/r2_kernel_facts.json (the per-launch ncu facts bench.py quotes in
`roofline`) from an .ncu-rep of one k_world_step launch.  The facts carry a hash of the kernel
sources they were captured from; bench.py drops them (`"stale": true`) when the sources changed.
usage: python scripts/ncu_facts.py gpurun_out/prof_TAG.ncu-rep [summary file] > profiles/r2_kernel_facts.json"""
import csv, hashlib, io, json, os, subprocess, sys
rep = sys.argv[1]
summary = sys.argv[2] if len(sys.argv) > 2 else "profiles/r2_k_world_step_ncu_full.txt"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
h = hashlib.sha256()
for f in ("umrs_kernels.cu", "umrs_device.cuh"):
    h.update(open(os.path.join(ROOT, "ut_multirobot_sim_b200", "csrc", f), "rb").read())
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units, val = rows[0], rows[1], rows[2]
col = {h: i for i, h in enumerate(hdr)}


def get(name, scale_to=None):
    v = float(val[col[name]].replace(",", ""))
    u = units[col[name]]
    if scale_to == "byte":
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}[u.split("/")[0]]
    if scale_to == "us":
        v *= {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}[u]
    return v


worlds = 8192
facts = {
    "kernel": "k_world_step",
    "source": "%s (ncu --set full, one launch, 8192 worlds x 1081 rays, 1024-seg map)" % summary,
    "source_sha256": h.hexdigest(),
    "worlds": worlds,
    "rays_per_launch": worlds * 1081,
    "dram_bytes_read": int(round(get("dram__bytes_read.sum", "byte"))),
    "dram_bytes_write": int(round(get("dram__bytes_write.sum", "byte"))),
    "inst_executed": int(get("smsp__inst_executed.sum")),
    "gpu_time_duration_us": get("gpu__time_duration.sum", "us"),
    "issue_active_pct": get("smsp__issue_active.avg.pct_of_peak_sustained_active"),
    "registers_per_thread": int(get("launch__registers_per_thread")),
    "dynamic_smem_bytes": int(round(get("launch__shared_mem_per_block_dynamic", "byte"))),
    "ctas_per_sm": int(min(get("launch__occupancy_limit_registers"), get("launch__occupancy_limit_shared_mem"))),
    "fma_pipe_pct": get("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
    "alu_pipe_pct": get("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
    "xu_pipe_pct": get("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
    "lsu_pipe_pct": get("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
    "fp64_pipe_pct": get("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active"),
    "smem_wavefronts_pct_of_peak": get("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed"),
    "warps_active_pct": get("sm__warps_active.avg.pct_of_peak_sustained_active"),
}
print(json.dumps(facts, indent=1))
