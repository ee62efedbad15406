"""Turns the .ncu-rep captures of profiles/capture_rN.sh into small CSV summaries under profiles/ (run where ncu exists):
    python profiles/summarise.py gpurun_out r2"""
import csv, glob, io, os, subprocess, sys
KEYS = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tmem.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active"]
src = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out"
tag = sys.argv[2] if len(sys.argv) > 2 else "r2"
out = os.path.dirname(os.path.abspath(__file__))
for rep in sorted(glob.glob(os.path.join(src, f"*_{tag}.ncu-rep"))):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    if len(rows) < 3:
        continue
    hdr, units = rows[0], rows[1]
    idx = [hdr.index(k) for k in KEYS if k in hdr]
    name = os.path.basename(rep).replace(".ncu-rep", "_ncu.csv")
    with open(os.path.join(out, name), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in idx]); w.writerow([units[i] for i in idx])
        for r in rows[2:]:
            w.writerow([r[i][:70] for i in idx])
    print("wrote", name)

# dram bytes per launch of the per-layer GEMM shapes -> gemm_traffic_r1.json (bench.py reports it as roofline.traffic)
step = os.path.join(out, f"step_{tag}_ncu.csv")
if os.path.exists(step):
    import json
    rows = list(csv.reader(open(step)))
    h = rows[0]
    ni, ri, wi = h.index("Kernel Name"), h.index("dram__bytes_read.sum"), h.index("dram__bytes_write.sum")
    scale = {"Mbyte": 1e6, "Gbyte": 1e9, "Kbyte": 1e3, "byte": 1.0}[rows[1][ri]]
    per = sorted((float(r[ri]) + float(r[wi])) * scale for r in rows[2:] if "gemm_tc" in r[ni])
    json.dump({"kernel": "gemm_tc_kernel / gemm_tc2_kernel",
               "source": f"profiles/step_{tag}_ncu.csv (ncu --set full, one launch each of the per-layer GEMM shapes of ViT-B/32 at batch 512)",
               "dram_bytes_per_launch_mean": sum(per) / len(per), "per_launch": per}, open(os.path.join(out, f"gemm_traffic_{tag}.json"), "w"), indent=1)
    print(f"wrote gemm_traffic_{tag}.json", [round(x / 1e6) for x in per])
