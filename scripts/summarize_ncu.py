"""Summarise ncu outputs brought back in gpurun_out/ into profiles/ (tracked):
  python scripts/summarize_ncu.py launches gpurun_out/launches_r01.csv  > profiles/r01_launches.txt
  python scripts/summarize_ncu.py step     gpurun_out/launches_r01.csv ["label"] > profiles/r01_launches_step.txt
  python scripts/summarize_ncu.py kernel   gpurun_out/gemm_r01.ncu-rep  > profiles/r01_gemm.txt"""
import collections, csv, subprocess, sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "lts__t_sector_hit_rate.pct",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "sm__cycles_elapsed.max",
        "launch__shared_mem_per_block_dynamic", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum"]

def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    h, data = rows[hdr], rows[hdr + 1:]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in data:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
        k = r[ki].split("(")[0][:64]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    print("# %s: %d launches, %.1f us total (ncu per-launch times are cold-cache and serialised: compare SHARES)" % (path, len(data), tot))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-66s n=%4d %10.1f us %5.1f%%" % (k, v[0], v[1], 100 * v[1] / tot))

def step(path, label="cfg3, B=512"):
    """one steady-state training step out of a launch list: the launches between the last two adam kernels"""
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = [i for i, r in enumerate(rows) if r[0] == "ID"][0]
    h, data = rows[hdr], rows[hdr + 1:]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    adam = [i for i, r in enumerate(data) if "adam_kernel" in r[ki]]
    data = data[adam[-2] + 1:adam[-1] + 1]
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in data:
        v = float(r[vi].replace(",", ""))
        v = v / 1e3 if r[ui] == "ns" else (v * 1e3 if r[ui] == "ms" else v)
        k = r[ki].split("(")[0][:72]
        agg[k][0] += 1
        agg[k][1] += v
    tot = sum(v[1] for v in agg.values())
    gemm = sum(v[1] for k, v in agg.items() if "gemm_tf32_kernel" in k or "gemm_f16_kernel" in k)
    print("# one training step (%s) out of %s: %d launches, %.1f us total, GEMMs %.1f us = %.1f%% (ncu per-launch times are "
          "cold-cache and serialised, and the step's side streams run one kernel at a time under ncu: compare SHARES)" % (label, path, len(data), tot, gemm, 100 * gemm / tot))
    for k, v in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%-72s n=%4d %10.1f us %5.1f%%" % (k, v[0], v[1], 100 * v[1] / tot))

def kernel(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, units = rows[0], rows[1]
    for r in rows[2:]:
        print("# kernel:", r[h.index("Kernel Name")][:100])
        for i, c in enumerate(h):
            if c in KEYS or c == "smsp__issue_active.avg.pct_of_peak_sustained_active":
                print("%-72s %14s %s" % (c, r[i], units[i]))
        # warp stall reasons (warps stalled per issue-active cycle), largest first
        st = [(float(r[i].replace(",", "")), c) for i, c in enumerate(h) if c.startswith("smsp__average_warps_issue_stalled_") and c.endswith("_per_issue_active.ratio") and r[i]]
        print("# stall reasons (warps per issue-active cycle):")
        for v, c in sorted(st, reverse=True)[:8]:
            print("%-72s %14.3f" % (c.replace("smsp__average_warps_issue_stalled_", "stall_").replace("_per_issue_active.ratio", ""), v))

if __name__ == "__main__":
    {"launches": launches, "kernel": kernel, "step": step}[sys.argv[1]](*sys.argv[2:])
