"""Summarise an .ncu-rep: headline metrics, stall totals, hottest instructions."""
import csv, io, subprocess, sys
rep = sys.argv[1]; ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 25
def page(name):
    """rep may be an .ncu-rep or the prefix of exported <prefix>_raw.csv / <prefix>_source.csv"""
    if rep.endswith(".ncu-rep"):
        return subprocess.run(["ncu", "-i", rep, "--page", name, "--csv"], capture_output=True, text=True).stdout
    return open(f"{rep}_{name}.csv").read()
raw = page("raw")
rows = list(csv.reader(io.StringIO(raw))); hdr, units, vals = rows[0], rows[1], rows[2]
KEYS = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
        'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_sector_pipe_lsu_mem_global_op_ld_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'smsp__thread_inst_executed_per_inst_executed.ratio']
for k in KEYS:
    if k in hdr:
        i = hdr.index(k); print(f"{k} = {vals[i]} {units[i]}")
src = page("source")
rows = list(csv.reader(io.StringIO(src))); hdr, body = rows[1], rows[2:]
g = hdr.index
names = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
tot = sum(int(r[g("# Samples")] or 0) for r in body)
print("instructions", len(body), "samples", tot)
print("  ".join(f"{n[6:]}={sum(int(r[g(n)] or 0) for r in body)}" for n in names))
for i, r in sorted(enumerate(body), key=lambda x: -int(x[1][g("# Samples")] or 0))[:ntop]:
    st = {n[6:]: int(r[g(n)] or 0) for n in names}
    big = sorted(st.items(), key=lambda kv: -kv[1])[:2]
    print(i, r[g("# Samples")], big, r[g("Source")][:90])
