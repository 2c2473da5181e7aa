"""Summarise an .ncu-rep here (no GPU): key raw metrics + stall breakdown + hottest SASS lines."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h = rows[0]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__t_requests_pipe_lsu_mem_local_op_ld.sum",
        "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct"]
for r in rows[2:]:
    print("==", r[h.index("Kernel Name")][:100])
    for w in want:
        if w in h:
            print("  %-70s %s %s" % (w, r[h.index(w)], rows[1][h.index(w)]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr, data = None, []
for r in rows:
    if r and r[0] == "Address":
        if hdr is not None:
            break
        hdr = r
        continue
    if hdr and len(r) == len(hdr):
        data.append(r)
if hdr:
    ix = {n: i for i, n in enumerate(hdr)}
    stalls = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
    tot = {s: sum(int(r[ix[s]] or 0) for r in data) for s in stalls}
    T = sum(tot.values()) or 1
    print("stalls %:", {k[6:]: round(100 * v / T, 1) for k, v in sorted(tot.items(), key=lambda x: -x[1]) if v > 0.01 * T})
    print("SASS instructions:", len(data))
    data.sort(key=lambda r: -int(r[ix["# Samples"]] or 0))
    for r in data[:int(sys.argv[2]) if len(sys.argv) > 2 else 25]:
        print("  %6s %-70s %s" % (r[ix["# Samples"]], r[ix["Source"]][:70],
                                  {s[6:]: r[ix[s]] for s in stalls if int(r[ix[s]] or 0) > 15}))
