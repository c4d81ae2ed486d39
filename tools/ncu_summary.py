"""Prints key metrics and the per-instruction hot spots of an .ncu-rep (development aid)."""
import csv, subprocess, sys, io
rep = sys.argv[1]; kern = sys.argv[2] if len(sys.argv) > 2 else None
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sectors_op_read.sum','lts__t_sectors_op_write.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum','l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum','l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum','l1tex__t_requests_pipe_lsu_mem_global_op_st.sum','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__occupancy_limit_registers','launch__occupancy_limit_shared_mem','sm__throughput.avg.pct_of_peak_sustained_elapsed','gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__inst_executed.sum','lts__t_sectors_srcunit_tex_op_read.sum','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__inst_executed_pipe_lsu.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','smsp__pcsamp_warps_issue_stalled_long_scoreboard','smsp__pcsamp_warps_issue_stalled_lg_throttle','smsp__pcsamp_warps_issue_stalled_short_scoreboard','smsp__pcsamp_warps_issue_stalled_not_selected','smsp__pcsamp_warps_issue_stalled_wait','smsp__pcsamp_warps_issue_stalled_branch_resolving','smsp__pcsamp_warps_issue_stalled_math_pipe_throttle','smsp__pcsamp_warps_issue_stalled_mio_throttle','smsp__pcsamp_warps_issue_stalled_no_instructions','smsp__pcsamp_warps_issue_stalled_selected','smsp__pcsamp_sample_buffer']
idx = {h: i for i, h in enumerate(hdr)}
for r in rows[2:]:
    if kern and kern not in r[idx['Kernel Name']]: continue
    print('----')
    for w in want:
        if w in idx: print(w, '=', r[idx[w]][:90], units[idx[w]])
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"] + (["--kernel-name", "regex:" + kern] if kern else []), capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
h = None
data = []
for r in rows:
    if 'Source' in r and '# Samples' in r: h = r; continue
    if h and len(r) == len(h) and r[h.index('# Samples')].isdigit():
        data.append((int(r[h.index('# Samples')]), int(r[h.index('Instructions Executed')]), float(r[h.index('Avg. Threads Executed')]), r[h.index('Source')].strip()))
tot = sum(d[0] for d in data) or 1; toti = sum(d[1] for d in data) or 1
print("instructions", len(data), "warp-inst", toti)
reg = []
for n, (s, ie, at, srcl) in enumerate(data):
    key = 'full' if at >= 24 else ('mid' if at >= 8 else 'low')
    if reg and reg[-1][0] == key: reg[-1][1] += s; reg[-1][2] += ie; reg[-1][4] = n
    else: reg.append([key, s, ie, n, n])
for key, s, ie, a, b in reg:
    if 100 * s / tot > 1 or 100 * ie / toti > 1: print("%-5s instr %3d-%3d samples %5.1f%% inst %5.1f%%" % (key, a, b, 100 * s / tot, 100 * ie / toti))
top = sorted(range(len(data)), key=lambda n: -data[n][0])[:int(sys.argv[3]) if len(sys.argv) > 3 else 30]
for n in sorted(top): print("%3d %5.2f%% inst=%5.2f%% thr=%5.1f  %s" % (n, 100 * data[n][0] / tot, 100 * data[n][1] / toti, data[n][2], data[n][3][:90]))
