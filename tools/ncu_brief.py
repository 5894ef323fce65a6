#!/usr/bin/env python
"""Print the handful of ncu metrics used while tuning, plus a SASS region table (exec/warp, share) for one kernel.
    python tools/ncu_brief.py gpurun_out/prof.ncu-rep [kernel-regex]"""
import csv, subprocess, sys
rep = sys.argv[1]; kre = sys.argv[2] if len(sys.argv) > 2 else None
args = ["ncu", "-i", rep, "--page", "raw", "--csv"] + (["--kernel-name", "regex:" + kre] if kre else [])
rows = list(csv.reader(subprocess.run(args, capture_output=True, text=True).stdout.splitlines()))
hdr, units = rows[0], rows[1]
want = ['Kernel Name','gpu__time_duration.sum','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','l1tex__t_sector_hit_rate.pct','l1tex__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed','lts__throughput.avg.pct_of_peak_sustained_elapsed','l1tex__m_l1tex2xbar_req_cycles_active.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','launch__registers_per_thread','launch__grid_size','l1tex__data_pipe_lsu_wavefronts_mem_shared.sum','l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum','sm__cycles_elapsed.max','launch__shared_mem_per_block_dynamic','launch__occupancy_limit_shared_mem','launch__occupancy_limit_registers']
for r in rows[2:]:
    print('-----')
    for w in want:
        if w in hdr: print(' ', w, '=', r[hdr.index(w)], units[hdr.index(w)])
    st = [(h.split('stalled_')[1].replace('_per_issue_active.ratio',''), float(r[i])) for i,h in enumerate(hdr) if 'issue_stalled' in h and 'per_issue_active' in h and 'not_issued' not in h]
    print('  stalls/issue:', ' '.join('%s=%.2f'%(k,v) for k,v in sorted(st,key=lambda t:-t[1])[:8]))
if kre:
    args = ["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + kre]
    rows = list(csv.reader(subprocess.run(args, capture_output=True, text=True).stdout.splitlines()))
    hi = [i for i, r in enumerate(rows) if 'Source' in r and 'Instructions Executed' in r][0]
    hdr = rows[hi]; data = [r for r in rows[hi + 1:] if len(r) == len(hdr) and r[hdr.index('Instructions Executed')].isdigit()]
    for rep in (4, 2):  # the source page repeats the listing once per captured launch of the kernel
        n = len(data) // rep
        if n and len(data) % rep == 0 and [r[hdr.index('Source')] for r in data[:n]] == [r[hdr.index('Source')] for r in data[n:2 * n]]:
            data = data[:n]; break
    iS = hdr.index('Source'); iE = hdr.index('Instructions Executed')
    tot = sum(int(r[iE]) for r in data); mx = max(int(r[iE]) for r in data)
    nw = int(data[0][iE])
    print('total warp-instr', tot, ' warps', nw, ' instr/warp %.0f' % (tot / nw), ' static', len(data))
    segs = []; prev = None; start = 0; acc = 0
    for i, r in enumerate(data):
        lvl = int(r[iE]) / nw
        if prev is None: prev = lvl
        if abs(lvl - prev) > 0.2 * max(prev, 0.2):
            segs.append((start, i - 1, prev, acc)); start = i; acc = 0; prev = lvl
        acc += int(r[iE])
    segs.append((start, len(data) - 1, prev, acc))
    for s in segs:
        if s[3] / tot > float(sys.argv[3] if len(sys.argv)>3 else 0.015): print('  instr %4d-%4d exec/warp %5.2f share %4.1f%% | %s' % (s[0], s[1], s[2], 100 * s[3] / tot, data[s[0]][iS][:60]))
