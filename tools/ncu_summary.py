"""Condense an `ncu --page raw --csv` export: one line per launch with the metrics that matter here."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, data = rows[0], rows[2:]
c = lambda n: hdr.index(n)
keep = [('gpu__time_duration.sum', 'us'), ('launch__grid_size', 'grid'), ('launch__registers_per_thread', 'regs'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'occ%'), ('smsp__inst_executed.sum', 'winst'),
        ('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'fp64%'),
        ('SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg', 'dmma_cyc'),
        ('TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed', 'tensor%'),
        ('smsp__issue_active.avg.pct', 'issue%'), ('dram__bytes_read.sum', 'rdMB'), ('dram__bytes_write.sum', 'wrMB'),
        ('l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'bankconf'),
        ('smsp__thread_inst_executed_per_inst_executed.ratio', 'thr/inst'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'st_long'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'st_short'),
        ('smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'st_bar'),
        ('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'st_wait'),
        ('smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'st_math'),
        ('smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio', 'st_mio'),
        ('smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio', 'st_lg'),
        ('smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio', 'st_br'),
        ('smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio', 'st_noinst'),
        ('smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio', 'st_disp')]
keep = [(k, n) for k, n in keep if k in hdr]
minus = float(sys.argv[2]) if len(sys.argv) > 2 else 0.0
cyc = 'smsp__cycles_elapsed.avg' if 'smsp__cycles_elapsed.avg' in hdr else ('sm__cycles_elapsed.avg' if 'sm__cycles_elapsed.avg' in hdr else None)
dm = 'SM_C.TriageCompute.smsp__pipe_tensor_subpipe_dmma_cycles_active.avg'
for r in data:
    if float(r[c('gpu__time_duration.sum')]) < minus:
        continue
    out = [r[c('Kernel Name')][:34].ljust(34)]
    if cyc and dm in hdr:      # share of the elapsed cycles in which the FP64 tensor sub-pipe (DMMA) of an SM sub-partition was busy
        try:
            out.append('dmma%%=%.3g' % (100.0 * float(r[c(dm)].replace(',', '')) / float(r[c(cyc)].replace(',', ''))))
        except (ValueError, ZeroDivisionError):
            pass
    for k, n in keep:
        v = r[c(k)]
        try:
            v = '%.4g' % float(v)
        except ValueError:
            pass
        out.append('%s=%s' % (n, v))
    print(' '.join(out))
