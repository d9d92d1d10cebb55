"""Pick the counters we track out of an `ncu --page raw --csv` dump: python tools/ncusum.py raw.csv [out.csv]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
want = ['Kernel Name', 'Block Size', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'l1tex__t_sector_hit_rate.pct',
        'lts__t_sector_hit_rate.pct', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
        'l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed', 'l1tex__f_wavefronts.avg.pct_of_peak_sustained_elapsed',
        'l1tex__t_requests_pipe_tex_mem_texture.sum', 'l1tex__texin_sm2tex_req_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_tex.avg.pct_of_peak_sustained_active', 'l1tex__m_xbar2l1tex_read_bytes.sum',
        'gpc__cycles_elapsed.max', 'launch__occupancy_limit_registers', 'launch__waves_per_multiprocessor',
        'sm__maximum_warps_per_active_cycle_pct', 'sm__cycles_active.avg', 'sm__cycles_elapsed.avg', 'smsp__thread_inst_executed_per_inst_executed.ratio',
        'l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__inst_executed.avg.per_cycle_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'l1tex__t_sectors_pipe_tex_mem_texture.sum','l1tex__t_sectors_pipe_tex_mem_texture_lookup_miss.sum', 'l1tex__data_bank_reads.avg.pct_of_peak_sustained_elapsed']
want += [h for h in hdr if 'issue_stalled' in h and 'per_issue_active' in h]
out = csv.writer(open(sys.argv[2], 'w') if len(sys.argv) > 2 else sys.stdout)
out.writerow(['metric', 'unit'] + ['launch%d' % i for i in range(len(data))])
for i, h in enumerate(hdr):
    if h in want:
        out.writerow([h, units[i]] + [r[i] for r in data])
