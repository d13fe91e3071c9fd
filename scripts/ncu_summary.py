"""Summarise an .ncu-rep: key metrics per kernel + stall reasons / hottest SASS for one kernel.
usage: python scripts/ncu_summary.py REP [kernel-regex-for-source-page]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(io.StringIO(raw))); h = r[0]
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'l1tex__t_sector_hit_rate.pct',
        'launch__shared_mem_per_block_dynamic', 'launch__grid_size', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.avg',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_bytes.sum']
for row in r[2:]:
    print('----', row[h.index('Kernel Name')][:90])
    print('   ', ', '.join("%s=%s" % (n.split('.')[0].replace('__', '.'), row[h.index(n)]) for n in want if n in h))
if len(sys.argv) > 2:
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + sys.argv[2], "--launch-count", "1"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src))); hh = rows[1]; si = hh.index('# Samples')
    data = [x for x in rows[2:] if len(x) == len(hh) and x[si].isdigit()]
    reasons = [x for x in hh if x.startswith('stall_') and 'Not Issued' not in x]
    tot = sum(int(x[si]) for x in data)
    print('=====', rows[0][1][:80], 'samples', tot)
    for rs in sorted(reasons, key=lambda rs: -sum(int(x[hh.index(rs)]) for x in data))[:8]:
        t = sum(int(x[hh.index(rs)]) for x in data); print("  %-24s %5.1f%%" % (rs, 100 * t / tot))
    samples = [int(x[si]) for x in data]; ie = hh.index('Instructions Executed')
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    for i in sorted(sorted(range(len(data)), key=lambda i: -samples[i])[:n]):
        x = data[i]; top = max(reasons, key=lambda rs: int(x[hh.index(rs)]))
        print("  %6d  %-58s %7d %4.1f%% %-18s exec=%s" % (i, x[1].strip()[:58], samples[i], 100 * samples[i] / tot, top, x[ie]))
