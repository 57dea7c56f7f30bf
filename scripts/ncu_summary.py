"""Summarise ncu outputs: launch list shares, key raw metrics and the hottest SASS blocks (development aid)."""
import collections, csv, subprocess, sys

def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    h = rows[0]; ki = h.index('Kernel Name'); vi = h.index('Metric Value'); mi = h.index('Metric Name')
    agg = collections.defaultdict(lambda: [0, 0.0])
    for r in rows[1:]:
        if r[mi] != 'gpu__time_duration.sum': continue  # lists taken with several metrics: time only
        try: v = float(r[vi].replace(',', ''))
        except ValueError: continue
        k = r[ki].split('(')[0][-48:]
        agg[k][0] += 1; agg[k][1] += v
    tot = sum(v[1] for k, v in agg.items() if 'int32_peak' not in k)
    for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
        print("%-50s n=%3d total=%9.3f ms avg=%8.3f ms share(excl. peak)=%.4f" % (k, v[0], v[1] / 1e6, v[1] / 1e6 / v[0], v[1] / tot))

WANT = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_alu.sum.pct', 'sm__inst_executed_pipe_fma.sum.pct',
        'sm__inst_executed_pipe_lsu.sum.pct_of_peak_sustained_active', 'dram__bytes_read.sum ', 'dram__bytes_write.sum ',
        'smsp__issue_active.avg.pct', 'smsp__thread_inst_executed_per_inst_executed.ratio', 'launch__grid_size',
        'launch__occupancy_limit', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum ', 'lts__t_sector_hit_rate.pct',
        'smsp__average_warp_latency_issue_stalled', 'smsp__average_warps_issue_stalled']

def raw(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h, u, v = rows[0], rows[1], rows[-1]
    for i, n in enumerate(h):
        if any((n + ' ').startswith(w) for w in WANT): print("%-70s %-14s %s" % (n, u[i], v[i]))

def sass(rep, top=12):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = rows[1]; data = rows[2:]
    ia, isrc, ie, it, iss = (h.index(x) for x in ('Address', 'Source', 'Instructions Executed', 'Thread Instructions Executed', '# Samples'))
    tot = sum(int(r[ie]) for r in data); ts = sum(int(r[iss]) for r in data)
    blocks, cur = [], []
    for r in data:
        cur.append(r)
        if any(x in r[isrc] for x in ('BRA', 'EXIT', 'RET', 'BSYNC', 'CALL')): blocks.append(cur); cur = []
    if cur: blocks.append(cur)
    res = []
    for b in blocks:
        e = sum(int(r[ie]) for r in b); t = sum(int(r[it]) for r in b); s = sum(int(r[iss]) for r in b)
        res.append((e, t, s, len(b), b[0][ia][-5:], sum('VIADDMNMX' in r[isrc] for r in b)))
    res.sort(reverse=True)
    print("total warp-instr %d samples %d" % (tot, ts))
    for e, t, s, n, a, v in res[:top]:
        print("addr %s n=%4d viaddmnmx=%2d instr share %.3f thr/instr %.1f samples share %.3f" % (a, n, v, e / tot, t / max(e, 1), s / ts))

if __name__ == '__main__':
    for a in sys.argv[1:]:
        if a.endswith('.csv'): launches(a)
        else: raw(a); sass(a)
