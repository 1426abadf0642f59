"""Turn ncu output into the summaries kept under profiles/.

    python scripts/ncu_summaries.py shares  <launches.csv>  "<title line>"  > profiles/rNN_launch_shares.md
    python scripts/ncu_summaries.py kernels <report.ncu-rep> "<title line>" > profiles/rNN_ncu_kernels.md
    python scripts/ncu_summaries.py traffic <report.ncu-rep>                > profiles/traffic.json

`launches.csv` is the `--metrics gpu__time_duration.sum --csv --log-file` launch list; `report.ncu-rep` a
`--set full` capture (read here with `ncu -i ... --page raw --csv`).
"""
import csv
import json
import subprocess
import sys
from collections import OrderedDict, defaultdict

WANTED = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum',
          'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
          'smsp__warps_eligible.avg.per_cycle_active', 'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread',
          'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
          'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
          'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
          'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
          'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio']


def shares(path, title):
    rows = [r for r in csv.reader(l for l in open(path, newline='') if l.startswith('"'))]
    head = rows[0]
    k_name, k_val, k_metric = head.index('Kernel Name'), head.index('Metric Value'), head.index('Metric Name')
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        if r[k_metric] != 'gpu__time_duration.sum':
            continue
        name = r[k_name].split('(')[0][:70]
        tot[name] += float(r[k_val].replace(',', ''))
        cnt[name] += 1
    total = sum(tot.values())
    print('# ' + title)
    print('kernel | launches | total ns | ns / launch | share\n---|---|---|---|---')
    for name in sorted(tot, key=tot.get, reverse=True)[:24]:
        print('%s | %d | %d | %d | %.1f%%' % (name, cnt[name], tot[name], tot[name] / cnt[name], 100 * tot[name] / total))


def raw_rows(rep):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(l for l in out.splitlines() if l.startswith('"')))
    return rows[0], rows[1], rows[2:]


def kernels(rep, title):
    head, units, rows = raw_rows(rep)
    k_name, k_grid, k_block = head.index('Kernel Name'), head.index('Grid Size'), head.index('Block Size')
    seen = OrderedDict()
    k_dur = head.index('gpu__time_duration.sum')
    for r in rows:                                           # the LONGEST captured launch of each kernel (the optimiser runs twice:
        name = r[k_name].split('(')[0]                       # the planes' pass and the decoder's few blocks)
        if name not in seen or float(r[k_dur].replace(',', '')) > float(seen[name][k_dur].replace(',', '')):
            seen[name] = r
    print('# ' + title)
    for name, r in seen.items():
        print('\n## %s  grid %s block %s' % (name, r[k_grid], r[k_block]))
        for m in WANTED:
            if m in head:
                print('%s = %s %s' % (m, r[head.index(m)], units[head.index(m)]))


def traffic(rep):
    head, units, rows = raw_rows(rep)
    k_name = head.index('Kernel Name')
    rd, wr = head.index('dram__bytes_read.sum'), head.index('dram__bytes_write.sum')
    scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
    out = OrderedDict()
    for r in rows:                                           # the launch with the most traffic of each kernel
        name = r[k_name].split('(')[0].replace('void ', '').replace('mne::', '')
        out[name] = max(out.get(name, 0), int(float(r[rd]) * scale[units[rd]] + float(r[wr]) * scale[units[wr]]))
    print(json.dumps({'dram_bytes_per_launch': out, 'source': 'ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, '
                      'largest captured launch of each kernel (%s)' % rep}, indent=1))


if __name__ == '__main__':
    mode = sys.argv[1]
    if mode == 'shares':
        shares(sys.argv[2], sys.argv[3])
    elif mode == 'kernels':
        kernels(sys.argv[2], sys.argv[3])
    elif mode == 'traffic':
        traffic(sys.argv[2])
    else:
        raise SystemExit(__doc__)
