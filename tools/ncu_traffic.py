"""Summarise one `ncu --set full` capture of a kernel's launches: DRAM bytes, duration and pipe utilisation per launch.

    python tools/ncu_traffic.py gpurun_out/sytrd_bench.ncu-rep profiles/r01_sytrd_traffic.json

Reads the report with `ncu -i ... --page raw --csv` (no GPU needed) and writes a small JSON that bench.py picks up
for `roofline.traffic` (dram__bytes_read.sum + dram__bytes_write.sum averaged per launch).
"""
import csv
import io
import json
import subprocess
import sys

UNIT = {'byte': 1.0, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9, 'Tbyte': 1e12,
        'ns': 1e-6, 'us': 1e-3, 'ms': 1.0, 's': 1e3, 'nsecond': 1e-6, 'usecond': 1e-3, 'msecond': 1.0, 'second': 1e3}
WANT = {
    'dram__bytes_read.sum': 'dram_read_bytes',
    'dram__bytes_write.sum': 'dram_write_bytes',
    'gpu__time_duration.sum': 'duration_ms',
    'lts__t_sector_hit_rate.pct': 'l2_hit_pct',
    'smsp__issue_active.avg.pct_of_peak_sustained_active': 'issue_active_pct',
    'sm__inst_executed_pipe_tensor_op_gmma.avg.pct_of_peak_sustained_active': 'tensor_pipe_pct',
    'sm__pipe_tensor_subpipe_tf32_cycles_active.avg.pct_of_peak_sustained_active': 'tensor_tf32_pct',
    'sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active': 'uniform_pipe_pct',
}


def main() -> None:
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    head, units = rows[0], rows[1]
    launches = []
    for r in rows[2:]:
        rec = {'kernel': r[head.index('Kernel Name')].split('(')[0], 'grid': r[head.index('Grid Size')]}
        for name, key in WANT.items():
            if name in head:
                i = head.index(name)
                try:
                    v = float(r[i].replace(',', '')) * UNIT.get(units[i], 1.0)
                    if v == v:
                        rec[key] = v
                except ValueError:
                    pass
        launches.append(rec)
    # ncu occasionally returns nan for a counter of one launch: such launches are left out of the mean and counted
    def dram(l):
        return l.get('dram_read_bytes', float('nan')) + l.get('dram_write_bytes', float('nan'))
    valid = [l for l in launches if dram(l) == dram(l)]
    for l in launches:
        if dram(l) != dram(l):
            l.pop('dram_read_bytes', None), l.pop('dram_write_bytes', None)
            l['dram_counters'] = 'nan in this capture'
    n = max(1, len(valid))
    total = sum(dram(l) for l in valid)
    summary = {
        'source': rep, 'launches': len(launches), 'launches_with_dram_counters': len(valid),
        'dram_bytes_total': total, 'dram_bytes_per_launch': total / n,
        'duration_ms_total_under_ncu': sum(l.get('duration_ms', 0.0) for l in launches),
        'per_launch': launches,
    }
    with open(out, 'w') as f:
        json.dump(summary, f, indent=1)
    print(json.dumps({k: v for k, v in summary.items() if k != 'per_launch'}))


if __name__ == '__main__':
    main()
