"""Summaries of ncu outputs (read on the CPU box): launch list CSV -> per-kernel table; .ncu-rep -> key raw
metrics + SASS opcode mix + stall totals.  Usage: ncu_summary.py launches <csv> | rep <file.ncu-rep> [superblocks]"""
import collections
import csv
import io
import re
import subprocess
import sys

RAW = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
       'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active',
       'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'launch__shared_mem_per_block_dynamic',
       'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_registers', 'sm__throughput.avg.pct_of_peak_sustained_elapsed',
       'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum', 'sm__cycles_elapsed.max',
       'smsp__issue_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct',
       'sm__inst_executed_pipe_tensor.sum', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
       'smsp__cycles_active.avg', 'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active',
       'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum']


def launches(path):
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hi = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    h, rows = rows[hi], rows[hi + 1:]
    ki, vi, ui = h.index('Kernel Name'), h.index('Metric Value'), h.index('Metric Unit')
    d = collections.defaultdict(list)
    for r in rows:
        d[re.sub(r'\(.*', '', r[ki])[-70:]].append(float(r[vi].replace(',', '')))
    tot = sum(sum(v) for v in d.values())
    print(f"{'kernel':70s} {'n':>5s} {'avg':>9s} {'min':>9s} {'max':>9s} {'share':>6s}  ({rows[0][ui]})")
    for k, v in sorted(d.items(), key=lambda kv: -sum(kv[1])):
        print(f"{k:70s} {len(v):5d} {sum(v)/len(v):9.1f} {min(v):9.1f} {max(v):9.1f} {100*sum(v)/tot:5.1f}%")


def rep(path, nsb=None):
    raw = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    h = rows[0]
    print('kernels:', [re.sub(r'\(.*', '', r[h.index('Kernel Name')])[-60:] for r in rows[2:]])
    for w in RAW:
        if w in h:
            i = h.index(w)
            print(f'{w:70s}', [r[i] for r in rows[1:]])
    src = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--print-source', 'sass'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(src)))
    hidx = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
    h = rows[hidx[0]]
    end = hidx[1] - 1 if len(hidx) > 1 else len(rows)
    body = [r for r in rows[hidx[0] + 1:end] if len(r) == len(h)]
    ie, si, s = h.index('Instructions Executed'), h.index('# Samples'), h.index('Source')
    tot, tots = sum(int(r[ie]) for r in body), sum(int(r[si]) for r in body)
    print(f'first launch: {len(body)} SASS lines, {tot} warp instructions, {tots} samples' +
          (f', {tot * 32 / nsb:.0f} thread-instr per superblock' if nsb else ''))
    op, ops = collections.Counter(), collections.Counter()
    for r in body:
        m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_.]+)', r[s])
        k = m.group(2).split('.')[0] if m else r[s][:10]
        op[k] += int(r[ie]); ops[k] += int(r[si])
    for k, v in op.most_common(18):
        print(f'  {k:10s} {100*v/tot:5.1f}% of instr' + (f' {v*32/nsb:6.1f}/sb' if nsb else '') + f'  {100*ops[k]/max(tots,1):5.1f}% of samples')
    st = {c: sum(int(r[h.index(c)]) for r in body) for c in h if c.startswith('stall_') and 'Not Issued' not in c}
    print('  stalls:', {k[6:]: v for k, v in sorted(st.items(), key=lambda kv: -kv[1]) if v})
    top = sorted(body, key=lambda r: -int(r[si]))[:12]
    for r in top:
        print(f'   {r[si]:>5s} smp  ex {r[ie]:>8s}  {r[s].strip()[:80]}')


if __name__ == '__main__':
    if sys.argv[1] == 'launches':
        launches(sys.argv[2])
    else:
        rep(sys.argv[2], float(sys.argv[3]) if len(sys.argv) > 3 else None)
