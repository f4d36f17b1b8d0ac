"""Summarise an .ncu-rep: headline metrics per kernel and executed-instruction / stall share per
CUDA source line (needs -lineinfo; lines come from `nvdisasm -g` of the cubin in the .so).
Usage: python profiles/ncu_lines.py gpurun_out/prof.ncu-rep <mangled kernel substring> [top]"""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

csv.field_size_limit(10 ** 9)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, 'collie_b200', 'lib', 'libcollie_b200.so')

HEAD = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'lts__t_bytes.sum', 'smsp__inst_executed.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_tensor.sum']


def run(cmd):
    return subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True).stdout


def main():
    rep, kern = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
    raw = list(csv.reader(run(['ncu', '-i', rep, '--page', 'raw', '--csv']).splitlines()))
    hdr, units = raw[0], raw[1]
    ki = hdr.index('Kernel Name')
    for r in raw[2:]:
        print('==', r[ki][:100])
        for h in HEAD + [x for x in hdr if 'issue_stalled' in x and x.endswith('per_issue_active.ratio')]:
            if h in hdr:
                v = r[hdr.index(h)]
                try:
                    if float(v) == 0:
                        continue
                except ValueError:
                    pass
                print(f'   {h:88s} {v[:24]:>24} {units[hdr.index(h)]}')
    # per-line
    tmp = tempfile.mkdtemp()
    subprocess.run(['cuobjdump', '-xelf', 'all', LIB], cwd=tmp, stdout=subprocess.DEVNULL)
    off2line = {}
    name = None
    for f in os.listdir(tmp):
        txt = run(['nvdisasm', '-g', '-c', os.path.join(tmp, f)])
        if kern not in txt:
            continue
        on, cur = False, None
        for ln in txt.splitlines():
            m = re.match(r'\s*\.text\.(\S+):', ln) or re.match(r'\s*\.section\s+\.text\.(\S+),', ln)
            if m:
                on = kern in m.group(1)
                if on:
                    name = m.group(1)
                continue
            if not on:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', ln)
            if m:
                off2line.setdefault(int(m.group(1), 16), cur)
        if name:
            break
    demangled = re.match(r'[A-Za-z_0-9]*?([a-z_0-9]+?)I', kern)
    kregex = sys.argv[4] if len(sys.argv) > 4 else (demangled.group(1) if demangled else kern)
    src = list(csv.reader(run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-name',
                               'regex:' + kregex]).splitlines()))
    hi = next(i for i, r in enumerate(src) if 'Address' in r)
    h = src[hi]
    ai, ii, wi = h.index('Address'), h.index('Instructions Executed'), \
        h.index('Warp Stall Sampling (All Samples)')
    base, by, st = None, collections.Counter(), collections.Counter()
    for r in src[hi + 1:]:
        if len(r) <= ii:
            if by:
                break
            continue
        try:
            a = int(r[ai], 16) if r[ai].startswith('0x') else int(r[ai])
        except ValueError:
            continue
        base = a if base is None else base
        k = off2line.get(a - base)
        by[k] += int(r[ii] or 0)
        st[k] += int(r[wi] or 0)
    tot, tw = sum(by.values()) or 1, sum(st.values()) or 1
    files = {}
    print(f'-- {name}: {tot} warp instructions in the first profiled launch')
    for k, n in sorted(by.items(), key=lambda kv: -(kv[1] / tot + st[kv[0]] / tw))[:top]:
        f, l = k if k else ('?', 0)
        if f not in files:
            path = os.path.join(ROOT, 'collie_b200', 'csrc', f)
            files[f] = open(path).read().split('\n') if os.path.exists(path) else []
        text = files[f][l - 1].strip()[:80] if 0 < l <= len(files[f]) else ''
        print(f'{100 * n / tot:5.1f}% inst {100 * st[k] / tw:5.1f}% stall  {f}:{l}: {text}')


if __name__ == '__main__':
    main()
