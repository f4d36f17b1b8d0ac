"""Instruction / stall share per phase of k_fused_step_v2 from an ncu report (uses ncu_lines)."""
import collections, contextlib, importlib, io, re, sys
sys.path.insert(0, 'profiles')
rep, kern, per = sys.argv[1], sys.argv[2], float(sys.argv[3])
sys.argv = ['x', rep, kern, '600', 'k_fused_step_v2']
m = importlib.import_module('ncu_lines')
buf = io.StringIO()
with contextlib.redirect_stdout(buf):
    m.main()
lines = [l for l in buf.getvalue().split('\n') if '% inst' in l]
src = open('collie_b200/csrc/train_step_v2.cu').read().split('\n')
def find(s):
    for i, l in enumerate(src):
        if s in l:
            return i + 1
marks = [('phaseA', find('phase A: the id table')), ('issue-lambda', find('auto issue = [&]')),
         ('prologue', find('for (int k = 0; k < kStages - 1; ++k) issue(k);')),
         ('finish-lambda', find('auto finish_user')), ('loop-top/user', find('for (int le = 0; le < n_e; ++le) {')),
         ('wait', find('issue(le + kStages - 1);')), ('dots', find('all N + 1 dot products at once')),
         ('select', find('if (j0 == 0) {')), ('loss', find('loss value and d loss / d score')),
         ('backward', find('backward: lane owns float4')), ('tail', find('the stage may be refilled only after'))]
marks = [(k, v) for k, v in marks if v]
acc, st = collections.Counter(), collections.Counter()
for l in lines:
    mm = re.match(r'\s*([\d.]+)% inst\s+([\d.]+)% stall\s+(\S+):(\d+):', l)
    if not mm:
        continue
    pi, ps, f, ln = float(mm.group(1)), float(mm.group(2)), mm.group(3), int(mm.group(4))
    key = f
    if f == 'train_step_v2.cu':
        key = 'v2:helpers'
        for k, a in marks:
            if ln >= a:
                key = 'v2:' + k
    acc[key] += pi
    st[key] += ps
for k, v in acc.most_common():
    print(f'{k:28s} inst {v:5.1f}%  stall {st[k]:5.1f}%   ~{v * per / 100:.0f} instr/interaction')
