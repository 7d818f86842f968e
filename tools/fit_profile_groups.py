"""Group the ncu source-page samples of fit_kernel by code region (worker pixel passes, barriers, master's scalar code)."""
import csv, sys, collections
sys.path.insert(0, __file__.rsplit('/', 1)[0])
import ncu_lines as nl
src, sass, func = sys.argv[1:4]
evals = float(sys.argv[4]) if len(sys.argv) > 4 else 20000 * 30
import re
bounds = {}
for i, ln in enumerate(open('/root/repo/shopt_b200/csrc/lbfgs_fit.cu'), 1):
    for name in ('scalar_values(', 'scalar_partials(', 'e_to_g(', 'exp_profile(double', 'lds_f64(', 'warp_reduce4(', 'void eval_pixels(',
                 'FG eval_fg(', 'void worker_loop(', 'double eps_of(', 'struct LineSearch', 'double dot3(', 'struct FitArgs'):
        if name in ln and ('__device__' in ln or ln.startswith('struct')):
            bounds.setdefault(name, i)
order = sorted(bounds.items(), key=lambda kv: kv[1])
labels = {'scalar_values(': 'M scalar_values', 'scalar_partials(': 'M scalar_partials', 'e_to_g(': 'M e_to_g', 'exp_profile(double': 'W exp_profile',
          'lds_f64(': 'W misc', 'warp_reduce4(': 'W warp_reduce', 'void eval_pixels(': 'W eval_pixels', 'FG eval_fg(': 'M eval_fg',
          'void worker_loop(': 'W worker_loop (barrier wait)', 'double eps_of(': 'M linesearch', 'struct LineSearch': 'M linesearch',
          'double dot3(': 'M lbfgs', 'struct FitArgs': 'kernel body / staging'}
def grp(key):
    if key is None: return 'none'
    f, l, _ = key
    if f != 'lbfgs_fit.cu': return 'lib:' + f
    if l < order[0][1]: return 'M ieee-ops'
    lab = None
    for name, start in order:
        if l >= start: lab = labels[name]
    return lab
lines = nl.load_lines(sass, func)
rows = list(csv.reader(open(src))); hdr = rows[1]
ia, isamp, iex = hdr.index("Address"), hdr.index("# Samples"), hdr.index("Instructions Executed")
groups = {}; tot = 0; base = None
for r in rows[2:]:
    if len(r) <= iex or not r[ia].startswith('0x'): continue
    a = int(r[ia], 16)
    if base is None: base = a
    k = grp(lines.get(a - base)); s = int(r[isamp] or 0); e = int(r[iex] or 0)
    g = groups.setdefault(k, [0, 0]); g[0] += s; g[1] += e; tot += s
for k, (s, e) in sorted(groups.items(), key=lambda kv: -kv[1][0]):
    print("%6.2f%% samples %8.1f inst/eval/warp  %s" % (100 * s / tot, e / evals / (8 if k.startswith('W') else 1), k))
