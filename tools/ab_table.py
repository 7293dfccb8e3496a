"""Developer helper: tabulate a tools/gpu_call*.sh A/B log (Msamples/s per configuration, relative to the first)."""
import re, sys
cfg = None; rows = {}; diag = {}
for line in open(sys.argv[1]):
    if line.startswith('=='):
        cfg = re.sub(r'MCPT_LIB=\S*libmcpt(_\w+)?\.so', lambda m: 'LIB' + (m.group(1) or ''), line[3:].strip()); rows[cfg] = []; diag[cfg] = []; continue
    m = re.search(r'([\d.]+) Msamples/s', line)
    if m: rows[cfg].append(float(m.group(1)))
    if line.startswith('[diag]') or 'nodes/ray' in line: diag[cfg].append(line.strip())
base = next(iter(rows.values()))
for c, v in rows.items():
    print('%-55s' % c, ' '.join('%7.1f' % x for x in v), ' | ', ' '.join('%+5.1f%%' % ((x / b - 1) * 100) for x, b in zip(v, base)))
for c in diag:
    if diag[c]:
        print('==', c)
        for l in diag[c]: print('  ', l)
