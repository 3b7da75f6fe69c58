"""Summarise `ncu --page source --print-source cuda,sass --csv` of a warp-specialised kernel per ROLE: every SASS instruction
is assigned to the role whose source-line range (in the kernel's .cu file) the nearest instruction at a lower or equal address
maps to (inlined helpers from other headers inherit the role of their surroundings).
usage: ncu_roles.py source_cs.csv file.cu 'A:244-329,P:330-360,...' [trials]"""
import csv, sys, collections, re
path, cu, spec = sys.argv[1], sys.argv[2], sys.argv[3]
trials = float(sys.argv[4]) if len(sys.argv) > 4 else 65536.0
roles = []
for part in spec.split(','):
    n, r = part.split(':'); lo, hi = r.split('-'); roles.append((n, int(lo), int(hi)))
rows = list(csv.reader(open(path)))
cur_file = None; hdr = None; ins = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path': cur_file = r[1]; continue
    if len(r) == 2: continue
    if r and r[0] == 'Line No': hdr = r; continue
    if hdr is None or len(r) < 10: continue
    if r[0] != '': cur_line = int(r[0]); continue
    if not r[2].startswith('0x'): continue
    d = dict(zip(hdr[2:], r[2:]))
    ins.append(dict(addr=int(r[2], 16), file=cur_file, line=cur_line, sass=r[3].strip(), n=int(r[7] or 0), samples=int(r[6] or 0),
                    stalls={k: int(v) for k, v in zip(hdr, r) if k.startswith('stall_') and 'Not Issued' not in k and v not in ('', '-', '0')}))
def role_of(i):
    if i['file'].endswith(cu):
        for n, lo, hi in roles:
            if lo <= i['line'] <= hi: return n
    return None
# an instruction inside an inlined helper is listed once per frame of its inline stack: keep one record per address, with the
# role of whichever frame lies in the kernel's file and the (file, line) of the innermost frame seen first
uniq = {}
for i in ins:
    u = uniq.setdefault(i['addr'], i)
    r = role_of(i)
    if r is not None and u.get('drole') is None: u['drole'] = r
ins = sorted(uniq.values(), key=lambda x: x['addr'])
last = 'prologue'
for i in ins:
    if i.get('drole') is not None: last = i['drole']
    i['role'] = last
# mbarrier polling (mibmi_ptx.cuh mbar_test / mbar_wait*): the instrumented pass that counts instructions runs slower than the
# real kernel, so its poll counts are inflated; they are reported separately
def is_poll(i): return i['file'].endswith('mibmi_ptx.cuh') and 25 <= i['line'] <= 76
tot_n = sum(i['n'] for i in ins); tot_s = sum(i['samples'] for i in ins)
print(f'non-poll warp-instructions per trial: {sum(i["n"] for i in ins if not is_poll(i)) / trials:.0f}; poll (instrumented pass): {sum(i["n"] for i in ins if is_poll(i)) / trials:.0f}; poll samples {100 * sum(i["samples"] for i in ins if is_poll(i)) / tot_s:.1f}%')
for i in ins: i['poll'] = is_poll(i)
print(f'total warp-instructions {tot_n} = {tot_n / trials:.0f} per trial; samples {tot_s}')
agg = collections.defaultdict(lambda: dict(n=0, s=0, st=collections.Counter(), op=collections.Counter()))
for i in ins:
    a = agg[i['role']]; a['n'] += i['n']; a['s'] += i['samples']; a['st'].update(i['stalls'] if not i['poll'] else {}); a['np'] = a.get('np', 0) + (0 if i['poll'] else i['n']); a['ps'] = a.get('ps', 0) + (i['samples'] if i['poll'] else 0)
    if i['poll']: continue
    a['op'][i['sass'].split()[0] if not i['sass'].startswith('@') else i['sass'].split()[1]] += i['n']
for rname, a in sorted(agg.items(), key=lambda kv: -kv[1]['n']):
    st = ', '.join(f'{k[6:]} {100 * v / max(a["s"] - a["ps"], 1):.0f}%' for k, v in a['st'].most_common(7))
    ops = ', '.join(f'{k} {v / trials:.0f}' for k, v in a['op'].most_common(12))
    print(f'{rname:9s} {a["np"] / trials:7.0f} non-poll instr/trial, poll samples {100 * a["ps"] / max(a["s"], 1):.0f}% of its samples;  samples {100 * a["s"] / tot_s:4.1f}%  non-poll stalls: {st}\n          ops/trial: {ops}')
if '--lines' in sys.argv:
    byl = collections.defaultdict(lambda: [0, 0])
    for i in ins:
        k = (i['role'], i['file'].split('/')[-1], i['line']); byl[k][0] += i['n']; byl[k][1] += i['samples']
    for k, v in sorted(byl.items(), key=lambda kv: -kv[1][1])[:40]:
        print(f'{k[0]:9s} {k[1]}:{k[2]:<5d} instr/trial {v[0] / trials:7.1f}  samples {100 * v[1] / tot_s:5.2f}%')
