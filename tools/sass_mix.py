"""Histogram of the innermost hot loops of every force kernel in a .so/.cubin (SASS)."""
import collections, re, subprocess, sys
path = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    name = f.split('\n')[0].strip()
    if 'force_kernel' not in name: continue
    dem = subprocess.run(["cu++filt", name], capture_output=True, text=True).stdout.strip()
    ins = []
    for l in f.split('\n'):
        m = re.search(r'/\*([0-9a-f]{4,5})\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)(.*?);', l)
        if m: ins.append((int(m.group(1), 16), m.group(2), m.group(3)))
    loops = []
    for addr, op, rest in ins:
        if op.startswith('BRA'):
            t = re.search(r'0x([0-9a-f]+)', rest)
            if t and int(t.group(1), 16) < addr:
                loops.append((int(t.group(1), 16), addr))
    for (tgt, addr) in loops:
        if any(t2 >= tgt and a2 <= addr and (t2, a2) != (tgt, addr) for (t2, a2) in loops):
            continue  # not innermost
        body = [i for i in ins if tgt <= i[0] <= addr]
        c = collections.Counter(o.split('.')[0] for _, o, _ in body)
        if c.get('DFMA', 0) + c.get('FFMA', 0) + c.get('FFMA2', 0) < 6: continue
        best = (len(body), c)
        if best:
            c = best[1]
            fp64 = c.get('DFMA', 0) + c.get('DMUL', 0) + c.get('DADD', 0)
            n_int = c.get('MUFU', 0)
            other = best[0] - fp64
            tag = re.search(r'<(.*?)>', dem).group(1) if '<' in dem else dem
            if fp64:
                print(f"{tag:34s} loop={best[0]:4d} pairs={n_int:3d} fp64/pair={fp64/n_int:.2f} other/pair={other/n_int:.2f} "
                      f"model cycles/pair={(2*fp64+other)/n_int:.2f}  ", {k: v for k, v in c.items() if k not in ('DFMA','DMUL','DADD')})
            else:
                print(f"{tag:34s} loop={best[0]:4d} pairs={n_int:3d} instr/pair={best[0]/n_int:.2f} ", dict(c))
