"""Counts the Blackwell-specific SASS mnemonics inside hist_scan_kernel and prints the TMA gather loop verbatim.
Usage: python profiles/sass_evidence.py > profiles/r2_hist_sass.txt   (runs cuobjdump on the in-tree libemrf.so)"""
import collections
import os
import re
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
txt = subprocess.run(["cuobjdump", "-sass", os.path.join(ROOT, "euromillioner_b200", "libemrf.so")], capture_output=True, text=True).stdout
funcs = re.split(r'\n\s*Function : ', txt)
out = ["# SASS evidence for the TMA row staging of hist_scan_kernel (round 2)", "",
       "`cuobjdump -sass euromillioner_b200/libemrf.so`, instruction mnemonics counted per function (sm_100a).",
       "`UTMALDG.2D.GATHER4` = cp.async.bulk.tensor.2d ... tile::gather4 (TMA), `SYNCS.*` = mbarrier (init / arrive.expect_tx /",
       "try_wait), `ATOMS` = shared-memory atomic add (the TMA loop's adds are `ATOMS.ADD` coming from red.shared.add.u32; the rest belong",
       "to the dense-node, generic and LDG/STS paths), `UBLKCP` = cp.async.bulk (none: rows are gathered four at a time through the",
       "tensor map).", "",
       "| function | instructions | UTMALDG | SYNCS | ATOMS | LDGSTS | UBLKCP | R2UR | ELECT |", "|---|---:|---:|---:|---:|---:|---:|---:|---:|"]
tot = collections.Counter()
n_inst = 0
for f in funcs[1:]:
    name = f.split('\n', 1)[0].strip()
    if 'hist_scan_kernel' not in name:
        continue
    ins = re.findall(r'/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)', f)
    c = collections.Counter(i.split('.')[0] for i in ins)
    m = re.search(r'hist_scan_kernelILi(\d)E([jt])E', name)
    label = f"hist_scan_kernel<{m.group(1)}, {'uint32_t' if m.group(2) == 'j' else 'uint16_t'}>" if m else name[:60]
    out.append(f"| `{label}` | {len(ins)} | {c['UTMALDG']} | {c['SYNCS']} | {c['ATOMS']} | {c['LDGSTS']} | {c['UBLKCP']} | {c['R2UR']} | {c['ELECT']} |")
    n_inst += 1
    for k in ('UTMALDG', 'SYNCS', 'ATOMS', 'LDGSTS', 'UBLKCP'):
        tot[k] += c[k]
out += ["", f"Totals over the {n_inst} instantiations: UTMALDG {tot['UTMALDG']}, SYNCS {tot['SYNCS']}, ATOMS {tot['ATOMS']}, LDGSTS {tot['LDGSTS']}, "
        f"UBLKCP {tot['UBLKCP']}.",
        "Template parameter = 16-byte chunks per TreePoint row: 2 / 4 / 8 (rows of 32 / 64 / 128 bytes) carry the TMA loop (and, behind",
        "EMRF_NO_TMA=1, the LDG/STS staging it replaced); 0 is the generic byte-gather path for more than 128 features.", ""]
for f in funcs[1:]:
    name = f.split('\n', 1)[0].strip()
    if 'hist_scan_kernelILi4EjE' in name:
        lines = [l for l in f.split('\n') if re.search(r'/\*[0-9a-f]{4}\*/', l)]
        strip = lambda l: re.sub(r'\s*/\* 0x[0-9a-f]+ \*/\s*$', '', l).rstrip()
        idx = [i for i, l in enumerate(lines) if 'UTMALDG' in l]
        if idx:
            a, b = max(0, idx[0] - 24), min(len(lines), idx[0] + 3)
            out += ["## The gather loop of `hist_scan_kernel<4, uint32_t>` (rows of 64 bytes), verbatim", "", "```"] + [strip(l) for l in lines[a:b]] + ["```"]
        idx = [i for i, l in enumerate(lines) if 'SYNCS.PHASECHK' in l]
        if idx:
            out += ["", "and the wait on the stage's mbarrier with the first feature reads and adds behind it:", "", "```"]
            out += [strip(l) for l in lines[idx[0] - 1: idx[0] + 16]] + ["```"]
        break
print('\n'.join(out))
