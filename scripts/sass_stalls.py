#!/usr/bin/env python
"""Static schedule of the sweep kernel: decodes the stall counts of the SASS control words (cuobjdump -sass) and prints,
per library given on the command line, instructions and stall cycles per unrolled cell body of sweep_kernel<3,TWDD>."""
import re, subprocess, sys
def load(lib):
    out=subprocess.run(f"cuobjdump -sass {lib}",shell=True,capture_output=True,text=True).stdout
    lines=out.split('\n'); ins=[]; on=False
    for i,l in enumerate(lines):
        if 'Function :' in l: on = ('sweep_kernelILi3ELi1ELb0' in l or ('sweep_kernelILi3ELi1E' in l and 'ELb' not in l))
        if not on: continue
        m=re.match(r'\s+/\*([0-9a-f]{4})\*/\s+(.*?);\s+/\* (0x[0-9a-f]+) \*/',l)
        if m:
            hi=int(re.search(r'/\* (0x[0-9a-f]+) \*/',lines[i+1]).group(1),16)
            ins.append((m.group(2).strip(),hi))
    return ins
def dec(hi): return (hi>>41)&0xf, (hi>>45)&1, (hi>>46)&7, (hi>>49)&7, (hi>>52)&0x3f, (hi>>58)&0xf
for lib in sys.argv[1:]:
    ins=load(lib)
    mufu=[i for i,(t,h) in enumerate(ins) if t.startswith('MUFU.RCP64H')]
    # cell bodies: every 4th MUFU marks the psi division; take spans between consecutive psi-MUFUs
    psi=mufu[3::4]
    spans=[(psi[k],psi[k+1]) for k in range(len(psi)-1) if psi[k+1]-psi[k]<130]
    tot=[]
    for a,b in spans:
        st=sum(dec(h)[0] for t,h in ins[a:b]); n=b-a
        waits=sum(1 for t,h in ins[a:b] if dec(h)[4])
        reuse=sum(bin(dec(h)[5]).count('1') for t,h in ins[a:b])
        tot.append((n,st,waits,reuse))
    print(lib, 'cells',len(tot), 'instr/cell %.1f'%(sum(t[0] for t in tot)/len(tot)), 'stall-sum/cell %.1f'%(sum(t[1] for t in tot)/len(tot)), 'waitmask instrs %.1f'%(sum(t[2] for t in tot)/len(tot)), 'reuse flags %.1f'%(sum(t[3] for t in tot)/len(tot)))
