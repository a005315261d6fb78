"""Compile ONE K1S instantiation without a GPU and count what is in its column loop (register moves per row, DPX count, loop
bytes): scripts/sass_probe.sh <name> [-DPR=32 -DPG=32 -DPM=false -DPS=1 -DK1S_...] builds /tmp/probe/<name>.sass and calls this.
  python scripts/sass_loop_moves.py <name> [min DPX in the loop]"""
import re,sys
from collections import Counter
name=sys.argv[1]; minv=int(sys.argv[2]) if len(sys.argv)>2 else 90
L=[l.strip() for l in open(f'/tmp/probe/{name}.sass')]
ins=[]
for l in L:
    m=re.match(r'/\*([0-9a-f]+)\*/\s+(.*);',l)
    if m:
        a=int(m.group(1),16)
        if ins and a==0: break
        ins.append((a,m.group(2).strip()))
lo=hi=None
for a,i in ins:
    m=re.search(r'BRA(?:\.\w+)*\s+(?:\S+,\s*)?0x([0-9a-f]+)',i)
    if m and int(m.group(1),16)<a:
        t=int(m.group(1),16); n=sum(1 for b,j in ins if t<=b<=a and 'VIADDMNMX' in j)
        if n>=minv and (lo is None or a-t<hi-lo): lo,hi=t,a
body=[j for a,j in ins if lo<=a<=hi]
dpx=sum(1 for j in body if 'VIADDMNMX' in j)
mv=sum(1 for j in body if re.match(r'(@\S+ )?(IMAD\.MOV\.U32|MOV) R\d+, (RZ, RZ, )?R\d+',j))
imm=sum(1 for j in body if re.match(r'(@\S+ )?(IMAD\.MOV\.U32|MOV|HFMA2) R\d+, (-?RZ, RZ, )',j) and not re.match(r'(@\S+ )?(IMAD\.MOV\.U32|MOV) R\d+, (RZ, RZ, )?R\d+',j))
print(f'{name}: loop {hi-lo+16} B, {len(body)} instrs, VIADDMNMX {dpx}, reg-reg moves {mv} ({mv/dpx*3:.2f} per row), imm moves {imm}')
