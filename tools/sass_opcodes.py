"""Per-kernel SASS opcode counts of the built library (profiles/r2_sass_opcodes.txt):  python tools/sass_opcodes.py > profiles/r2_sass_opcodes.txt"""
import os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = os.path.join(ROOT, 'tacotron-2_b200', 'wavenet_vocoder', 'libwavenet_b200.so')
sass = subprocess.run(['cuobjdump', '-sass', lib], capture_output=True, text=True, check=True).stdout
names = subprocess.run(['c++filt'], input='\n'.join(re.findall(r'Function : (\S+)', sass)), capture_output=True, text=True).stdout.split('\n')
ops = ['UTCHMMA', 'UTCQMMA', 'LDTM', 'STTM', 'UTCBAR', 'UBLKCP', 'UTMALDG', 'UTMASTG', 'FFMA2', 'FFMA', 'HMMA', 'MUFU.EX2', 'MUFU.RCP', 'SYNCS', 'ST.ASYNC', 'STAS']
print('# cuobjdump -sass tacotron-2_b200/wavenet_vocoder/libwavenet_b200.so (round 2, final build), instruction counts per kernel')
print('# UTCHMMA = tcgen05.mma, LDTM / STTM = tcgen05.ld / .st, UTCBAR = tcgen05.commit, UBLKCP = cp.async.bulk, STAS = st.async (DSMEM), SYNCS = mbarrier ops')
print('{:<72s}'.format('kernel') + ''.join('{:>9s}'.format(o) for o in ops))
for i, body in enumerate(re.split(r'Function : \S+', sass)[1:]):
    name = re.sub(r'\(anonymous namespace\)::', '', names[i])
    name = re.sub(r'\(.*$', '', name)
    counts = []
    for o in ops:
        pat = r'\b' + re.escape(o) + (r'\b' if o not in ('FFMA',) else r'(?!2)\b')
        counts.append(len(re.findall(pat, body)))
    print('{:<72s}'.format(name[:71]) + ''.join('{:>9d}'.format(c) for c in counts))
