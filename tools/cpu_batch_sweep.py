"""CPU batch sweep (no GPU): batches of 2-4 files, some damaged / CAVLC / cut / bit-flipped, through the slice-parallel compress bodies; every
file's payloads must equal what the sequential chain gives for that file alone (estimator state is per file, recode.cpp:1314).
python tools/cpu_batch_sweep.py <seed> <batches>"""
import os, sys, random, subprocess, tempfile
ROOT=os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, os.path.join(ROOT,'tests'))
from conftest import damage_slice, as_cavlc, drop_slice_nal
EMUL=os.path.join(ROOT,'build/avr_emul'); GEN=os.path.join(ROOT,'avrecode-ms_b200/bin/avrgen')
rnd=random.Random(int(sys.argv[1])); n=int(sys.argv[2]); bad=0
names=["ip_basic","multislice_multiref","dct8x8","dense_full_blocks","skippy_idr","b_slices","b_direct_multiref"]
td=tempfile.mkdtemp()
def variant(i):
    name=rnd.choice(names); d=open(os.path.join(ROOT,'tests/golden',name+'.h264'),'rb').read()
    op=rnd.choice(['plain','plain','damage','cavlc','drop','flip'])
    if op=='damage': d=damage_slice(d, rnd.randrange(2), rnd.randrange(100))
    elif op=='cavlc': d=as_cavlc(d)
    elif op=='drop': d=drop_slice_nal(d, rnd.randrange(2))
    elif op=='flip':
        d=bytearray(d)
        for _ in range(3): d[rnd.randrange(len(d))]^=1<<rnd.randrange(8)
        d=bytes(d)
    p=os.path.join(td,f'f{i}.h264'); open(p,'wb').write(d); return p
for it in range(n):
    fs=[variant(i) for i in range(rnd.randrange(2,5))]
    for k in range(len(fs)):
        order=[fs[k]]+[f for j,f in enumerate(fs) if j!=k]
        pi=subprocess.run([EMUL,'pipeline']+order,capture_output=True)
        ch=subprocess.run([EMUL,'chain',fs[k]],capture_output=True)
        if ch.returncode!=0: continue        # file refused as a whole (literal) or chain flags it: not part of a device batch
        if pi.returncode!=0 or pi.stdout!=ch.stdout:
            bad+=1; print('BAD',it,k,pi.returncode,pi.stderr[-200:]); 
            for j,f in enumerate(order): os.system(f'cp {f} /tmp/batch_{sys.argv[1]}_{it}_{k}_{j}.h264')
            break
print('done',n,'bad',bad)
