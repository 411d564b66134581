import sys, numpy as np
sys.path.insert(0,'.')
import jplspatial_b200 as J
from oracle import refpy
from tests import util
from tests.test_gpu_render import make_scene, run_reference, run_gpu
ref=refpy.Ref(); ctx=J.Context(0)
N,C,L,blocks,frames=96,1,12,6,512
sc=make_scene(N,C,L,blocks,frames,util.SEED+N)
rper,rbus=run_reference(ref,sc,N,C,L,blocks,frames)
gper,gbus,_=run_gpu(ctx,sc,N,C,L,blocks,frames,J.MODE_PARITY)
bad=(gper.view(np.uint32)!=rper.view(np.uint32))
print("bad per block:",bad.reshape(blocks,-1).sum(1))
bs=bad.any(axis=(2,3))
for b in range(blocks):
    idx=np.nonzero(bs[b])[0]
    print("block",b,"bad sources",idx[:40], len(idx))
b,s=np.argwhere(bs)[0]
fr=np.nonzero(bad[b,s].any(1))[0]
print("first bad block/source",b,s,"frames",fr[:10],len(fr),"channels bad at first frame",np.nonzero(bad[b,s,fr[0]])[0])
print("delay samples cur/tgt", sc['delay'][b,s]*48000, sc['delay'][b+1,s]*48000)
print("gpu",gper[b,s,fr[0]], "\nref",rper[b,s,fr[0]])
