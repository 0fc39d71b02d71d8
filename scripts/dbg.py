import sys, time; sys.path.insert(0,'.')
from nori_b200.gpu import Engine
from nori_b200.scene import FlatScene
s=FlatScene.load('tests/golden/scenes/cbox_mis.npz')
e=Engine(0); print(e.load_scene(s), flush=True)
for spp in [16, 40, 100]:
    t=time.time(); _,st=e.render(sample_first=0,sample_last=spp,download=False); print(spp, time.time()-t, st, flush=True)
