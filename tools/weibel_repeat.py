import sys, numpy as np
sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
from vpic_b200 import b200_backend, load_b200
from weibel import run_weibel, gold_lines
L=load_b200(); L.vpic_b200_init(0,0,1); L.vpic_b200_set_coherence(1)
gold=np.array([[float(x) for x in ln.split()] for ln in gold_lines()])
for t in range(4):
    _,text=run_weibel(b200_backend())
    got=np.array([[float(x) for x in ln.split()] for ln in text.splitlines()])
    pe=np.abs(got[:,7:9]-gold[:,7:9])/gold[:,7:9]
    fe=[np.abs(got[:51,c]-gold[:51,c]).max()/max(gold[:51,c].max(),1e-30) for c in range(1,7)]
    es=np.abs(got[:,1:4].sum(1)-gold[:,1:4].sum(1))/np.maximum(gold[:,1:4].sum(1),1e-30)
    bs=np.abs(got[:,4:7].sum(1)-gold[:,4:7].sum(1))/np.maximum(gold[:,4:7].sum(1),1e-30)
    print("run",t,"particle rel max all",pe.max(0),"first50",pe[:51].max(0),"field first50",np.round(fe,7),"Esum",es.max(),"Bsum",bs.max())
