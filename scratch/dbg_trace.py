import sys; sys.path.insert(0,'.')
import numpy as np
import ipu_path_trace_b200 as pt
from oracle import oracle as O
orc=O.Oracle(); model=pt.synthetic_model()
for (W,H,kw) in [(320,240,{}),(64,64,dict(aa_noise_type=2))]:
    s=O.Settings(width=W,height=H,seed=77,env_map_rotation_degrees=33.0,**kw)
    t=pt.PathTracer(W,H,**kw); t.init_nif_weights(model); t.init_render_settings(seed=77,env_map_rotation_degrees=33.0)
    items=O.make_worklist(W,H); t.setup(items)
    for sample in (0,5):
        esc,uv,thr,length,cam=t.trace_wavefront(sample)
        e2,uv2,t2,l2,cam2=orc.trace_philox(items,sample,s)
        bad=np.where((esc!=e2)|(length!=l2))[0]
        print(W,H,sample,'flips',len(bad))
        for i in bad[:12]: print('  item',i,'gpu',esc[i],length[i],thr[i],'cpu',e2[i],l2[i],t2[i])
        m=(esc==e2)&(length==l2)&(esc==1)
        d=np.abs(uv[m]-uv2[m]); j=np.argsort(-d.max(1))[:8]
        print('  uv worst',d.max(), [ (uv[m][k].tolist(),uv2[m][k].tolist()) for k in j[:4]])
        dt=np.abs(thr[m]-t2[m])/np.maximum(np.abs(t2[m]),1e-7); print('  thr worst rel',dt.max())
        print('  lengths hist gpu',np.bincount(length,minlength=11),'cpu',np.bincount(l2,minlength=11))
