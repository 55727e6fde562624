import time, numpy as np, torch, sys
sys.path.insert(0, '.')
from audio_style_transfer_b200.step import StyleStep
from audio_style_transfer_b200 import synth
for name,(cont,style,gatys) in {'cfg0':([25],list(range(10)),False),'cfg1':([29],list(range(30)),True)}.items():
    T=16384
    st=StyleStep(T,cont,style,gatys,precision='fp32')
    NL=st.n_layers
    st.set_weights(synth.make_weights(NL,0,0.0))
    x=torch.randn(1,T,device='cuda')*30
    Ec,Gn=st.embeds(x)
    st.set_targets(Ec*0.9,Gn,100.0)
    lo=torch.empty(1,4,device='cuda'); g=torch.empty(1,T,device='cuda')
    for _ in range(3): st.step(x,lo,g)
    torch.cuda.synchronize(); t0=time.time()
    for _ in range(10): st.step(x,lo,g)
    torch.cuda.synchronize(); dt=(time.time()-t0)/10
    print(name,'fp32 ms/step',dt*1e3,'launches',st.launches_per_step,'phases',st.step_timed(x,lo,g), 'ws MB', st.workspace_bytes/1e6)
    xh=x.cpu().numpy().astype(np.float64)[0]
    st.step_host(xh); t0=time.time()
    for _ in range(10): st.step_host(xh)
    print(name,'host ms/step',(time.time()-t0)/10*1e3)
    st.close()
