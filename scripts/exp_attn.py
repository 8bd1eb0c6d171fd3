import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from petals_b200.ops import attention as A
dev="cuda"; dt=torch.float16; nh=nkv=32; D=128; s_max=2048
inv=A.rope_inv_freq(D,10000.0,device=dev)
for past in (128, 512, 1500):
    for splits in (4, 16):
        L=32
        qkv=torch.randn(1,(nh+2*nkv)*D,device=dev).to(dt); out=torch.empty(1,nh*D,device=dev,dtype=dt)
        caches=[(torch.randn(nkv,s_max,D,device=dev).to(dt),torch.randn(nkv,s_max,D,device=dev).to(dt)) for _ in range(L)]
        ptrs=[torch.tensor([[k.data_ptr(),v.data_ptr()]],dtype=torch.int64,device=dev) for k,v in caches]
        pos=torch.tensor([past],dtype=torch.int32,device=dev)
        ws=A.DecodeAttnWorkspace(dev,1,nh,nkv,D,splits)
        def run():
            for p in ptrs: A.attn_decode(qkv,out,p,pos,inv,ws,nh,nkv,D,s_max,pdl=True)
        run(); torch.cuda.synchronize()
        g=torch.cuda.CUDAGraph(); s=torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            with torch.cuda.graph(g,stream=s): run()
        torch.cuda.current_stream().wait_stream(s); g.replay(); torch.cuda.synchronize()
        e0,e1=torch.cuda.Event(True),torch.cuda.Event(True); e0.record()
        for _ in range(20): g.replay()
        e1.record(); torch.cuda.synchronize()
        print(f"past={past} splits={splits}: {e0.elapsed_time(e1)*1e3/(20*L):.2f} us/layer", flush=True)
