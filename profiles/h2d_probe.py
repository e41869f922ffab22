import torch, time
dev=torch.device('cuda',0)
h=torch.empty(38_400_000//4, dtype=torch.float32).pin_memory()
d=torch.empty_like(h, device=dev)
def bw(nstreams, reps=20):
    streams=[torch.cuda.Stream() for _ in range(nstreams)]
    chunks=h.chunk(nstreams); dch=d.chunk(nstreams)
    torch.cuda.synchronize()
    a=torch.cuda.Event(enable_timing=True); b=torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        for s,hc,dc in zip(streams,chunks,dch):
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s): dc.copy_(hc, non_blocking=True)
        for s in streams: torch.cuda.current_stream().wait_stream(s)
    b.record(); torch.cuda.synchronize()
    return h.numel()*4*reps/(a.elapsed_time(b)*1e-3)/1e9
for n in (1,2,4): print(n,'streams', round(bw(n),1),'GB/s')
# D2H concurrently?
import subprocess
print(subprocess.run(['nvidia-smi','--query-gpu=pcie.link.gen.current,pcie.link.width.current,pcie.link.gen.max','--format=csv'],capture_output=True,text=True).stdout)
