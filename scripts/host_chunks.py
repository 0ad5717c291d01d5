import torch, sys
sys.path.insert(0, "/root/repo")
from mini_flash_attention_b200 import flash_attn_func_host
B,S,H,D=48,4096,24,128
q=torch.randn(B,S,H,D,dtype=torch.bfloat16).pin_memory(); k=torch.randn_like(q).pin_memory(); v=torch.randn_like(q).pin_memory()
o=torch.empty_like(q).pin_memory()
for cb,depth in [(6,3),(4,3),(3,3),(2,3),(3,4),(2,4),(1,4)]:
    for _ in range(2): flash_attn_func_host(q,k,v,False,out=o,chunk_batch=cb,depth=depth)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): flash_attn_func_host(q,k,v,False,out=o,chunk_batch=cb,depth=depth)
    e1.record(); torch.cuda.synchronize()
    print(cb, depth, round(e0.elapsed_time(e1)/5,2), "ms", flush=True)
