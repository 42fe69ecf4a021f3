import sys; sys.path.insert(0,'.')
import torch
from alphatriangle_b200 import _lib
lib=_lib.load()
src=torch.zeros(1<<16,dtype=torch.uint8,device='cuda'); out=torch.zeros(148,dtype=torch.float32,device='cuda')
print("per four MMAs: 1 = try_wait, 2 = commit, 4 = fence::after, 8 = epilogue warps poll a barrier; cycles per MMA of issuer 0")
for issuers in (2,):
    for flags in (3, 11):
        for epi in (0, 8, 16):
            for _ in range(2):
                assert lib.at_probe_mma_rate(128,4096,issuers,32+flags,epi,src.data_ptr(),148,out.data_ptr(),None)==0
                torch.cuda.synchronize()
            o=out.cpu(); print(f"issuers={issuers} flags={flags} epi_warps={epi}: {o.mean():.1f} {o.min():.1f} {o.max():.1f}")
