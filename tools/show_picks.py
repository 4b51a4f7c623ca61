import sys; sys.path.insert(0,'/root/repo')
import torch, cpmcu_b200 as B
from cpmcu_b200 import capi as L
torch.zeros(1,device='cuda')
for cfg in list(range(1005,1013))+list(range(2001,2081)):
    print(cfg, L.config_name(cfg), [L.lib().b200_w4a16_max_clusters(cfg, m, 128, 0) for m in (1,16,32,64)])
for (k,n) in [(4096,4608),(4096,4096),(4096,32768),(16384,4096)]:
    print(k,n,[ (m, L.pick_config(m,n,k), L.config_name(L.pick_config(m,n,k))) for m in (1,16,32,64,128)])
