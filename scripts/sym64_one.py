"""One symmetric fp64 evaluation at the given n (ncu target)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn, random_q
n = int(sys.argv[1]) if len(sys.argv) > 1 else 17
bn, lat, obs, x = random_binary_bn(n, seed=0)
q = random_q(1 << n, seed=0).cuda()
t = ops.SteinTables.from_network(bn, x, lat, precision="fp64")
for _ in range(2):
    u, part = t.ksd_sym64(q)
torch.cuda.synchronize()
print(part.item())
