"""One build of the joint and score tables at the given n with S64 kept (ncu target for K1/K2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tensornetworks_b200 import ops
from tensornetworks_b200.synthetic import random_binary_bn
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
bn, lat, obs, x = random_binary_bn(n, seed=0)
for _ in range(3):
    t = ops.SteinTables.from_network(bn, x, lat, precision="fp32", keep_scores=True)
torch.cuda.synchronize()
print(float(t.P.sum()))
