"""Where does CUDA-graph capture of the distributed step stall?  (torchrun, short timeouts)"""
import faulthandler, os, sys
faulthandler.dump_traceback_later(45, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
def log(*a):
    if rank == 0: print(*a, flush=True)
# 1. plain NCCL capture
x = torch.ones(1024, device=dev); out = torch.empty(1024 * world, device=dev)
s = torch.cuda.Stream(); s.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(s):
    for _ in range(3): dist.all_gather_into_tensor(out, x)
torch.cuda.current_stream().wait_stream(s); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    dist.all_gather_into_tensor(out, x)
g.replay(); torch.cuda.synchronize()
log("1 plain NCCL capture+replay ok", out.sum().item())
# 2. engine pieces
import ntxent_b200
from ntxent_b200.distributed import RowPartitionedNTXent
M, D, tau = 4096, 128, 0.1
rows = M // world
z = torch.from_numpy(np.random.default_rng(0).standard_normal((M, D), dtype=np.float32)).to(torch.bfloat16)
zl = z[rank * rows:(rank + 1) * rows].to(dev)
eng = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev)
for _ in range(3): eng.forward_backward(zl)
torch.cuda.synchronize(); log("2 eager ok")
g2 = torch.cuda.CUDAGraph()
with torch.cuda.graph(g2):
    loss, state = eng.forward(zl)
log("3 captured forward")
g2.replay(); torch.cuda.synchronize(); log("4 replayed forward", loss.item())
g3 = torch.cuda.CUDAGraph()
with torch.cuda.graph(g3):
    loss, state = eng.forward(zl)
    dz = eng.backward(state, None)
log("5 captured fwd+bwd")
g3.replay(); torch.cuda.synchronize(); log("6 replayed fwd+bwd", loss.item(), dz.float().abs().max().item())
del g, g2, g3
torch.cuda.synchronize()
import gc; gc.collect()
log('7 graphs released')
dist.destroy_process_group()
log('8 destroyed')
