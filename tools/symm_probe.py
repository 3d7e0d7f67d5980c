"""Does the symmetric-memory push path work on this box?  (torchrun, short timeout)"""
import faulthandler, os, sys
faulthandler.dump_traceback_later(50, exit=True)
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
def log(*a):
    if rank == 0: print(*a, flush=True)
import ntxent_b200
from ntxent_b200.distributed import RowPartitionedNTXent
M, D, tau = 4096, 128, 0.1
rows = M // world
z = torch.from_numpy(np.random.default_rng(0).standard_normal((M, D), dtype=np.float32)).to(torch.bfloat16)
zl = z[rank * rows:(rank + 1) * rows].to(dev)
ref = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev)
l0, d0 = ref.forward_backward(zl); torch.cuda.synchronize(); log("nccl ok", l0.item())
eng = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev, comm="symm")
log("symm buffers ok; multicast:", eng.symm.hdl.has_multicast_support if hasattr(eng.symm.hdl, "has_multicast_support") else "?")
for i in range(3):
    l1, d1 = eng.forward_backward(zl); torch.cuda.synchronize()
    log("symm step", i, l1.item(), "dz maxdiff:", (d0.float() - d1.float()).abs().max().item(), "of", d0.float().abs().max().item())
eng_g = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev, comm="symm", use_cuda_graph=True)
for i in range(5):
    l2, d2 = eng_g.forward_backward(zl); torch.cuda.synchronize()
    log("symm graph step", i, l2.item(), "dz equal to eager symm:", torch.equal(d1, d2))
eng_g.release()
dist.barrier(); torch.cuda.synchronize()
log("done")
dist.destroy_process_group()
