"""2-rank check: CUDA-graph capture of the step including the NCCL gradient exchange. Run under torchrun with a timeout."""
import faulthandler, os, sys
faulthandler.dump_traceback_later(45, exit=True)
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np, torch, torch.distributed as dist
from deeplearning4j_b200 import network, parallel, zoo
rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import test_gpu_network as T
m = T._tiny_resnet() if len(sys.argv) < 2 else zoo.resnet50()
B, K = (8, 10) if len(sys.argv) < 2 else (64, 1000)
net = network.Network(m, B, device=local)
pw = parallel.ParallelWrapper(net)
x, lab = T._data(m, B, K, seed=rank + 1)
print(rank, "eager", flush=True)
for _ in range(2): pw.fit(x, lab)
net.sync(); dist.barrier(); torch.cuda.synchronize()
print(rank, "capture", flush=True)
net.capture(hook=pw._hook, include_input=True)
print(rank, "replay", flush=True)
for _ in range(3): s = net.fit_graph(x, lab)
net.sync()
print(rank, "score", float(s.cpu()[0]), "param checksum", float(net.params.double().sum().cpu()), flush=True)
dist.barrier(); torch.cuda.synchronize()
sys.stdout.flush(); os._exit(0)  # destroy_process_group() blocks while a captured graph references the communicator
