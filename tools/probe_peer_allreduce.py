"""Two or more ranks (torchrun): ppo_allreduce_sumsq over symmetric memory against NCCL all_reduce, eagerly and
replayed from a CUDA graph, with timings. `python -m torch.distributed.run --nproc-per-node 2 tools/probe_peer_allreduce.py`"""
import ctypes
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import faulthandler  # noqa: E402

import torch  # noqa: E402
import torch.distributed as td  # noqa: E402
import torch.distributed._symmetric_memory as symm  # noqa: E402

from diffusion_piano_b200.ppo import _kernels  # noqa: E402

faulthandler.dump_traceback_later(90, exit=True)
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
td.init_process_group("nccl", device_id=dev)
sys.path.insert(0, str(Path(__file__).resolve().parent))
from experiments.load import load_experiments  # noqa: E402

lib = load_experiments()  # product kernels + tools/experiments (this probe's kernel is not in the product library)
n = 227_000 // 4 * 4
grad = symm.empty(n, dtype=torch.float32, device=dev)
flags = symm.empty(lib.ppo_allreduce_flag_words(), dtype=torch.int32, device=dev)
flags.zero_()
hg = symm.rendezvous(grad, td.group.WORLD)
hf = symm.rendezvous(flags, td.group.WORLD)
print(rank, "multicast_ptr", hex(hg.multicast_ptr), "ptrs", [hex(p) for p in hg.buffer_ptrs][:2], flush=True)
out = torch.zeros(n, device=dev)
epoch = torch.ones(1, dtype=torch.int32, device=dev)
err = torch.zeros(1, dtype=torch.int32, device=dev)
scratch = torch.zeros(1024, device=dev)
ticket = torch.zeros(1, dtype=torch.int32, device=dev)
ctrl = torch.zeros(ctypes.sizeof(_kernels.PPOCtrl), dtype=torch.uint8, device=dev)
td.barrier()
torch.cuda.synchronize()


MODE = int(os.environ.get("PEER_AR_MODE", "0"))  # 256: timing experiment without the closing barrier (unsafe)


def call():
    _kernels.check(lib.ppo_allreduce_sumsq(hg.buffer_ptrs_dev, hf.buffer_ptrs_dev, out.data_ptr(), n, rank, world,
                                           epoch.data_ptr(), err.data_ptr(), scratch.data_ptr(), ticket.data_ptr(),
                                           ctrl.data_ptr(), MODE, torch.cuda.current_stream().cuda_stream))


g = torch.Generator(device=dev).manual_seed(rank)
for it in range(3):
    grad.copy_(torch.randn(n, device=dev, generator=g))
    ref = grad.clone()
    td.all_reduce(ref)
    call()
    torch.cuda.synchronize()
    c = _kernels.PPOCtrl.from_buffer_copy(ctrl.cpu().numpy().tobytes())
    diff = (out - ref).abs().max().item()
    print(rank, "iter", it, "max diff vs NCCL", diff, "norm2", c.norm2[0], float((ref.double() ** 2).sum()),
          "err", int(err.item()), "epoch", int(epoch.item()), flush=True)
    assert diff < 1e-5 and int(err.item()) == 0
# graph replay + timing against NCCL
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    call()
torch.cuda.synchronize()
td.barrier()
gr = torch.cuda.CUDAGraph()
with torch.cuda.graph(gr, capture_error_mode="thread_local"):
    for _ in range(20):
        call()
gn = torch.cuda.CUDAGraph()
ref = grad.clone()
with torch.cuda.graph(gn, capture_error_mode="thread_local"):
    for _ in range(20):
        td.all_reduce(ref)
for name, graph in (("peer kernel (all-reduce + sumsq)", gr), ("NCCL all_reduce", gn)):
    graph.replay()
    torch.cuda.synchronize()
    td.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        graph.replay()
    e1.record()
    torch.cuda.synchronize()
    if rank == 0:
        print(f"{name}: {1e3 * e0.elapsed_time(e1) / 200:.2f} us per call ({n * 4 / 1e6:.2f} MB, {world} GPUs)", flush=True)
print(rank, "err", int(err.item()), "epoch", int(epoch.item()), flush=True)
td.barrier()
sys.stdout.flush()
os._exit(0)
