"""Latency of one in-kernel cross-rank exchange through the peer scratch (dexgg_p2p_small): 40 back-to-back launches per step,
CUDA-event time per launch.  torchrun --nproc-per-node N tools/p2p_probe.py"""
import ctypes, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
from dex_gan_grasp_b200 import _native
from dex_gan_grasp_b200._native import SlabComm

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
L = _native.lib()
mine, handle = ctypes.c_void_p(), (ctypes.c_ubyte * 64)()
_native.check(L.dexgg_peer_alloc(L.dexgg_slab_scratch_bytes(), ctypes.byref(mine), handle))
handles = [None] * world
dist.all_gather_object(handles, bytes(handle))
ptrs = []
for r, h in enumerate(handles):
    if r == rank:
        ptrs.append(mine.value)
    else:
        q = ctypes.c_void_p()
        _native.check(L.dexgg_peer_open((ctypes.c_ubyte * 64).from_buffer_copy(h), ctypes.byref(q)))
        ptrs.append(q.value)
table = torch.tensor(ptrs, dtype=torch.int64, device=dev)
parity = torch.zeros(1, dtype=torch.int32, device=dev)
buf = torch.ones(64, device=dev)
dist.barrier()
st = torch.cuda.current_stream().cuda_stream
res = {}
for n_in in (64,):
    times = []
    for rep in range(6):
        _native.check(L.dexgg_slab_step_begin(mine, parity.data_ptr(), st))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(40):
            c = SlabComm()
            c.peers, c.world, c.rank, c.slot, c.parity = table.data_ptr(), world, rank, i, parity.data_ptr()
            _native.check(L.dexgg_p2p_small(ctypes.byref(c), buf.data_ptr(), n_in, buf.data_ptr(), None, st))
        b.record()
        b.synchronize()
        times.append(a.elapsed_time(b) / 40 * 1e3)
    res[n_in] = times
# the same through NCCL
t2 = []
for rep in range(4):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for i in range(40):
        dist.all_reduce(buf)
    b.record()
    b.synchronize()
    t2.append(a.elapsed_time(b) / 40 * 1e3)
if rank == 0:
    print(json.dumps({"world": world, "p2p_small_us_per_launch": res, "nccl_allreduce_us_per_call": t2, "check": float(buf[0])}))
import threading, time
threading.Thread(target=lambda: (time.sleep(10), os._exit(0)), daemon=True).start()
dist.destroy_process_group()
