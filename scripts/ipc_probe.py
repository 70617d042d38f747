"""2 ranks: rank 1 exports a layers tensor over CUDA IPC, rank 0 reads it from its own GPU with a kernel."""
import os, sys, ctypes as C
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from soillib_b200 import _lib
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = _lib.load()
pad = torch.zeros(1000, device="cuda")                      # make the tensor a sub-allocation
L = torch.arange(64 * 32 * 2, dtype=torch.float32, device="cuda").reshape(64, 32, 2) + 1000 * rank
handle, off = C.create_string_buffer(64), C.c_uint64(0)
_lib.check(lib.soilb200_ipc_export(L.data_ptr(), handle, C.byref(off)), "export")
got = [None] * world
dist.all_gather_object(got, (handle.raw, int(off.value)))
if rank == 0:
    h, o = got[1]
    base = C.c_void_p()
    _lib.check(lib.soilb200_ipc_open(h, C.byref(base)), "open")
    out = torch.zeros((64, 32), device="cuda")
    _lib.check(lib.soilb200_layer_merge(out.data_ptr(), base.value + o, 64, 32, None), "merge")
    torch.cuda.synchronize()
    want = (torch.arange(64 * 32 * 2, dtype=torch.float32).reshape(64, 32, 2) + 1000).sum(-1)
    print("peer read ok:", bool(torch.equal(out.cpu(), want)), "offset", o, flush=True)
dist.barrier()
dist.destroy_process_group()
