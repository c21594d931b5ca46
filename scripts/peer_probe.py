"""2-GPU probe: kernel loads from a peer GPU's memory mapped with CUDA IPC (torchrun, 2 ranks)."""
import ctypes
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "dgl-1_b200"))
from dgl_b200 import _lib, ops  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
lib = _lib.get_lib()
n, f = 1 << 20, 128
x = (torch.arange(n * f, device=dev, dtype=torch.float32).reshape(n, f) + rank * 0.5)
st = x.untyped_storage()
meta = st._share_cuda_()
handle, size, off = meta[1], meta[2], meta[3]
objs = [None] * world
dist.all_gather_object(objs, (bytes(handle), int(size), int(off), int(x.storage_offset())))
dist.barrier()
peer = (rank + 1) % world
h, size, off, toff = objs[peer]
print(rank, "torch handle bytes", len(h), "type byte", h[0] if len(h) == 65 else None, flush=True)
if len(h) == 66:      # torch prefixes the cudaIpcMemHandle_t with (version, handle type) bytes; type "c" = cudaMalloc
    print(rank, "prefix", h[0], h[1], flush=True)
    assert h[1] == ord("c"), "expandable-segment handles are not cudaIpcMemHandle_t"
    h = h[2:]
elif len(h) == 65:
    h = h[1:]
ptr = ctypes.c_void_p()
_lib.check_call(lib.dglb200_ipc_open(ctypes.c_char_p(h), ctypes.byref(ptr)))
base = ptr.value + off + toff * 4
print(rank, "mapped peer", peer, hex(base), flush=True)
idx = torch.randint(0, n, (200000,), device=dev)
out = torch.empty(idx.numel(), f, device=dev)
bad = torch.zeros(1, dtype=torch.int32, device=dev)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for it in range(3):
    ev0.record()
    _lib.check_call(lib.dglb200_gather_rows(idx.numel(), ctypes.c_void_p(idx.data_ptr()), n, ctypes.c_void_p(base),
                                            f * 4, ctypes.c_void_p(out.data_ptr()), ctypes.c_void_p(bad.data_ptr()),
                                            ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)))
    ev1.record()
    torch.cuda.synchronize()
ms = ev0.elapsed_time(ev1)
want = (idx[:, None] * f + torch.arange(f, device=dev)[None, :]).float() + peer * 0.5
print(rank, "peer gather ok:", bool(torch.equal(out, want)), "%.3f ms  %.1f GB/s" % (ms, idx.numel() * f * 4 / ms / 1e6), flush=True)
dist.barrier()
_lib.check_call(lib.dglb200_ipc_close(ptr))
dist.destroy_process_group()
