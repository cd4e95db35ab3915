"""Feasibility probe for peer memory between the ranks of one box (run under torchrun, >= 2 GPUs):
(a) torch symmetric memory rendezvous (peer buffer pointers + signal pads), (b) legacy CUDA IPC handles through
libcudart.  Prints one line per mechanism and rank.     torchrun --nproc-per-node 2 tools/peer_probe.py"""
import ctypes as C
import os
import sys
import traceback

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device(f"cuda:{local}")
dist.init_process_group("nccl", device_id=dev)
other = (rank + 1) % world


def probe_symm():
    import torch.distributed._symmetric_memory as sm
    t = sm.empty(1 << 20, dtype=torch.int32, device=dev)
    t.zero_()
    h = sm.rendezvous(t, dist.group.WORLD)
    print(f"[{rank}] symm: buffer_ptrs={[hex(p) for p in h.buffer_ptrs]} signal_pads={[hex(p) for p in h.signal_pad_ptrs]} "
          f"pad_size={h.signal_pad_size} mc_ptr={hex(h.multicast_ptr) if h.multicast_ptr else None}", flush=True)
    h.barrier()
    peer = h.get_buffer(other, (1 << 20,), torch.int32)
    peer.fill_(rank + 1)
    torch.cuda.synchronize()
    h.barrier()
    torch.cuda.synchronize()
    want = ((rank - 1) % world) + 1
    ok = bool((t == want).all().item())
    # bandwidth of a peer write
    src = torch.ones(64 << 20, dtype=torch.int32, device=dev)
    big = sm.empty(64 << 20, dtype=torch.int32, device=dev)
    hb = sm.rendezvous(big, dist.group.WORLD)
    pb = hb.get_buffer(other, (64 << 20,), torch.int32)
    hb.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    pb.copy_(src)
    e0.record()
    for _ in range(5):
        pb.copy_(src)
    e1.record()
    torch.cuda.synchronize()
    gbs = 5 * 256e6 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    hb.barrier()
    print(f"[{rank}] symm: peer write visible={ok}, peer copy {gbs:.0f} GB/s", flush=True)


def probe_ipc():
    rt = C.CDLL("libcudart.so.12") if os.path.exists("/usr/local/cuda/lib64/libcudart.so.12") else C.CDLL("libcudart.so")
    ptr = C.c_void_p()
    assert rt.cudaMalloc(C.byref(ptr), 1 << 22) == 0
    class Handle(C.Structure):
        _fields_ = [("reserved", C.c_byte * 64)]
    handle = Handle()
    rc = rt.cudaIpcGetMemHandle(C.byref(handle), ptr)
    print(f"[{rank}] ipc: cudaIpcGetMemHandle rc={rc}", flush=True)
    handles = [None] * world
    dist.all_gather_object(handles, bytes(handle))
    peer = C.c_void_p()
    h2 = Handle.from_buffer_copy(handles[other])
    rt.cudaIpcOpenMemHandle.argtypes = [C.POINTER(C.c_void_p), Handle, C.c_uint]
    rc = rt.cudaIpcOpenMemHandle(C.byref(peer), h2, 1)
    print(f"[{rank}] ipc: cudaIpcOpenMemHandle rc={rc} peer={hex(peer.value or 0)}", flush=True)
    if rc == 0:
        rc = rt.cudaMemset(peer, rank + 1, 1 << 22)
        torch.cuda.synchronize()
        dist.barrier()
        host = (C.c_ubyte * 16)()
        rt.cudaMemcpy(host, ptr, 16, 2)
        print(f"[{rank}] ipc: memset peer rc={rc}; own buffer now starts with {list(host)[:4]} (want {((rank - 1) % world) + 1})", flush=True)
        dist.barrier()
        rt.cudaIpcCloseMemHandle(peer)
    rt.cudaFree(ptr)


for name, fn in (("symm", probe_symm), ("ipc", probe_ipc)):
    try:
        fn()
    except Exception:  # noqa: BLE001
        print(f"[{rank}] {name}: FAILED\n{traceback.format_exc()}", flush=True)
    dist.barrier()
dist.destroy_process_group()
