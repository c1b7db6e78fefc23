"""Probe: does legacy CUDA IPC (cudaIpcGetMemHandle / cudaIpcOpenMemHandle) work between the ranks of one box?
Run under torchrun with 2+ ranks.  Prints one line per rank."""
import ctypes as C
import glob
import os

import torch
import torch.distributed as dist


class Handle(C.Structure):
    _fields_ = [('reserved', C.c_char * 64)]


def main():
    rank, world = int(os.environ['RANK']), int(os.environ['WORLD_SIZE'])
    torch.cuda.set_device(rank)
    dist.init_process_group('nccl', device_id=torch.device('cuda:%d' % rank))
    torch.zeros(1, device='cuda')
    cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), 'lib', 'libcudart*.so*')) + \
        glob.glob('/usr/local/cuda/lib64/libcudart.so*')
    rt = C.CDLL(cands[0])
    n = 1 << 20
    ptr = C.c_void_p()
    assert rt.cudaMalloc(C.byref(ptr), C.c_size_t(n)) == 0
    assert rt.cudaMemset(ptr, C.c_int(rank + 1), C.c_size_t(n)) == 0
    assert rt.cudaDeviceSynchronize() == 0
    handle = Handle()
    rc = rt.cudaIpcGetMemHandle(C.byref(handle), ptr)
    mine = bytes(C.string_at(C.byref(handle), 64))
    got = [None] * world
    dist.all_gather_object(got, (rc, mine))
    ok = True
    for peer in range(world):
        if peer == rank:
            continue
        prc, ph = got[peer]
        h = Handle.from_buffer_copy(ph)
        pp = C.c_void_p()
        rt.cudaIpcOpenMemHandle.argtypes = [C.POINTER(C.c_void_p), Handle, C.c_uint]
        r = rt.cudaIpcOpenMemHandle(C.byref(pp), h, C.c_uint(1))
        val = -1
        if r == 0:
            buf = torch.zeros(n, dtype=torch.uint8, device='cuda')
            r2 = rt.cudaMemcpy(C.c_void_p(buf.data_ptr()), pp, C.c_size_t(n), C.c_int(3))
            rt.cudaDeviceSynchronize()
            val = int(buf[12345]) if r2 == 0 else -2
        ok = ok and r == 0 and val == peer + 1
        print('rank %d: get rc %d, open peer %d rc %d, read value %d (expect %d)' % (rank, rc, peer, r, val, peer + 1), flush=True)
    dist.barrier()
    print('rank %d: IPC %s' % (rank, 'OK' if ok else 'FAILED'), flush=True)
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
