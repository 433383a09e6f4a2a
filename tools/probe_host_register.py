"""Probe: page-locking (cudaHostRegister) a 20 GB /dev/shm mapping: one call vs 1 GB chunks, one vs two processes."""
import os, sys, time, ctypes, glob, torch, multiprocessing as mp

def cudart():
    import nvidia.cuda_runtime as m
    p = glob.glob(os.path.join(list(m.__path__)[0], "lib", "libcudart.so*"))
    return ctypes.CDLL(p[0])

def worker(rank, path, n, chunk, q, prefault):
    torch.cuda.init(); torch.ones(1, device="cuda")
    rt = cudart()
    rt.cudaHostRegister.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_uint]
    t = torch.from_file(path, shared=True, size=n // 8, dtype=torch.float64)
    if prefault:
        t0 = time.time(); t.view(torch.uint8)[::4096] = 0; pf = time.time() - t0
    else: pf = 0
    t0 = time.time(); rcs = []
    for o in range(0, n, chunk):
        rc = rt.cudaHostRegister(t.data_ptr() + o, min(chunk, n - o), 3)
        rcs.append(rc)
        if rc: rt.cudaGetLastError(); break
    dt = time.time() - t0
    q.put((rank, n >> 30, chunk >> 30, "prefault" if prefault else "", rcs[-1], len(rcs), round(pf, 2), round(dt, 2)))
    time.sleep(2)
    for i, rc in enumerate(rcs):
        if rc == 0: rt.cudaHostUnregister(ctypes.c_void_p(t.data_ptr() + i * chunk))

if __name__ == "__main__":
    mp.set_start_method("spawn")
    n = 20 << 30
    for nproc, chunk, prefault in ((1, n, False), (1, 1 << 30, False), (1, n, True), (2, 1 << 30, True), (2, 4 << 30, False)):
        path = "/dev/shm/pf_probe.bin"
        with open(path, "wb") as fh: fh.truncate(n)
        q = mp.Queue()
        ps = [mp.Process(target=worker, args=(r, path, n, chunk, q, prefault)) for r in range(nproc)]
        for p in ps: p.start()
        for p in ps: p.join(120)
        while not q.empty(): print(q.get(), flush=True)
        os.remove(path)
    os.system("ulimit -a | grep -i lock; cat /proc/meminfo | grep -i -E 'unevict|mlock|shmem'")
