#!/usr/bin/env python
"""Device -> pinned host copies of a 1080p f64 frame (49.8 MB): one cudaMemcpyAsync against chunks, one or two streams.
usage: python tools/dma_probe.py"""
import ctypes
import time

import torch

rt = ctypes.CDLL("libcudart.so.12") if True else None


def main():
    n = 1920 * 1080 * 24
    dev = torch.empty(n, dtype=torch.uint8, device="cuda")
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]
    for chunk_kb, streams in ((0, 1), (256, 1), (1024, 1), (1024, 2), (2048, 1), (4096, 1), (4096, 2), (8192, 1)):
        chunk = n if chunk_kb == 0 else chunk_kb << 10
        best = 1e9
        for rep in range(8):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            off, k = 0, 0
            while off < n:
                b = min(chunk, n - off)
                st = (s1, s2)[k % streams]
                rt.cudaMemcpyAsync(host.data_ptr() + off, dev.data_ptr() + off, b, 2, st.cuda_stream)
                off += b
                k += 1
            s1.synchronize()
            s2.synchronize()
            best = min(best, time.perf_counter() - t0)
        print("chunk %5d KB x %d stream(s): %.3f ms  %.1f GB/s" % (chunk_kb, streams, best * 1e3, n / best / 1e9), flush=True)


if __name__ == "__main__":
    main()
